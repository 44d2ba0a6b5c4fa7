#!/usr/bin/env python
"""Regenerate artifacts/sass/ from the library that is actually loaded: `cuobjdump -sass _lib/libdlion.so`, one
listing per `__global__` kernel of csrc/ (the bf16 instantiation that the GPT-2 / Llama training path launches; the
fused exchange kernel also in fp32), plus SUMMARY.md with per-kernel mnemonic counts.  SUMMARY.md carries the build
stamp of the library it was made from (_lib/stamp.txt = hash of csrc/, the header and the compile flags);
tests/test_sass_artifacts.py fails when the two differ, so a listing cannot go stale unnoticed.

    python scripts/dump_sass.py            # rebuilds the library if needed, then rewrites artifacts/sass/
"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
OUT = ROOT / "artifacts" / "sass"

# kernel name (as in csrc) -> list of (file suffix, regex on the mangled name) to keep as full listings
WANT = {
    "lion_fused_kernel": [("bf16", r"lion_fused_kernelI13__nv_bfloat16"), ("fp32", r"lion_fused_kernelIfE")],
    "lion_local_kernel": [("bf16", r"lion_local_kernelI13__nv_bfloat16")],
    "sign_pack_kernel": [("bf16", r"sign_pack_kernelI13__nv_bfloat16")],
    "vote_update_kernel": [("bf16", r"vote_update_kernelI13__nv_bfloat16")],
    "multi_gather_kernel": [("bf16", r"multi_gather_kernelI13__nv_bfloat16S1_")],
    "sumsq_kernel": [("bf16", r"sumsq_kernelI13__nv_bfloat16")],
    "clip_scale_kernel": [("bf16", r"clip_scale_kernelI13__nv_bfloat16")],
    "norm_fwd_kernel": [("bf16_ln768_dropout", r"norm_fwd_kernelI13__nv_bfloat16Li1ELi96ELb0ELb1"),
                        ("bf16_rms4096", r"norm_fwd_kernelI13__nv_bfloat16Li4ELi256ELb1ELb0")],
    "norm_fwd_tma_kernel": [("bf16_ln768_dropout", r"norm_fwd_tma_kernelI13__nv_bfloat16Li3ELb0ELb1")],
    "norm_bwd_kernel": [("bf16_ln768_dropout_dxsum", r"norm_bwd_kernelI13__nv_bfloat16Li1ELi96ELb0ELi1ELb1ELb1"),
                        ("bf16_rms4096", r"norm_bwd_kernelI13__nv_bfloat16Li4ELi256ELb1ELi1ELb0ELb0")],
    "norm_bwd_finalize_kernel": [("bf16", r"norm_bwd_finalize_kernelI13__nv_bfloat16")],
    "colsum_kernel": [("bf16_plain_fused", r"colsum_kernelI13__nv_bfloat16Lb1ENS_11PlainSource"),
                      ("bf16_gelu_bwd", r"colsum_kernelI13__nv_bfloat16Lb1ENS_13GeluBwdSource"),
                      ("bf16_qkv_pack", r"colsum_kernelI13__nv_bfloat16Lb1ENS_13QkvPackSource")],
    "gelu_fwd_kernel": [("bf16", r"gelu_fwd_kernelI13__nv_bfloat16")],
    "gelu_bwd_kernel": [("bf16", r"gelu_bwd_kernelI13__nv_bfloat16")],
    "swiglu_fwd_kernel": [("bf16", r"swiglu_fwd_kernelI13__nv_bfloat16")],
    "swiglu_bwd_kernel": [("bf16", r"swiglu_bwd_kernelI13__nv_bfloat16")],
    "rope_kernel": [("bf16", r"rope_kernelI13__nv_bfloat16")],
    "ce_rowres_kernel": [("bf16_fwd_bwd", r"ce_rowres_kernelI13__nv_bfloat16Li1E")],
    "embed_fwd_kernel": [("bf16", r"embed_fwd_kernelI13__nv_bfloat16")],
    "embed_bwd_sorted_kernel": [("bf16", r"embed_bwd_sorted_kernelI13__nv_bfloat16")],
    "ce_kernel": [("bf16_fwd_bwd", r"ce_kernelI13__nv_bfloat16Li1E"), ("bf16_rowscale", r"ce_kernelI13__nv_bfloat16Li2E")],
    "lmhead_lse_kernel": [("tcgen05", r"lmhead_lse_kernel")],
    "lmhead_merge_kernel": [("fp32", r"lmhead_merge_kernel")],
    "dgrad_dgelu_kernel": [("bf16_tcgen05", r"dgrad_dgelu_kernelI13__nv_bfloat16")],
    "dgrad_dgelu_pair_kernel": [("bf16_cta_group2", r"dgrad_dgelu_pair_kernelI13__nv_bfloat16")],
    "dgelu_bias_finalize_kernel": [("bf16", r"dgelu_bias_finalize_kernelI13__nv_bfloat16")],
    "q4_quantize_kernel": [("bf16", r"q4_quantize_kernelI13__nv_bfloat16")],
    "q4_dequantize_kernel": [("bf16", r"q4_dequantize_kernelI13__nv_bfloat16")],
}
MNEMONICS = re.compile(r"^\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)")
KEEP = re.compile(r"^(UBLKCP|UTMA|UTC|LDTM|STTM|SYNCS|LDG|STG|LDS|STS|ATOM|RED|MEMBAR|CCTL|ERRBAR|BAR|F2F|F2FP|MUFU|HMMA|"
                  r"NANOSLEEP|REDUX|SHFL|LDGSTS|DEPBAR)")


def split_functions(text):
    cur, name = [], None
    for line in text.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                yield name, cur
            name, cur = m.group(1), []
        elif name:
            cur.append(line)
    if name:
        yield name, cur


def main():
    from distributed_lion_pytorch_b200 import build as b
    lib = b.build(force=False)
    stamp = (b.LIBDIR / "stamp.txt").read_text().strip()
    text = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True, check=True).stdout
    funcs = dict(split_functions(text))
    OUT.mkdir(parents=True, exist_ok=True)
    for old in OUT.glob("*.sass"):
        old.unlink()
    missing, lines = [], []
    for kern, variants in WANT.items():
        for suffix, pat in variants:
            hit = [n for n in funcs if re.search(pat, n)]
            if not hit:
                missing.append(f"{kern} ({pat})")
                continue
            name = sorted(hit)[0]
            (OUT / f"{kern}_{suffix}.sass").write_text(f"// {name}\n// library build stamp {stamp}\n" + "\n".join(funcs[name]) + "\n")
    if missing:
        raise SystemExit("no SASS found for: " + ", ".join(missing))
    for name in sorted(funcs):
        if "dlion" not in name or ("13__nv_bfloat16" not in name and "lion_fused_kernelIfE" not in name and "lmhead" not in name):
            continue
        cnt = collections.Counter()
        n = 0
        for line in funcs[name]:
            m = MNEMONICS.match(line)
            if m:
                n += 1
                if KEEP.match(m.group(1)):
                    cnt[m.group(1)] += 1
        lines.append(f"* `{name}` — {n} instr — " + ", ".join(f"{k}×{v}" for k, v in sorted(cnt.items())))
    (OUT / "SUMMARY.md").write_text(
        "# SASS mnemonic summary (cuobjdump -sass _lib/libdlion.so, sm_100a)\n\n"
        f"build-stamp: {stamp}\n\n"
        "Regenerate with `python scripts/dump_sass.py`; `tests/test_sass_artifacts.py` fails when the stamp above is not\n"
        "the stamp of the library built from the current `csrc/`.  Per kernel (bf16 instantiations; the fused kernel also\n"
        "fp32): total instructions and the memory / barrier / conversion mnemonics.  `UBLKCP.S.G` / `UBLKCP.G.S` =\n"
        "cp.async.bulk (TMA 1-D) global→shared / shared→global, `SYNCS.*` = mbarrier ops, `*.STRONG.SYS` = system-scope\n"
        "accesses to peer-visible memory, `MEMBAR.*.SYS` = fence.acq_rel.sys, `F2FP.*.PACK_AB` = packed bf16 rounding;\n"
        "`UTMALDG.2D` = cp.async.bulk.tensor.2d (TMA tile load), `UTCHMMA` = tcgen05.mma kind::f16, `UTCBAR` = tcgen05.commit,\n"
        "`LDTM` = tcgen05.ld (TMEM to registers), `UTCATOMSWS` = tcgen05.alloc / dealloc (the LM-head kernel).\n"
        "Full listings of the kernels on the training path are the `*.sass` files beside this one.\n\n"
        + "\n".join(lines) + "\n")
    print(f"wrote {len(list(OUT.glob('*.sass')))} listings + SUMMARY.md for stamp {stamp[:16]}…")


if __name__ == "__main__":
    main()

"""The committed SASS listings must come from the library that the current sources build (the listings
had gone stale).  CPU-only: compares the build stamp recorded in artifacts/sass/SUMMARY.md with the stamp
`distributed_lion_pytorch_b200.build` derives from csrc/, include/dlion.h and the compile flags."""
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_sass_summary_matches_the_current_sources():
    from distributed_lion_pytorch_b200 import build as b
    summary = (ROOT / "artifacts" / "sass" / "SUMMARY.md").read_text()
    m = re.search(r"^build-stamp: ([0-9a-f]{64})$", summary, re.M)
    assert m, "artifacts/sass/SUMMARY.md carries no build stamp: run python scripts/dump_sass.py"
    want = b._stamp(b._sources())
    assert m.group(1) == want, "csrc/ changed since the SASS listings were made: run python scripts/dump_sass.py"


def test_every_training_path_kernel_has_a_listing_with_the_expected_hardware_paths():
    d = ROOT / "artifacts" / "sass"
    names = {p.name for p in d.glob("*.sass")}
    for k in ("lion_fused_kernel_bf16", "lion_fused_kernel_fp32", "lion_local_kernel_bf16", "sign_pack_kernel_bf16",
              "vote_update_kernel_bf16", "multi_gather_kernel_bf16", "sumsq_kernel_bf16", "norm_fwd_kernel_bf16_rms4096",
              "norm_bwd_kernel_bf16_rms4096", "colsum_kernel_bf16_gelu_bwd", "colsum_kernel_bf16_qkv_pack",
              "swiglu_fwd_kernel_bf16", "swiglu_bwd_kernel_bf16", "gelu_bwd_kernel_bf16", "rope_kernel_bf16",
              "ce_kernel_bf16_fwd_bwd", "ce_kernel_bf16_rowscale", "q4_quantize_kernel_bf16", "q4_dequantize_kernel_bf16"):
        assert k + ".sass" in names, k
    fused = (d / "lion_fused_kernel_bf16.sass").read_text()
    for mnemonic in ("UBLKCP.S.G", "UBLKCP.G.S", "SYNCS.ARRIVE.TRANS64", "STG.E.128.STRONG.SYS", "LDG.E.64.STRONG.SYS"):
        assert mnemonic in fused, mnemonic          # TMA in both directions, mbarriers, multimem / system-scope traffic

#!/usr/bin/env python
"""Which library attention path is fastest for the two model shapes, measured the way the models use
it: packed qkv projection output [B, S, 3E] in, [B, S, E] out, backward down to d(qkv) — so layout copies
and the concat of dq/dk/dv count.  CUDA-event times, median of N, L2 flushed between iterations."""
import argparse
import json
import sys
from pathlib import Path

import torch
import torch.nn.functional as F
from torch.nn.attention import SDPBackend, sdpa_kernel

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def timeit(fn, iters=10, warm=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for i in range(iters):
        flush.fill_(i)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def sdpa_path(qkv, B, S, H, D, p, backend):
    q, k, v = (t.transpose(1, 2) for t in qkv.view(B, S, 3, H, D).unbind(2))
    if backend is None:
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True, dropout_p=p)
    else:
        with sdpa_kernel(backend):
            y = F.scaled_dot_product_attention(q, k, v, is_causal=True, dropout_p=p)
    return y.transpose(1, 2).reshape(B, S, H * D)


def fa_path(qkv, B, S, H, D, p):
    from flash_attn import flash_attn_qkvpacked_func
    return flash_attn_qkvpacked_func(qkv.view(B, S, 3, H, D), dropout_p=p, causal=True).reshape(B, S, H * D)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    rows = []
    for tag, B, S, H, D, p in (("gpt2 B20 H12 S1024 D64 p0.1", 20, 1024, 12, 64, 0.1),
                               ("gpt2 B20 H12 S1024 D64 p0", 20, 1024, 12, 64, 0.0),
                               ("llama7b B4 H32 S1024 D128 p0", 4, 1024, 32, 128, 0.0)):
        qkv = torch.randn(B, S, 3 * H * D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        dy = torch.randn(B, S, H * D, device="cuda", dtype=torch.bfloat16)
        variants = [("sdpa default", lambda: sdpa_path(qkv, B, S, H, D, p, None)),
                    ("sdpa cudnn", lambda: sdpa_path(qkv, B, S, H, D, p, SDPBackend.CUDNN_ATTENTION)),
                    ("sdpa flash", lambda: sdpa_path(qkv, B, S, H, D, p, SDPBackend.FLASH_ATTENTION)),
                    ("sdpa efficient", lambda: sdpa_path(qkv, B, S, H, D, p, SDPBackend.EFFICIENT_ATTENTION)),
                    ("flash_attn qkvpacked", lambda: fa_path(qkv, B, S, H, D, p))]
        ref = None
        for name, fwd in variants:
            try:
                def step():
                    qkv.grad = None
                    fwd().backward(dy)
                us = timeit(step)
                us_f = timeit(lambda: fwd())
                with torch.no_grad():
                    y = fwd().float() if p == 0 else None
                err = None
                if y is not None:
                    ref = y if ref is None else ref
                    err = float((y - ref).abs().max())
                rows.append({"shape": tag, "path": name, "fwd_bwd_us": round(us, 1), "fwd_us": round(us_f, 1),
                             "max_abs_diff_vs_first": err})
                print(f"{tag:34s} {name:22s} fwd+bwd {us:8.1f} us   fwd {us_f:8.1f} us   diff {err}")
            except Exception as e:      # noqa: BLE001 — a backend that cannot run this shape is a result too
                rows.append({"shape": tag, "path": name, "error": str(e)[:200]})
                print(f"{tag:34s} {name:22s} unavailable: {str(e)[:120]}")
    if a.out:
        Path(a.out).write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()

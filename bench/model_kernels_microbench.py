#!/usr/bin/env python
"""Microbenchmark of the model-side kernels (N3, N4, N5, colsum) at the GPT-2-small and Llama-2-7B
shapes of the BASELINE configs: CUDA-event time (median of N, L2 flushed between iterations),
algorithmic HBM bytes, achieved GB/s and fraction of the measured copy peak."""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion_pytorch_b200.ops import _cabi  # noqa: E402
from distributed_lion_pytorch_b200.ops.cross_entropy import _ce_kernel  # noqa: E402
from distributed_lion_pytorch_b200.ops.linear import column_sum  # noqa: E402
from distributed_lion_pytorch_b200.ops.norm import _FusedNorm  # noqa: E402
from distributed_lion_pytorch_b200.ops.rope import _launch as rope_launch, rope_tables  # noqa: E402


def timeit(fn, iters=12, warm=3):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    peak = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"] if (ROOT / "MEASURED_PEAKS.json").exists() else 6650.0
    dt = torch.bfloat16
    rows_out = []

    def row(name, us, nbytes):
        gbs = nbytes / us / 1e3
        rows_out.append({"kernel": name, "us": round(us, 2), "MB": round(nbytes / 1e6, 1), "GBps": round(gbs, 1),
                         "frac_of_copy_peak": round(gbs / peak, 3)})
        print(f"{name:58s} {us:9.1f} us {nbytes / 1e6:9.1f} MB {gbs:8.1f} GB/s {gbs / peak:6.1%}")

    for tag, rows, cols, rms in (("gpt2 [20480x768] LN", 20480, 768, False), ("llama7b [4096x4096] RMS", 4096, 4096, True)):
        x = torch.randn(rows, cols, device="cuda").to(dt)
        r = torch.randn(rows, cols, device="cuda").to(dt)
        w = torch.ones(cols, device="cuda", dtype=dt)
        b = None if rms else torch.zeros(cols, device="cuda", dtype=dt)
        nb = rows * cols * 2
        for p in (0.0, 0.1):
            if rms and p > 0:
                continue
            us = timeit(lambda: _FusedNorm.apply(x, r, w, b, 1e-5, rms, p, 7))
            row(f"norm_fwd {tag} dropout={p}", us, 4 * nb)
            xg = x.clone().requires_grad_()
            rg = r.clone().requires_grad_()
            wg = w.clone().requires_grad_()
            bg = None if b is None else b.clone().requires_grad_()
            y, s = _FusedNorm.apply(xg, rg, wg, bg, 1e-5, rms, p, 7)
            dy, ds = torch.randn_like(y), torch.randn_like(y)
            us = timeit(lambda: torch.autograd.grad((y, s), (xg, rg, wg) + (() if bg is None else (bg,)), (dy, ds),
                                                    retain_graph=True))
            row(f"norm_bwd+finalize {tag} dropout={p}", us, (4 + (1 if p > 0 else 0)) * nb)
    for tag, rows, cols in (("gpt2 c_fc dY [20480x3072]", 20480, 3072), ("gpt2 c_proj dY [20480x768]", 20480, 768)):
        dy = torch.randn(rows, cols, device="cuda").to(dt)
        us = timeit(lambda: column_sum(dy, dt))
        row(f"colsum+finalize {tag}", us, rows * cols * 2)
    for tag, rows, V in (("gpt2 [4096x50257]", 4096, 50257), ("llama [4096x32000]", 4096, 32000)):
        Vp = (V + 63) // 64 * 64
        logits = torch.randn(rows, Vp, device="cuda").to(dt)
        labels = torch.randint(0, V, (rows,), device="cuda")
        losses = torch.empty(rows, device="cuda")
        us = timeit(lambda: _ce_kernel(logits, labels, losses, V, -100, True))
        row(f"ce_fwd_bwd {tag}", us, 3 * rows * Vp * 2)
    cos, sin = rope_tables(1024, 128, device="cuda")
    q = torch.randn(4, 1024, 32 * 128, device="cuda").to(dt)
    k = torch.randn(4, 1024, 32 * 128, device="cuda").to(dt)
    us = timeit(lambda: rope_launch(q, k, cos, sin, 1024, 128, False))
    row("rope llama7b q,k [4x1024x32x128]", us, 2 * 2 * q.numel() * 2)
    from distributed_lion_pytorch_b200.ops import quant4
    K = 6       # calls per timed region on K different weight sets: a single call is shorter than its host launch path
    for tag, shape in (("llama7b q_proj [4096x4096]", (4096, 4096)), ("llama7b up_proj [11008x4096]", (11008, 4096))):
        wts = [(torch.randn(shape, device="cuda") * 0.02).to(dt) for _ in range(K)]
        n = wts[0].numel()
        us = timeit(lambda: [quant4.quantize_4bit(w_) for w_ in wts]) / K
        row(f"q4_quantize nf4 {tag} (mean of {K} back-to-back)", us, n * 2 + n // 2 + n // 16)
        qs = [quant4.quantize_4bit(w_) for w_ in wts]
        bufs = [torch.empty(n, device="cuda", dtype=dt) for _ in range(K)]
        us = timeit(lambda: [quant4.dequantize_4bit(p_, a_, shape, dt, out=b_) for (p_, a_), b_ in zip(qs, bufs)]) / K
        row(f"q4_dequantize nf4 {tag} (mean of {K} back-to-back)", us, n // 2 + n // 16 + n * 2)
        del wts, qs, bufs
    if a.out:
        Path(a.out).write_text(json.dumps({"peak_gbs": peak, "rows": rows_out}, indent=1))


if __name__ == "__main__":
    main()

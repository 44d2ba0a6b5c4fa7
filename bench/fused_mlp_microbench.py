#!/usr/bin/env python
"""tcgen05 dgrad GEMM + GELU backward + bias gradient (one kernel) against the two launches it replaces (cuBLAS GEMM, then
the one-pass activation-backward / column-sum kernel), at GPT-2 MLP shapes.  CUDA events, L2 flushed between iterations.
    python bench/fused_mlp_microbench.py [--iters 20] [--out profiles/fused_mlp_microbench.json]"""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion_pytorch_b200.ops import _cabi, linear as lin, mlp  # noqa: E402
from distributed_lion_pytorch_b200.utils.profiling import ClockSampler  # noqa: E402

SHAPES = [("gpt2 (20 x 1024 tokens)", 20480, 3072, 768), ("gpt2-medium (8 x 1024)", 8192, 4096, 1024),
          ("gpt2-xl (4 x 1024)", 4096, 6400, 1600)]


def timed(fn, iters, flush):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        torch.cuda._sleep(400_000)          # ~0.2 ms of GPU spin: the host queues the launches behind it, so the events
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)   # bracket GPU time only
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    L = _cabi.lib()
    sampler = ClockSampler(0)
    sampler.start()
    rows = []
    for name, m, n, k in SHAPES:
        g = torch.Generator(device=dev).manual_seed(0)
        dy = torch.randn(m, k, device=dev, generator=g).to(torch.bfloat16)
        w = (torch.randn(n, k, device=dev, generator=g) * k ** -0.5).to(torch.bfloat16)
        pre = torch.randn(m, n, device=dev, generator=g).to(torch.bfloat16)
        bias_grad = torch.zeros(n, device=dev, dtype=torch.bfloat16)
        dpre = torch.empty_like(pre)
        partial = torch.empty(int(L.dlion_colsum_parts(m, n)), n, dtype=torch.float32, device=dev)

        def split():
            dact = dy @ w.t()
            _cabi.check(L.dlion_gelu_tanh_bwd_colsum(pre.data_ptr(), dact.data_ptr(), dpre.data_ptr(), m, n, partial.data_ptr(),
                                                     lin._tickets(dev).data_ptr(), bias_grad.data_ptr(), 1,
                                                     _cabi.dtype_code(pre.dtype), _cabi.stream_ptr(dev)), "gelu_bwd_colsum")

        t_fused = timed(lambda: mlp.dgrad_dgelu(dy, w, pre, bias_grad_out=bias_grad, accumulate=True), a.iters, flush)
        t_split = timed(split, a.iters, flush)
        t_gemm = timed(lambda: dy @ w.t(), a.iters, flush)
        t_fused_nobias = timed(lambda: mlp.dgrad_dgelu(dy, w, pre, want_bias_grad=False), a.iters, flush)
        flops = 2.0 * m * n * k
        row = {"shape": name, "m": m, "n": n, "k": k, "fused_us": round(t_fused, 1), "fused_kernel_only_us": round(t_fused_nobias, 1), "gemm_plus_kernel_us": round(t_split, 1),
               "cublas_gemm_only_us": round(t_gemm, 1), "fused_tflops": round(flops / t_fused / 1e6, 1),
               "speedup": round(t_split / t_fused, 3), "hbm_bytes_min": 2 * (m * k + n * k + 2 * m * n)}
        rows.append(row)
        print(json.dumps(row), flush=True)
    clocks = sampler.stop()
    print(json.dumps({"clocks": clocks}))
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"rows": rows, "clocks": clocks}, indent=1))


if __name__ == "__main__":
    main()

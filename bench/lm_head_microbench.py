#!/usr/bin/env python
"""tcgen05 LM-head log-sum-exp kernel against the library path it replaces on the no-grad side (cuBLAS logits block +
CE kernel), per BASELINE shape.  CUDA events, L2 flushed between iterations, clocks sampled; one JSON line per shape.
    python bench/lm_head_microbench.py [--iters 10] [--out profiles/lm_head_microbench.json]"""
import argparse
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion_pytorch_b200.ops import cross_entropy as ce, lm_head  # noqa: E402
from distributed_lion_pytorch_b200.utils.profiling import ClockSampler  # noqa: E402

SHAPES = [("gpt2 micro-batch", 20480, 50257, 768), ("gpt2 one block", 4096, 50257, 768),
          ("llama-2-7b dpo micro-batch", 8192, 32000, 4096), ("llama-2-7b sft micro-batch", 4096, 32000, 4096)]


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
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows_out = []
    sampler = ClockSampler(0)
    sampler.start()
    for name, n, v, k in SHAPES:
        g = torch.Generator(device=dev).manual_seed(1)
        h = torch.randn(n, k, device=dev, generator=g).to(torch.bfloat16)
        w = (torch.randn(v, k, device=dev, generator=g) * k ** -0.5).to(torch.bfloat16)
        labels = torch.randint(0, v, (n,), device=dev, generator=g)
        flops = 2.0 * n * v * k
        with torch.no_grad():
            t_fused = timed(lambda: lm_head.lm_head_log_probs(h, w, labels, want_lse=False), a.iters, flush)
            os.environ["DLION_LMHEAD_FUSED"] = "0"
            t_split = timed(lambda: ce.linear_log_probs(h, w, labels), a.iters, flush)
            os.environ.pop("DLION_LMHEAD_FUSED")
            wp = ce._pad_rows(w)
            t_gemm = timed(lambda: h[:4096] @ wp.t(), a.iters, flush) * (n / 4096)
            lp_f = lm_head.lm_head_log_probs(h, w, labels, want_lse=False)[0]
            os.environ["DLION_LMHEAD_FUSED"] = "0"
            lp_s = ce.linear_log_probs(h, w, labels)
            os.environ.pop("DLION_LMHEAD_FUSED")
        row = {"shape": name, "rows": n, "vocab": v, "hidden": k, "fused_ms": round(t_fused, 4), "split_ms": round(t_split, 4),
               "cublas_logits_gemm_only_ms": round(t_gemm, 4), "fused_tflops": round(flops / t_fused / 1e9, 1),
               "split_tflops": round(flops / t_split / 1e9, 1), "speedup": round(t_split / t_fused, 3),
               "max_abs_diff_vs_split": float((lp_f - lp_s).abs().max()), "splits": lm_head.plan(n, v)[0]}
        rows_out.append(row)
        print(json.dumps(row), flush=True)
        del h, w, wp
    clocks = sampler.stop()
    res = {"rows": rows_out, "clocks": clocks}
    print(json.dumps({"clocks": clocks}))
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Optimizer-step microbenchmark on ONE GPU (BASELINE config 5, single-GPU part).

Times, with CUDA events after warm-up and an L2 flush between iterations:
  * N6  fused local Lion step                      (ours, W=1)
  * sign_pack + vote_update                        (the nccl_1bit pair, without the all_gather)
  * N1  fused exchange kernel, W ranks emulated as CTA groups of one launch
  * the reference's update_fn ATen sequence per parameter tensor (baseline/_ref)
and reports each as achieved HBM GB/s and as a fraction of MEASURED_PEAKS.json's copy bandwidth.
Bytes per element: read g, m, p + write p, m = 5 * sizeof(T)  (+ W/8 gathered bits).
"""
import argparse
import importlib.util
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion_pytorch_b200.ops import lion_ops  # noqa: E402
from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange  # noqa: E402


def peak_gbs():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        return json.loads(f.read_text())["hbm_gbs"], "measured"
    return 6650.0, "fallback"


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
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def gpt2_shapes():
    shapes = [(50257, 768), (1024, 768)]
    for _ in range(12):
        shapes += [(768,), (768,), (768, 2304), (2304,), (768, 768), (768,), (768,), (768,), (768, 3072), (3072,),
                   (3072, 768), (768,)]
    shapes += [(768,), (768,)]
    return shapes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=124_440_000)
    ap.add_argument("--dtype", default="bfloat16")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    dtype = getattr(torch, a.dtype)
    es = torch.empty(0, dtype=dtype).element_size()
    tile = lion_ops.fused_tile(dtype)
    n = a.n // 8192 * 8192
    peak, src = peak_gbs()
    res = {"n": n, "dtype": a.dtype, "peak_gbs": peak, "peak_source": src, "rows": []}

    def row(name, ms, ms_min, nbytes, note=""):
        gbs = nbytes / ms / 1e6
        res["rows"].append({"name": name, "ms_median": round(ms, 4), "ms_min": round(ms_min, 4),
                            "algorithmic_GB": round(nbytes / 1e9, 4), "GBps": round(gbs, 1),
                            "frac_of_peak": round(gbs / peak, 3), "note": note})
        print(f"{name:44s} {ms:9.3f} ms  {gbs:8.1f} GB/s  {gbs / peak:6.1%} of {src} copy peak  {note}")

    p = torch.randn(n, device="cuda").to(dtype)
    g = torch.randn(n, device="cuda").to(dtype)
    m = torch.zeros(n, device="cuda", dtype=dtype)
    ms, mn = timeit(lambda: lion_ops.local_step(p, g, m, 1e-4, 0.1, 0.9, 0.99), a.iters)
    row("N6 lion_local (W=1)", ms, mn, 5 * es * n)
    W = 8
    words = torch.empty(n // 8, dtype=torch.uint8, device="cuda")
    ms, mn = timeit(lambda: lion_ops.sign_pack(m, g, 0.9, words), a.iters)
    row("sign_pack", ms, mn, 2 * es * n + n // 8)
    gathered = torch.randint(0, 256, (W, n // 8), dtype=torch.uint8, device="cuda")
    ms2, mn2 = timeit(lambda: lion_ops.vote_update(p, g, m, gathered, W, None, 1e-4, 0.1, 0.99), a.iters)
    row("vote_update (W=8 gathered)", ms2, mn2, 5 * es * n + W * n // 8)
    row("nccl_1bit pair w/o all_gather (W=8)", ms + ms2, mn + mn2, 7 * es * n + (W + 1) * n // 8,
        "2 launches, g and m read twice")
    del gathered, words
    for Wv in (2, 4, 8):
        nv = (n // Wv) // tile * tile
        ex = EmulatedExchange(Wv, nv, dtype, "cuda")
        P = torch.randn(Wv, nv, device="cuda").to(dtype)
        G = torch.randn(Wv, nv, device="cuda").to(dtype)
        M = torch.zeros(Wv, nv, device="cuda", dtype=dtype)
        ms, mn = timeit(lambda: ex.step(P, G, M, 1e-4, 0.1, 0.9, 0.99), a.iters)
        ex.check_status()
        row(f"N1 fused, {Wv} ranks emulated on one GPU", ms, mn, Wv * nv * (5 * es) + Wv * Wv * nv // 8,
            f"grid {ex.grid}x{Wv} CTAs; aggregate over the emulated ranks")
        del ex, P, G, M
    # the reference's own op sequence (W=1: update_fn), per parameter tensor like its step() loop
    f = ROOT / "baseline" / "_ref" / "distributed_lion.py"
    if f.exists() and a.dtype == "bfloat16":
        spec = importlib.util.spec_from_file_location("ref_distributed_lion", f)
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
        ps = [torch.nn.Parameter(torch.randn(s, device="cuda").to(dtype)) for s in gpt2_shapes()]
        for q in ps:
            q.grad = torch.randn_like(q)
        opt = ref.Lion(ps, lr=1e-4, weight_decay=0.1)
        tot = sum(q.numel() for q in ps)
        ms, mn = timeit(lambda: opt.step(), a.iters)
        row("reference Lion.step (W=1, 148 tensors, ATen)", ms, mn, 5 * es * tot, "same GPT-2-small parameter census")
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()

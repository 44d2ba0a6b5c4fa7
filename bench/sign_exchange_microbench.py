#!/usr/bin/env python
"""Sign-exchange microbenchmark (BASELINE config 5), one process per GPU under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 \
        bench/sign_exchange_microbench.py --sizes 1,16,124 [--out profiles/x.json]

Per bucket size (millions of bf16 parameters) times one optimizer step, device-timed with CUDA
events, max over ranks, L2 flushed between iterations:
  fused       ours: one persistent kernel, words over NVLink peer stores (no NCCL)
  nccl_1bit   ours: pack kernel + NCCL all_gather of the true 1-bit payload + vote/update kernel
  allreduce   plain DDP-style: NCCL bf16 all-reduce of the gradient + fused local Lion kernel
  reference   the reference's update_fn_distributed on the same flat tensor (one big tensor: its
              best case — it really runs once per parameter tensor, 148x for GPT-2 small)
and prints each as time and as a fraction of the fused roofline max(HBM time, NVLink time).

`*_full` rows time what a trainer step runs around the exchange as well — gradient norm, clip, update, gradient
zero-fill: for `fused_full` that is `sumsq` + ONE fused kernel (clip and zero-fill ride inside it; roofline
14 B/param: g read twice and cleared once, m and p read and written once); for the baselines it is norm + clip
kernel + their step + memset.  Every row carries the SM clocks and throttle reasons sampled while it ran.
"""
import argparse
import gc
import importlib.util
import json
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion import Lion  # noqa: E402
from distributed_lion_pytorch_b200.ops import lion_ops  # noqa: E402
from distributed_lion_pytorch_b200.ops import bucket_ops  # noqa: E402
from distributed_lion_pytorch_b200.parallel import runtime as rt  # noqa: E402
from distributed_lion_pytorch_b200.utils.profiling import ClockSampler  # noqa: E402


BURST = 1      # --burst K: K back-to-back steps per timed sample (amortises host launch path and barrier skew)


def timeit(fn, iters, warm, device, prep=None):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    for _ in range(warm):
        if prep is not None:
            prep()
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    ts = []
    for i in range(iters):
        if prep is not None:
            prep()                      # e.g. refill the gradients a zero-filling step consumed (untimed)
        flush.fill_(i)
        dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(BURST):
            fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / BURST], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1,16,124")
    ap.add_argument("--iters", type=int, default=8)
    ap.add_argument("--out", default=None)
    ap.add_argument("--burst", type=int, default=1, help="steps per timed sample; > 1 reports the pipelined per-step time")
    ap.add_argument("--skip-reference-above", type=float, default=200.0)
    a = ap.parse_args()
    global BURST
    BURST = max(1, a.burst)
    device = rt.init_distributed()
    rank, W = rt.rank(), rt.world_size()
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    hbm = peaks.get("hbm_gbs", 6650.0)
    nvl = 770.0
    ref = None
    f = ROOT / "baseline" / "_ref" / "distributed_lion.py"
    if f.exists():
        spec = importlib.util.spec_from_file_location("ref_distributed_lion", f)
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
    rows = []
    for mp in [float(x) for x in a.sizes.split(",")]:
        n = int(mp * 1e6) // 8192 * 8192
        res = {"n": n, "world": W, "burst": BURST}
        sampler = ClockSampler(torch.cuda.current_device())
        sampler.start()
        import os
        for comm in ("fused", "fused_multicast", "nccl_1bit"):
            os.environ["DLION_MULTICAST"] = "1" if comm == "fused_multicast" else "0"
            p = torch.nn.Parameter(torch.randn(n, device=device, dtype=torch.bfloat16))
            p.grad = torch.randn(n, device=device, dtype=torch.bfloat16)
            opt = Lion([p], lr=1e-4, weight_decay=0.1, comm="fused" if comm.startswith("fused") else comm)
            t = timeit(lambda: opt.step(), a.iters, 3, device)
            unavailable = False
            if comm.startswith("fused"):
                opt._backend._ctx.check_status()
                if comm == "fused_multicast" and not opt._backend._ctx.uses_multicast:
                    t, unavailable = None, True    # fabric exposes no multicast pointer
            res[comm] = t
            # the trainer's sequence around the exchange: norm -> (clip) -> step -> zero_grad
            if not unavailable:
                opt.zero_grads_in_step = True
                noise = torch.randn(n, device=device, dtype=torch.bfloat16)

                def full_step():
                    opt.clip_grad_norm_(1.0)
                    opt.step()
                    opt.zero_grad()
                res[comm + "_full"] = timeit(full_step, a.iters, 3, device,
                                             prep=(lambda: p.grad.copy_(noise)) if BURST == 1 else None)
                if comm.startswith("fused"):
                    opt._backend._ctx.check_status()
                del noise
            del opt, p
            gc.collect()
            torch.cuda.empty_cache()
        p = torch.randn(n, device=device, dtype=torch.bfloat16)
        g = torch.randn(n, device=device, dtype=torch.bfloat16)
        m = torch.zeros_like(p)

        def ddp_step():
            dist.all_reduce(g)
            g.div_(W)
            lion_ops.local_step(p, g, m, 1e-4, 0.1, 0.9, 0.99)
        res["allreduce"] = timeit(ddp_step, a.iters, 3, device)

        def ddp_full():
            dist.all_reduce(g)
            g.div_(W)
            bucket_ops.clip_grad_norm_flat_([g], 1.0)
            lion_ops.local_step(p, g, m, 1e-4, 0.1, 0.9, 0.99)
            g.zero_()
        noise = torch.randn(n, device=device, dtype=torch.bfloat16)
        res["allreduce_full"] = timeit(ddp_full, a.iters, 3, device, prep=(lambda: g.copy_(noise)) if BURST == 1 else None)
        del noise
        if ref is not None and mp <= a.skip_reference_above:
            res["reference"] = timeit(lambda: ref.update_fn_distributed(p, g, m, 1e-4, 0.1, 0.9, 0.99), max(a.iters // 2, 2),
                                      1, device)
        del p, g, m
        gc.collect()
        torch.cuda.empty_cache()
        t_hbm = (10 * n + W * n / 8) / (hbm * 1e9) * 1e3
        t_nvl = ((W - 1) * n / 8) / (nvl * 1e9) * 1e3
        res["roofline_ms"] = max(t_hbm, t_nvl)
        res["fused_frac_of_roofline"] = res["roofline_ms"] / res["fused"]
        if res.get("fused_multicast"):
            res["fused_multicast_frac_of_roofline"] = res["roofline_ms"] / res["fused_multicast"]
        res["roofline_full_ms"] = max((14 * n + (W + 1) * n / 8) / (hbm * 1e9) * 1e3, t_nvl)
        for k in ("fused_full", "fused_multicast_full"):
            if res.get(k):
                res[k + "_frac_of_roofline"] = res["roofline_full_ms"] / res[k]
        clocks = sampler.stop()
        # worst case over ranks: the lowest median clock, the union of throttle reasons
        mhz = torch.tensor([float(clocks["sm_mhz"] or 0)], device=device)
        dist.all_reduce(mhz, op=dist.ReduceOp.MIN)
        allr = [None] * W
        dist.all_gather_object(allr, clocks["reasons"])
        res["clocks"] = {"sm_mhz_min_over_ranks": float(mhz), "sm_max_mhz": clocks["sm_max_mhz"],
                         "reasons": sorted({r for rs in allr for r in rs}), "samples": clocks["samples"]}
        rows.append(res)
        if rank == 0:
            print(json.dumps(res))
    if rank == 0 and a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"hbm_gbs": hbm, "nvlink_gbs": nvl, "rows": rows}, indent=1))
    rt.shutdown()


if __name__ == "__main__":
    main()

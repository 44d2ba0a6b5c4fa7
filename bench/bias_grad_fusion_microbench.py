#!/usr/bin/env python
"""Fused producer + bias-gradient kernels against the separate passes they replace, at the GPT-2 micro-batch shapes
(20 x 1024 tokens): CUDA-event times (median of 20, 256 MB L2 flush between samples), achieved GB/s on the algorithmic
bytes and the fraction of the measured copy peak.  DLION_COLSUM_K = 1 | 2 | 4 picks the rows a thread keeps in flight.

    python bench/bias_grad_fusion_microbench.py [--out profiles/x.json]
"""
import argparse
import ctypes as C
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion_pytorch_b200.ops import _cabi  # noqa: E402
from distributed_lion_pytorch_b200.ops import linear as lin  # noqa: E402

dev = torch.device("cuda")
flush = None


def timeit(fn, n=20):
    global flush
    if flush is None:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(3):
        fn()
    ts = []
    for i in range(n):
        flush.fill_(i)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return sorted(ts)[n // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    L = _cabi.lib()
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    hbm = peaks.get("hbm_gbs", 6650.0)
    dt = torch.bfloat16
    code = _cabi.dtype_code(dt)
    st = _cabi.stream_ptr(dev)
    rows = 20 * 1024
    res = {"rows_in_flight": os.environ.get("DLION_COLSUM_K", "4"), "hbm_gbs": hbm}

    def row(name, us, nbytes):
        res[name] = {"us": round(us, 1), "gbs": round(nbytes / us / 1e3, 1), "frac_of_copy_peak": round(nbytes / us / 1e3 / hbm, 3)}

    # ---- c_fc: GELU backward + bias gradient, [20480, 3072]
    cols = 3072
    pre = torch.randn(rows, cols, device=dev, dtype=dt)
    dy = torch.randn(rows, cols, device=dev, dtype=dt)
    dpre = torch.empty_like(pre)
    out = torch.zeros(cols, device=dev, dtype=dt)
    partial = torch.empty(int(L.dlion_colsum_parts(rows, cols)), cols, dtype=torch.float32, device=dev)
    tick = lin._tickets(dev)
    nb = pre.numel() * 2
    row("gelu_bwd", timeit(lambda: L.dlion_gelu_tanh_bwd(pre.data_ptr(), dy.data_ptr(), dpre.data_ptr(), pre.numel(), code, st)), 3 * nb)
    row("colsum_3072", timeit(lambda: L.dlion_colsum_fused(dpre.data_ptr(), rows, cols, partial.data_ptr(), tick.data_ptr(),
                                                           out.data_ptr(), 1, code, st)), nb)
    row("gelu_bwd_colsum", timeit(lambda: L.dlion_gelu_tanh_bwd_colsum(pre.data_ptr(), dy.data_ptr(), dpre.data_ptr(), rows, cols,
                                                                       partial.data_ptr(), tick.data_ptr(), out.data_ptr(), 1,
                                                                       code, st)), 3 * nb)
    del pre, dy, dpre
    # ---- c_attn: dq, dk, dv -> d(qkv) + bias gradient, B=20 S=1024 H=12 D=64
    B, S, H, D = 20, 1024, 12, 64
    cols = 3 * H * D
    gs = [torch.randn(B, H, S, D, device=dev, dtype=dt) for _ in range(3)]
    dqkv = torch.empty(B, S, cols, device=dev, dtype=dt)
    out = torch.zeros(cols, device=dev, dtype=dt)
    partial = torch.empty(int(L.dlion_colsum_parts(B * S, cols)), cols, dtype=torch.float32, device=dev)
    strides = (C.c_int64 * 9)(*[g.stride(i) for g in gs for i in range(3)])
    nb = dqkv.numel() * 2
    row("stack_qkv", timeit(lambda: torch.stack([g.transpose(1, 2) for g in gs], dim=2, out=dqkv.view(B, S, 3, H, D))), 2 * nb)
    row("colsum_2304", timeit(lambda: L.dlion_colsum_fused(dqkv.data_ptr(), B * S, cols, partial.data_ptr(), tick.data_ptr(),
                                                           out.data_ptr(), 1, code, st)), nb)
    row("qkv_pack_colsum", timeit(lambda: L.dlion_qkv_pack_colsum(gs[0].data_ptr(), gs[1].data_ptr(), gs[2].data_ptr(), strides,
                                                                  dqkv.data_ptr(), B, S, H, D, partial.data_ptr(),
                                                                  tick.data_ptr(), out.data_ptr(), 0, code, st)), 2 * nb)
    del gs, dqkv
    # ---- c_proj: bias gradient of a [20480, 768] gradient (now a by-product of the norm backward)
    cols = 768
    x = torch.randn(rows, cols, device=dev, dtype=dt)
    out = torch.zeros(cols, device=dev, dtype=dt)
    partial = torch.empty(int(L.dlion_colsum_parts(rows, cols)), cols, dtype=torch.float32, device=dev)
    row("colsum_768", timeit(lambda: L.dlion_colsum_fused(x.data_ptr(), rows, cols, partial.data_ptr(), tick.data_ptr(),
                                                          out.data_ptr(), 1, code, st)), x.numel() * 2)
    print(json.dumps(res))
    if a.out:
        Path(a.out).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""GPT-2-small MLP (20480 tokens, 768 -> 3072 -> 768, bf16): the cuBLASLt-epilogue form against the
GEMM + separate GELU / column-sum kernels it replaces — numerics against an fp32 reference, then
CUDA-event times of forward and forward+backward (median of N, L2 flushed between iterations)."""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion_pytorch_b200.ops import lt_gemm  # noqa: E402
from distributed_lion_pytorch_b200.ops.activation import gelu_tanh  # noqa: E402
from distributed_lion_pytorch_b200.ops.linear import fused_linear  # noqa: E402


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
    torch.manual_seed(0)
    dt, dev = torch.bfloat16, "cuda"
    M, E, F_ = 20480, 768, 3072
    x = (torch.randn(M, E, device=dev)).to(dt).requires_grad_()
    params = [(torch.randn(E, F_, device=dev) * 0.02).to(dt), (torch.randn(F_, device=dev) * 0.02).to(dt),
              (torch.randn(F_, E, device=dev) * 0.02).to(dt), (torch.randn(E, device=dev) * 0.02).to(dt)]
    for p in params:
        p.requires_grad_()
        p.grad = torch.zeros_like(p)            # the steady state: gradient buffers exist (flat bucket views)
    dy = torch.randn(M, E, device=dev).to(dt)

    def unfused(x):
        w1, b1, w2, b2 = params
        return fused_linear(gelu_tanh(fused_linear(x, w1, b1, True)), w2, b2, True)

    def fused(x):
        w1, b1, w2, b2 = params
        h = lt_gemm.linear_gelu(x, w1, b1)
        return None if h is None else fused_linear(h, w2, b2, True)

    def ref32():
        xs = x.detach().float().requires_grad_()
        ps = [p.detach().float().requires_grad_() for p in params]
        y = torch.nn.functional.gelu(xs @ ps[0] + ps[1], approximate="tanh") @ ps[2] + ps[3]
        y.backward(dy.float())
        return y, xs.grad, [p.grad for p in ps]

    print("cuBLASLt available:", lt_gemm.available())
    y_ref, dx_ref, g_ref = ref32()
    report = {"lt_available": lt_gemm.available(), "numerics": {}, "times_us": {}}
    for name, fn in (("unfused", unfused), ("fused", fused)):
        for p in params:
            p.grad.zero_()
        x.grad = None
        y = fn(x)
        if y is None:
            print(name, "unavailable")
            continue
        y.backward(dy)
        errs = {"y": float((y.float() - y_ref).abs().max() / y_ref.abs().max()),
                "dx": float((x.grad.float() - dx_ref).abs().max() / dx_ref.abs().max())}
        for k, p, g in zip(("dw_fc", "db_fc", "dw_proj", "db_proj"), params, g_ref):
            errs[k] = float((p.grad.float() - g).abs().max() / g.abs().max())
        report["numerics"][name] = errs
        print(name, "max rel err vs fp32:", {k: f"{v:.2e}" for k, v in errs.items()})

        def step():
            x.grad = None
            fn(x).backward(dy)
        t_f = timeit(lambda: fn(x))
        t_fb = timeit(step)
        report["times_us"][name] = {"fwd": round(t_f, 1), "fwd_bwd": round(t_fb, 1)}
        print(f"{name:8s} fwd {t_f:8.1f} us   fwd+bwd {t_fb:8.1f} us")
    # the three epilogue GEMMs one by one against what they replace
    from distributed_lion_pytorch_b200.ops.linear import column_sum
    from distributed_lion_pytorch_b200.ops import _cabi
    w1, b1, w2, b2 = [p.detach() for p in params]
    xd = x.detach()
    pre = torch.addmm(b1, xd, w1)
    act = torch.nn.functional.gelu(pre, approximate="tanh")
    gw2, gb2, gb1 = torch.zeros_like(w2), torch.empty_like(b2), torch.empty_like(b1)
    dpre, out_act, out_pre = torch.empty_like(pre), torch.empty_like(pre), torch.empty_like(pre)

    def gelu_bwd(p_, d_, o_):
        _cabi.check(_cabi.lib().dlion_gelu_tanh_bwd(p_.data_ptr(), d_.data_ptr(), o_.data_ptr(), p_.numel(),
                                                    _cabi.dtype_code(p_.dtype), _cabi.stream_ptr(p_.device)), "gelu_bwd")
    pieces = {
        "fwd GELU_AUX_BIAS": (lambda: lt_gemm.gemm(xd, w1, out_act, epilogue=lt_gemm.GELU_AUX_BIAS, bias=b1, aux=out_pre),
                              lambda: gelu_tanh(torch.addmm(b1, xd, w1))),
        "wgrad BGRAD_B beta=1": (lambda: lt_gemm.gemm(act, dy, gw2, trans_a=True, beta=1.0, epilogue=lt_gemm.BGRAD_B, bias_grad=gb2),
                                 lambda: (gw2.addmm_(act.t(), dy), column_sum(dy, dt))),
        "wgrad BGRAD_B beta=0": (lambda: lt_gemm.gemm(act, dy, gw2, trans_a=True, beta=0.0, epilogue=lt_gemm.BGRAD_B, bias_grad=gb2),
                                 lambda: (torch.mm(act.t(), dy, out=gw2), column_sum(dy, dt))),
        "wgrad NONE beta=1 (lt)": (lambda: lt_gemm.gemm(act, dy, gw2, trans_a=True, beta=1.0),
                                   lambda: gw2.addmm_(act.t(), dy)),
        "dgrad DGELU_BGRAD": (lambda: lt_gemm.gemm(dy, w2, dpre, trans_b=True, epilogue=lt_gemm.DGELU_BGRAD, aux=pre, bias_grad=gb1),
                              lambda: (gelu_bwd(pre, dy @ w2.t(), dpre), column_sum(dpre, dt))),
        "dgrad NONE (lt)": (lambda: lt_gemm.gemm(dy, w2, dpre, trans_b=True),
                            lambda: torch.mm(dy, w2.t(), out=dpre)),
    }
    report["pieces_us"] = {}
    for k, (f_lt, f_base) in pieces.items():
        ok = f_lt()
        t_lt = timeit(f_lt) if ok else None
        t_b = timeit(f_base)
        report["pieces_us"][k] = {"lt": t_lt, "separate": t_b}
        print(f"{k:26s} lt {t_lt if t_lt is None else round(t_lt, 1)} us   separate kernels {t_b:8.1f} us")
    # sweep the heuristic's candidates for the backward epilogues
    sweeps = {
        "dgrad DGELU_BGRAD": lambda r: lt_gemm.gemm(dy, w2, dpre, trans_b=True, epilogue=lt_gemm.DGELU_BGRAD, aux=pre, bias_grad=gb1, algo_rank=r),
        "dgrad DGELU": lambda r: lt_gemm.gemm(dy, w2, dpre, trans_b=True, epilogue=lt_gemm.DGELU, aux=pre, algo_rank=r),
        "wgrad BGRAD_B beta=1": lambda r: lt_gemm.gemm(act, dy, gw2, trans_a=True, beta=1.0, epilogue=lt_gemm.BGRAD_B, bias_grad=gb2, algo_rank=r),
        "wgrad NONE beta=1": lambda r: lt_gemm.gemm(act, dy, gw2, trans_a=True, beta=1.0, algo_rank=r),
    }
    report["sweeps_us"] = {}
    for k, f_r in sweeps.items():
        res = []
        for r in range(12):
            if not f_r(r):
                break
            res.append(round(timeit(lambda: f_r(r), iters=6, warm=2), 1))
        report["sweeps_us"][k] = res
        print(f"{k:24s} by heuristic rank: {res}")
    report["refused_by_library"] = [str(s) for s in lt_gemm._unsupported]
    print("refused:", report["refused_by_library"])
    if a.out:
        Path(a.out).write_text(json.dumps(report, indent=1))


if __name__ == "__main__":
    main()

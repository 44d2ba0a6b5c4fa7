"""Plain-PyTorch restatements of the optimizer kernels' arithmetic, for the GPU numerics tests.

Every intermediate is computed in float64 (so one fp32 ``fma`` = one rounding of an exact
product-sum) and rounded through the storage dtype at exactly the points where ATen's
elementwise kernels round — the same contract the CUDA kernels implement (csrc/common.cuh).
"""
import torch


def f32(x: float) -> float:
    return float(torch.tensor(x, dtype=torch.float32))


def rnd(x64: torch.Tensor, dtype) -> torch.Tensor:
    """double -> fp32 -> storage dtype -> double"""
    return x64.float().to(dtype).double()


def fma(a, b64, c64, dtype):
    """round_dtype(round_f32(a*b + c)) with a a python float already representable in fp32"""
    return rnd(a * b64 + c64, dtype)


def blend(m, g, beta1, dtype):
    b1, omb1 = f32(beta1), f32(1 - beta1)
    t = rnd(m.double() * b1, dtype)
    return fma(omb1, g.double(), t, dtype)


def momentum(m, g, beta2, dtype):
    b2, omb2 = f32(beta2), f32(1 - beta2)
    t = rnd(m.double() * b2, dtype)
    return fma(omb2, g.double(), t, dtype)


def apply(p, u64, lr, wd, dtype):
    decay, lr32 = f32(1 - lr * wd), f32(lr)
    p1 = rnd(p.double() * decay, dtype)
    return rnd(p1 - lr32 * u64, dtype)


def local_step(p, g, m, lr, wd, beta1, beta2):
    dt = p.dtype
    c = blend(m, g, beta1, dt)
    p2 = apply(p, torch.sign(c), lr, wd, dt)
    m2 = momentum(m, g, beta2, dt)
    return p2.to(dt), m2.to(dt)


def bits(m, g, beta1):
    return blend(m, g, beta1, m.dtype) > 0


def direction(cnt, L, vote, tie):
    cnt = cnt.double()
    if vote == "average":
        return ((2 * cnt - L).float() / float(L)).double()
    u = torch.where(2 * cnt > L, 1.0, -1.0).double()
    if tie == "zero":
        u = torch.where(2 * cnt == L, torch.zeros_like(u), u)
    return u


def distributed_step(ps, gs, ms, lr, wd, beta1, beta2, vote="majority", tie="neg", live_mask=None):
    """ps/gs/ms: lists (one per worker).  Returns (new ps, new ms, bits[W, n])."""
    W = len(ps)
    dt = ps[0].dtype
    B = torch.stack([bits(ms[r], gs[r], beta1) for r in range(W)])
    live = [r for r in range(W) if live_mask is None or (live_mask >> r) & 1]
    cnt = B[live].sum(0)
    u = direction(cnt, len(live), vote, tie)
    new_p = [apply(ps[r], u, lr, wd, dt).to(dt) for r in range(W)]
    new_m = [momentum(ms[r], gs[r], beta2, dt).to(dt) for r in range(W)]
    return new_p, new_m, B


def mismatch(a: torch.Tensor, b: torch.Tensor):
    """(fraction of differing elements, max abs difference)"""
    a64, b64 = a.double(), b.double()
    bad = a64 != b64
    return bad.double().mean().item(), (a64 - b64).abs().max().item()

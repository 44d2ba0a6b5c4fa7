"""GPU tier: the fused sign-pack-exchange-vote-update kernel (N1/N7) with all W ranks
emulated as CTA groups of ONE cooperative launch on one GPU — the full protocol (TMA ring,
peer-arena stores, release/acquire progress flags, parity double-buffering, vote, update)
runs exactly as on W GPUs, except that the "peer" pointers are local."""
import pytest
import torch

import gpuref
from distributed_lion_pytorch_b200.optim import semantics as sem

pytestmark = pytest.mark.gpu
HP = dict(lr=1e-3, wd=0.1, beta1=0.9, beta2=0.99)


def _stack(W, n, dtype, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    base = torch.randn(n, device="cuda", generator=g).to(dtype)
    P = base.repeat(W, 1).contiguous()
    G = torch.randn(W, n, device="cuda", generator=g).to(dtype)
    M = (torch.randn(W, n, device="cuda", generator=g) * 0.1).to(dtype)
    G[:, ::19] = 0
    M[:, ::19] = 0
    return P, G, M


def _check(P, M, rp, rm, dtype):
    W = P.shape[0]
    for r in range(W):
        fp, _ = gpuref.mismatch(P[r], rp[r])
        fm, _ = gpuref.mismatch(M[r], rm[r])
        assert fp <= 2e-5 and fm <= 2e-5, (r, fp, fm)
    for r in range(1, W):
        assert torch.equal(P[0], P[r]), "replicas diverged"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("W,grid", [(2, None), (4, 5), (8, None), (8, 3), (16, 2)])
def test_fused_emulated_majority_multi_step(dtype, W, grid):
    from distributed_lion_pytorch_b200.ops import lion_ops
    from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange
    tile = lion_ops.fused_tile(dtype)
    n = tile * 153                     # uneven tile counts across CTAs; small grids wrap the 8-stage ring
    ex = EmulatedExchange(W, n, dtype, "cuda", grid=grid)
    P, G, M = _stack(W, n, dtype, seed=W)
    for step in range(3):               # three launches: both parities, flag epochs advance
        rp, rm, _ = gpuref.distributed_step(list(P), list(G), list(M), **HP)
        ex.step(P, G, M, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], spin_limit=1 << 22)
        torch.cuda.synchronize()
        ex.check_status()
        _check(P, M, rp, rm, dtype)
        G = torch.randn_like(G.float()).to(dtype)


@pytest.mark.parametrize("vote,tie,mask", [("majority", "zero", None), ("average", "neg", None),
                                           ("majority", "neg", 0b0110), ("average", "neg", 0b1101)])
def test_fused_emulated_modes(vote, tie, mask):
    from distributed_lion_pytorch_b200.ops import lion_ops
    from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange
    W, dtype = 4, torch.bfloat16
    n = lion_ops.fused_tile(dtype) * 40
    ex = EmulatedExchange(W, n, dtype, "cuda")
    P, G, M = _stack(W, n, dtype, seed=11)
    rp, rm, _ = gpuref.distributed_step(list(P), list(G), list(M), vote=vote, tie=tie, live_mask=mask, **HP)
    ex.step(P, G, M, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], vote=vote, tie=tie, live_mask=mask,
            spin_limit=1 << 22)
    torch.cuda.synchronize()
    ex.check_status()
    _check(P, M, rp, rm, dtype)


def test_fused_small_bucket_and_single_tile():
    from distributed_lion_pytorch_b200.ops import lion_ops
    from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange
    W, dtype = 2, torch.float32
    n = lion_ops.fused_tile(dtype)       # one tile: most CTAs have nothing to do
    ex = EmulatedExchange(W, n, dtype, "cuda")
    P, G, M = _stack(W, n, dtype, seed=5)
    rp, rm, _ = gpuref.distributed_step(list(P), list(G), list(M), **HP)
    ex.step(P, G, M, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], spin_limit=1 << 22)
    torch.cuda.synchronize()
    ex.check_status()
    _check(P, M, rp, rm, dtype)


def test_fused_stochastic_equals_split_kernels():
    """N7: the fused kernel's Philox stream is keyed by (seed, step, rank, element), so it must
    produce exactly the ballots of sign_pack + vote_update."""
    from distributed_lion_pytorch_b200.ops import lion_ops
    from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange
    W, dtype = 4, torch.bfloat16
    n = lion_ops.fused_tile(dtype) * 24
    ex = EmulatedExchange(W, n, dtype, "cuda")
    P, G, M = _stack(W, n, dtype, seed=21)
    P2, M2 = P.clone(), M.clone()
    ex.step(P, G, M, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], max_grad_norm=1.0, seed=9, step=5,
            spin_limit=1 << 22)
    gathered = torch.empty(W, n // 8, dtype=torch.uint8, device="cuda")
    for r in range(W):
        lion_ops.sign_pack(M2[r], G[r], HP["beta1"], gathered[r], max_grad_norm=1.0, seed=9, step=5, rank=r)
    for r in range(W):
        lion_ops.vote_update(P2[r], G[r], M2[r], gathered, W, None, HP["lr"], HP["wd"], HP["beta2"])
    torch.cuda.synchronize()
    ex.check_status()
    assert torch.equal(P, P2) and torch.equal(M, M2)


def test_fused_reports_missing_rank_instead_of_hanging():
    """Fault injection: rank 1's CTAs never run (grid covers only rank 0) — the voter's bounded
    flag wait must give up, record who was missing, and let the kernel finish."""
    from distributed_lion_pytorch_b200.ops import _cabi, lion_ops
    dtype = torch.float32
    n = lion_ops.fused_tile(dtype) * 4
    W = 2
    arena = torch.zeros(2 * W * (n // 32), dtype=torch.int32, device="cuda")
    flags = torch.zeros(W * _cabi.FUSED_MAX_CTAS, dtype=torch.int64, device="cuda")
    status = torch.zeros(4, dtype=torch.int32, device="cuda")
    p = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda")
    # a real (non-emulated) rank 0 of a 2-rank world whose peer never shows up; both "peers" map to local memory
    lion_ops.fused_step(p, g, m, world=W, rank=0, arena_ptrs=[arena.data_ptr()] * W, flag_ptrs=[flags.data_ptr()] * W,
                        arena_words=n // 32, epoch=1, grid=4, lr=1e-3, wd=0.0, beta1=0.9, beta2=0.99,
                        status=status, spin_limit=20000)
    torch.cuda.synchronize()
    st = status.tolist()
    assert st[0] == 1 and st[2] == 1, st          # code 1 = flag timeout, source rank 1
    # and with the absent rank excluded from both masks the same launch completes cleanly
    status.zero_()
    lion_ops.fused_step(p, g, m, world=W, rank=0, arena_ptrs=[arena.data_ptr()] * W, flag_ptrs=[flags.data_ptr()] * W,
                        arena_words=n // 32, epoch=2, grid=4, lr=1e-3, wd=0.0, beta1=0.9, beta2=0.99,
                        status=status, spin_limit=20000, live_mask=0b01, wait_mask=0b01)
    torch.cuda.synchronize()
    assert status.tolist()[0] == 0

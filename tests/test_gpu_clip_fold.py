"""GPU tier: gradient clipping and the gradient zero-fill folded into the optimizer kernels.

Contract (include/dlion.h): with ``clip=(sumsq, max_norm)`` a kernel must produce — bit for bit — what
``dlion_clip_scale`` followed by the plain kernel produces, without rewriting g; with ``zero_g`` it must leave
g zeroed.  reference: HF Trainer's ``clip_grad_norm_`` before ``optimizer.step()`` and ``zero_grad()`` after it,
which upstream inherits (run_clm.py:588-622)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
HP = dict(lr=1e-3, wd=0.1, beta1=0.9, beta2=0.99)
DTYPES = [torch.bfloat16, torch.float32, torch.float16]


def _data(n, dtype, seed, scale=3.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    p = torch.randn(n, device="cuda", generator=g).to(dtype)
    gr = (torch.randn(n, device="cuda", generator=g) * scale).to(dtype)
    m = (torch.randn(n, device="cuda", generator=g) * 0.1).to(dtype)
    gr[::17] = 0
    m[::17] = 0
    return p, gr, m


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("n", [4099, (1 << 20) + 5])
@pytest.mark.parametrize("max_norm", [1.0, 1e9])          # clipping active / norm below the limit
def test_local_step_with_folded_clip_and_zero(dtype, n, max_norm):
    from distributed_lion_pytorch_b200.ops import bucket_ops, lion_ops
    p, g, m = _data(n, dtype, seed=n)
    p2, g2, m2 = p.clone(), g.clone(), m.clone()
    ss = bucket_ops.sumsq(g)
    # separate passes
    bucket_ops.clip_scale_(g2, ss, max_norm)
    lion_ops.local_step(p2, g2, m2, **{k: HP[k] for k in ("lr", "wd", "beta1", "beta2")})
    # folded
    g_before = g.clone()
    lion_ops.local_step(p, g, m, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], clip=(ss, max_norm), zero_g=False)
    torch.cuda.synchronize()
    assert torch.equal(p, p2) and torch.equal(m, m2)
    assert torch.equal(g, g_before), "the folded clip must not rewrite g"
    lion_ops.local_step(p, g, m, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], clip=(ss, max_norm), zero_g=True)
    torch.cuda.synchronize()
    assert int(g.count_nonzero()) == 0


@pytest.mark.parametrize("dtype", DTYPES)
def test_split_pair_with_folded_clip_and_zero(dtype):
    from distributed_lion_pytorch_b200.ops import bucket_ops, lion_ops
    W, n = 4, (1 << 18) + 24
    P, G, M = zip(*[_data(n, dtype, seed=100 + r) for r in range(W)])
    P = [P[0].clone() for _ in range(W)]
    ss = [bucket_ops.sumsq(G[r]) for r in range(W)]
    rows = (n + 31) // 32 * 4
    ref_words = torch.zeros(W, rows, dtype=torch.uint8, device="cuda")
    new_words = torch.zeros(W, rows, dtype=torch.uint8, device="cuda")
    G2 = [g.clone() for g in G]
    for r in range(W):
        bucket_ops.clip_scale_(G2[r], ss[r], 1.0)
        lion_ops.sign_pack(M[r], G2[r], HP["beta1"], ref_words[r])
        lion_ops.sign_pack(M[r], G[r], HP["beta1"], new_words[r], clip=(ss[r], 1.0))
    torch.cuda.synchronize()
    assert torch.equal(ref_words, new_words)
    for r in range(W):
        p2, m2 = P[r].clone(), M[r].clone()
        lion_ops.vote_update(p2, G2[r], m2, ref_words, W, None, HP["lr"], HP["wd"], HP["beta2"])
        lion_ops.vote_update(P[r], G[r], M[r], new_words, W, None, HP["lr"], HP["wd"], HP["beta2"],
                             clip=(ss[r], 1.0), zero_g=True)
        torch.cuda.synchronize()
        assert torch.equal(P[r], p2) and torch.equal(M[r], m2)
        assert int(G[r].count_nonzero()) == 0


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("W,grid", [(2, None), (4, 5), (8, 3)])
def test_fused_with_folded_clip_and_zero_equals_separate_passes(dtype, W, grid):
    """The fused kernel with clip + zero folded in == clip_scale, fused kernel, memset (emulated ranks).  Every rank
    clips with its OWN norm (gradients are local in Distributed Lion), passed here through one stacked call per
    variant: the emulated launch takes a single sumsq pointer, so all ranks share one norm in this test."""
    from distributed_lion_pytorch_b200.ops import bucket_ops, lion_ops
    from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange
    tile = lion_ops.fused_tile(dtype)
    n = tile * 77
    g = torch.Generator(device="cuda").manual_seed(W)
    base = torch.randn(n, device="cuda", generator=g).to(dtype)
    P = base.repeat(W, 1).contiguous()
    G = (torch.randn(W, n, device="cuda", generator=g) * 2).to(dtype)
    M = (torch.randn(W, n, device="cuda", generator=g) * 0.1).to(dtype)
    P2, G2, M2 = P.clone(), G.clone(), M.clone()
    ex_a = EmulatedExchange(W, n, dtype, "cuda", grid=grid)
    ex_b = EmulatedExchange(W, n, dtype, "cuda", grid=grid)
    for step in range(3):
        ss = bucket_ops.sumsq(G.view(-1))
        max_norm = float(ss.sqrt()) * 0.37
        bucket_ops.clip_scale_(G2.view(-1), ss, max_norm)
        ex_a.step(P2, G2, M2, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], spin_limit=1 << 22)
        ex_b.step(P, G, M, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], spin_limit=1 << 22, clip=(ss, max_norm),
                  zero_g=True)
        torch.cuda.synchronize()
        ex_a.check_status()
        ex_b.check_status()
        assert torch.equal(P, P2) and torch.equal(M, M2), step
        assert int(G.count_nonzero()) == 0, "zero_g must leave the gradient bucket zeroed"
        for r in range(1, W):
            assert torch.equal(P[0], P[r])
        G = (torch.randn(W, n, device="cuda", generator=g) * 2).to(dtype)
        G2 = G.clone()


def test_lion_clip_grad_norm_defers_and_matches_eager_clip():
    """Optimizer level: ``clip_grad_norm_`` + ``step()`` (folded) == torch's clip + ``step()`` on a twin model;
    after a step with ``zero_grads_in_step`` the gradient bucket is clean and ``zero_grad()`` launches nothing."""
    from distributed_lion import Lion
    from distributed_lion_pytorch_b200.ops import _cabi

    def make():
        torch.manual_seed(0)
        return torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.GELU(), torch.nn.Linear(256, 64)).cuda().bfloat16()

    m1, m2 = make(), make()
    o1 = Lion(m1.parameters(), lr=1e-3, weight_decay=0.1, zero_grads_in_step=True)
    o2 = Lion(m2.parameters(), lr=1e-3, weight_decay=0.1)
    x = torch.randn(32, 64, device="cuda", dtype=torch.bfloat16)
    for step in range(4):
        for m in (m1, m2):
            (m(x).float().pow(2).mean() * 50).backward()
        n1 = o1.clip_grad_norm_(1.0)
        if step > 0:
            assert o1._pending_clip is not None, "steady state must defer the scaling to step()"
        n2 = torch.nn.utils.clip_grad_norm_(list(m2.parameters()), 1.0)
        assert torch.allclose(n1.float().cpu(), n2.float().cpu(), rtol=1e-2)
        o1.step()
        o2.step()
        if step > 0:
            assert all(int(b.flat_g.count_nonzero()) == 0 for b in o1.buckets)
            before = _cabi.LAUNCHES
        o1.zero_grad()
        o2.zero_grad()
    # bf16 clip coefficient: torch multiplies by a tensor coefficient computed from a norm that sums in a different
    # order — compare to one bf16 ulp, and require bit-equality of the two Lion paths in the kernel tests above
    for a, b in zip(m1.parameters(), m2.parameters()):
        assert torch.allclose(a.float(), b.float(), rtol=2 ** -6, atol=2e-3)

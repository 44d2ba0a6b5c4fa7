"""GPU tier: opt-in stochastic rounding of the bf16 parameter write — the local, split (vote_update) and fused
kernels against the torch oracle (``optim.semantics``), bit for bit, and against each other."""
import pytest
import torch

from distributed_lion_pytorch_b200.optim import semantics as sem

pytestmark = pytest.mark.gpu
HP = dict(lr=1e-4, wd=0.1, beta1=0.9, beta2=0.99)


def _pgm(n, seed, scale=0.05):
    g = torch.Generator(device="cuda").manual_seed(seed)
    p = (torch.randn(n, device="cuda", generator=g) * scale).bfloat16()
    gr = torch.randn(n, device="cuda", generator=g).bfloat16()
    m = (torch.randn(n, device="cuda", generator=g) * 0.1).bfloat16()
    gr[::17] = 0
    m[::17] = 0
    return p, gr, m


@pytest.mark.parametrize("n,elem_base", [(8192 * 5, 0), (8192 * 3 + 8 * 7, 1 << 20), (64, 8)])
def test_local_kernel_matches_oracle_bit_for_bit(n, elem_base):
    from distributed_lion_pytorch_b200.ops import lion_ops
    key = sem.sr_key(3, 9)
    p, g, m = _pgm(n, 1)
    p_ref, m_ref = p.clone(), m.clone()
    sem.local_lion_step_(p_ref, g, m_ref, **HP, sr_key=key, elem_base=elem_base)
    lion_ops.local_step(p, g, m, **HP, sr_key=key, elem_base=elem_base)
    assert torch.equal(p, p_ref) and torch.equal(m, m_ref)
    # and it differs from round-to-nearest: most sub-ulp updates are no longer swallowed
    p_rn, _, m_rn = _pgm(n, 1)
    lion_ops.local_step(p_rn, g, m_rn, **HP)
    assert not torch.equal(p, p_rn) and torch.equal(m, m_rn)          # momentum is unaffected
    # a different step draws different noise
    p2, _, m2 = _pgm(n, 1)
    lion_ops.local_step(p2, g, m2, **HP, sr_key=sem.sr_key(3, 10), elem_base=elem_base)
    assert not torch.equal(p2, p)


def test_sr_ignored_for_fp32_and_fp16():
    from distributed_lion_pytorch_b200.ops import lion_ops
    for dtype in (torch.float32, torch.float16):
        p, g, m = (t.to(dtype) for t in _pgm(8192, 2))
        a, b = p.clone(), p.clone()
        lion_ops.local_step(a, g, m.clone(), **HP, sr_key=sem.sr_key(1, 1))
        lion_ops.local_step(b, g, m.clone(), **HP)
        assert torch.equal(a, b)


@pytest.mark.parametrize("W,mask", [(4, None), (8, 0b10110111)])
def test_split_and_fused_kernels_agree_with_oracle_under_sr(W, mask):
    from distributed_lion_pytorch_b200.ops import lion_ops
    from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange
    dtype = torch.bfloat16
    n = lion_ops.fused_tile(dtype) * 37
    g0 = torch.Generator(device="cuda").manual_seed(W)
    base = (torch.randn(n, device="cuda", generator=g0) * 0.05).to(dtype)
    P = base.repeat(W, 1).contiguous()
    G = torch.randn(W, n, device="cuda", generator=g0).to(dtype)
    M = (torch.randn(W, n, device="cuda", generator=g0) * 0.1).to(dtype)
    ex = EmulatedExchange(W, n, dtype, "cuda")
    for step in range(3):
        key = sem.sr_key(7, step)
        # oracle: all workers in one process, torch ops only
        ps, ms = [P[r].clone() for r in range(W)], [M[r].clone() for r in range(W)]
        gathered = torch.stack([sem.pack_bits(sem.local_bits(M[r], G[r], HP["beta1"])) for r in range(W)])
        cnt = sem.count_votes(gathered, n, mask)
        L = len(sem.live_ranks(W, mask))
        for r in range(W):
            sem.apply_vote_(ps[r], G[r], ms[r], cnt, L, HP["lr"], HP["wd"], HP["beta2"], sr_key=key)
        # split path kernels on rank 0's tensors
        p_s, m_s = P[0].clone(), M[0].clone()
        wire = torch.zeros(W, n // 8, dtype=torch.uint8, device="cuda")
        for r in range(W):
            lion_ops.sign_pack(M[r], G[r], HP["beta1"], wire[r])
        lion_ops.vote_update(p_s, G[0], m_s, wire, W, mask, HP["lr"], HP["wd"], HP["beta2"], sr_key=key)
        # fused kernel, W ranks emulated in one launch
        ex.step(P, G, M, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"], live_mask=mask, sr_key=key, spin_limit=1 << 22)
        torch.cuda.synchronize()
        ex.check_status()
        assert torch.equal(p_s, ps[0]) and torch.equal(m_s, ms[0])
        for r in range(W):
            assert torch.equal(P[r], ps[r]), f"fused kernel, rank {r}, step {step}"
            assert torch.equal(M[r], ms[r])
            assert torch.equal(P[r], P[0])
        G = torch.randn_like(G.float()).to(dtype)


def test_lion_optimizer_with_stochastic_rounding_on_gpu():
    from distributed_lion import Lion
    torch.manual_seed(0)
    w = torch.nn.Parameter(torch.ones(1 << 16, device="cuda", dtype=torch.bfloat16))
    opt = Lion([w], lr=1e-4, stochastic_rounding=True, seed=3)
    for _ in range(32):
        w.grad = torch.ones_like(w)
        opt.step()
    drift = 1.0 - float(w.detach().float().mean())
    assert abs(drift - 32 * 1e-4) < 0.15 * 32 * 1e-4, drift              # the mean follows the fp32 trajectory
    w2 = torch.nn.Parameter(torch.ones(1 << 16, device="cuda", dtype=torch.bfloat16))
    opt2 = Lion([w2], lr=1e-4)
    for _ in range(32):
        w2.grad = torch.ones_like(w2)
        opt2.step()
    assert float(w2.detach().float().mean()) == 1.0                       # round-to-nearest: nothing ever moves

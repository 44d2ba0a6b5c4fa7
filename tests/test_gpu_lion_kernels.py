"""GPU tier: each optimizer kernel against a plain-PyTorch restatement of the same op
(tests/gpuref.py), odd sizes and ragged tails included."""
import pytest
import torch

import gpuref
from distributed_lion_pytorch_b200.optim import semantics as sem

pytestmark = pytest.mark.gpu
DTYPES = [torch.float32, torch.bfloat16]
HP = dict(lr=1e-3, wd=0.1, beta1=0.9, beta2=0.99)


def _data(n, dtype, seed=0, zeros=True):
    g = torch.Generator(device="cuda").manual_seed(seed)
    p = torch.randn(n, device="cuda", generator=g).to(dtype)
    gr = torch.randn(n, device="cuda", generator=g).to(dtype)
    m = (torch.randn(n, device="cuda", generator=g) * 0.1).to(dtype)
    if zeros and n > 40:
        gr[::17] = 0
        m[::17] = 0          # exact-zero blend: sign(0) = 0 locally, bit 0 on the wire
    return p, gr, m


def _close(a, b, dtype, what):
    frac, mx = gpuref.mismatch(a, b)
    # the restatement rounds a double sum twice; allow a vanishing fraction of 1-ulp wobbles
    assert frac <= 2e-5, f"{what}: {frac:.2e} of elements differ (max {mx:.3e})"
    tol = 1e-6 if dtype == torch.float32 else 4e-2
    assert torch.allclose(a.float(), b.float(), rtol=tol, atol=1e-6), what


@pytest.mark.parametrize("dtype", DTYPES + [torch.float16])
@pytest.mark.parametrize("n", [8, 1000, 4099, (1 << 20) + 5])
def test_local_step_matches_reference(dtype, n):
    from distributed_lion_pytorch_b200.ops import lion_ops
    p, g, m = _data(n, dtype, seed=n)
    rp, rm = gpuref.local_step(p, g, m, **HP)
    lion_ops.local_step(p, g, m, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"])
    torch.cuda.synchronize()
    _close(p, rp, dtype, "p")
    _close(m, rm, dtype, "m")


@pytest.mark.parametrize("dtype", DTYPES)
def test_local_step_vs_aten_ops(dtype):
    """The kernel must agree with the reference's own ATen op sequence on the same GPU
    (distributed_lion.py:47-59) — exactly in fp32, to one bf16 ulp in bf16."""
    from distributed_lion_pytorch_b200.ops import lion_ops
    n = 1 << 18
    p, g, m = _data(n, dtype, seed=3)
    pa, ma = p.clone(), m.clone()
    sem.local_lion_step_(pa, g, ma, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"])
    lion_ops.local_step(p, g, m, HP["lr"], HP["wd"], HP["beta1"], HP["beta2"])
    fp, _ = gpuref.mismatch(p, pa)
    fm, _ = gpuref.mismatch(m, ma)
    print(f"[aten-parity {dtype}] p mismatch {fp:.3e}  m mismatch {fm:.3e}")
    if dtype == torch.float32:
        assert fp <= 1e-5 and fm <= 1e-5
    assert torch.allclose(p.float(), pa.float(), rtol=2 ** -7, atol=1e-6)
    assert torch.allclose(m.float(), ma.float(), rtol=2 ** -7, atol=1e-6)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("n", [1, 31, 32, 33, 4096, 100003])
def test_sign_pack_matches_reference(dtype, n):
    from distributed_lion_pytorch_b200.ops import lion_ops
    n_alloc = max(n, 8)
    p, g, m = _data(n_alloc, dtype, seed=n)
    g, m = g[:n], m[:n]
    out = torch.full(((n + 31) // 32 * 4,), 0xAA, dtype=torch.uint8, device="cuda")
    lion_ops.sign_pack(m, g, HP["beta1"], out)
    want = sem.pack_bits(gpuref.bits(m, g, HP["beta1"]))
    got = out[: want.numel()]
    diff = (got != want).sum().item()
    assert diff <= max(1, n // 100000), f"{diff} packed bytes differ"
    assert int(out[want.numel():].sum()) == 0            # pad bytes of the last word are zero
    # bits of the padding inside the last byte are zero too
    full = sem.unpack_bits(out, out.numel() * 8)
    assert not full[n:].any()


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("W,vote,tie,mask", [(2, "majority", "neg", None), (4, "majority", "zero", None),
                                             (8, "majority", "neg", None), (8, "average", "neg", None),
                                             (8, "majority", "neg", 0b10110101), (3, "average", "neg", 0b101),
                                             (16, "majority", "zero", None)])
def test_vote_update_matches_reference(dtype, W, vote, tie, mask):
    from distributed_lion_pytorch_b200.ops import lion_ops
    n = 50001
    p, g, m = _data(n, dtype, seed=W)
    gen = torch.Generator(device="cuda").manual_seed(W * 7)
    row = (n + 31) // 32 * 4
    gathered = torch.randint(0, 256, (W, row), dtype=torch.uint8, device="cuda", generator=gen)
    cnt = sem.count_votes(gathered, n, mask)
    L = len(sem.live_ranks(W, mask))
    u = gpuref.direction(cnt, L, vote, tie)
    rp = gpuref.apply(p, u, HP["lr"], HP["wd"], dtype).to(dtype)
    rm = gpuref.momentum(m, g, HP["beta2"], dtype).to(dtype)
    lion_ops.vote_update(p, g, m, gathered, W, mask, HP["lr"], HP["wd"], HP["beta2"], vote, tie)
    _close(p, rp, dtype, "p")
    _close(m, rm, dtype, "m")


def test_stochastic_pack_statistics_and_determinism():
    from distributed_lion_pytorch_b200.ops import lion_ops
    n = 1 << 20
    m = torch.zeros(n, device="cuda")
    g = torch.full((n,), 0.5, device="cuda")
    out1 = torch.empty(n // 8, dtype=torch.uint8, device="cuda")
    out2 = torch.empty_like(out1)
    out3 = torch.empty_like(out1)
    lion_ops.sign_pack(m, g, 0.9, out1, max_grad_norm=1.0, seed=7, step=3, rank=1)
    lion_ops.sign_pack(m, g, 0.9, out2, max_grad_norm=1.0, seed=7, step=3, rank=1)
    lion_ops.sign_pack(m, g, 0.9, out3, max_grad_norm=1.0, seed=7, step=4, rank=1)
    assert torch.equal(out1, out2) and not torch.equal(out1, out3)
    r = (1 + 1 / 0.9) * 1.0
    want = (0.05 + r) / (2 * r)
    got = sem.unpack_bits(out1, n).float().mean().item()
    assert abs(got - want) < 3e-3


def test_bucket_ops():
    from distributed_lion_pytorch_b200.ops import bucket_ops
    gen = torch.Generator(device="cuda").manual_seed(0)
    sizes = [7, 768, 50257 * 3, 1, 4096, 1000] * 20          # > 96 tensors: two launches
    srcs = [torch.randn(s, device="cuda", generator=gen).bfloat16() for s in sizes]
    flat = torch.zeros(sum((s + 127) // 128 * 128 for s in sizes), dtype=torch.bfloat16, device="cuda")
    dsts, off = [], 0
    for s in sizes:
        dsts.append(flat[off:off + s])
        off += (s + 127) // 128 * 128
    assert bucket_ops.multi_tensor_gather(srcs, dsts)
    for s, d in zip(srcs, dsts):
        assert torch.equal(s, d)
    # cast variant
    d32 = [torch.empty(s, device="cuda") for s in sizes[:5]]
    assert bucket_ops.multi_tensor_gather(srcs[:5], d32)
    for s, d in zip(srcs[:5], d32):
        assert torch.equal(s.float(), d)
    for dtype in DTYPES:
        x = torch.randn((1 << 22) + 3, device="cuda", generator=gen).to(dtype)
        for _ in range(2):                                     # scratch is reusable back to back
            ss = bucket_ops.sumsq(x)
            assert torch.allclose(ss, x.double().pow(2).sum().float().view(1), rtol=1e-5)
        y = x.clone()
        bucket_ops.clip_scale_(y, ss, 1.0)
        coef = min(1.0, 1.0 / (float(ss.sqrt()) + 1e-6))
        assert torch.allclose(y.float(), x.float() * coef, rtol=1e-2 if dtype == torch.bfloat16 else 1e-6, atol=1e-8)
        z = x.clone()
        bucket_ops.clip_scale_(z, ss, 1e9)                     # no clipping needed: untouched
        assert torch.equal(z, x)


def _guarded(n, dtype, fill, guard=4096):
    """A length-n view in the middle of a larger buffer whose margins hold a canary value."""
    buf = torch.full((n + 2 * guard,), fill, dtype=dtype, device="cuda")
    return buf, buf[guard:guard + n]


def _canaries_intact(buf, n, fill, guard=4096):
    return bool((buf[:guard] == fill).all()) and bool((buf[guard + n:] == fill).all())


@pytest.mark.parametrize("dtype", DTYPES)
def test_kernels_respect_buffer_bounds(dtype):
    """compute-sanitizer is closed on this pool, so out-of-bounds writes are caught the slow way:
    every output lives between canary margins that must survive the launch (ragged sizes included)."""
    from distributed_lion_pytorch_b200.ops import bucket_ops, lion_ops
    from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange
    for n in (8, 1003, 4096 * 3 + 8 * 5):
        pb, p = _guarded(n, dtype, 7.0)
        mb, m = _guarded(n, dtype, 7.0)
        g = torch.randn(n, device="cuda").to(dtype)
        p.copy_(torch.randn(n, device="cuda"))
        m.zero_()
        lion_ops.local_step(p, g, m, 1e-3, 0.1, 0.9, 0.99)
        wb, w = _guarded((n + 31) // 32 * 4, torch.uint8, 0x5A, guard=256)
        lion_ops.sign_pack(m, g, 0.9, w)
        gathered = torch.randint(0, 256, (4, (n + 31) // 32 * 4), dtype=torch.uint8, device="cuda")
        lion_ops.vote_update(p, g, m, gathered, 4, None, 1e-3, 0.1, 0.99)
        torch.cuda.synchronize()
        assert _canaries_intact(pb, n, 7.0) and _canaries_intact(mb, n, 7.0)
        assert bool((wb[:256] == 0x5A).all()) and bool((wb[256 + w.numel():] == 0x5A).all())
    # fused kernel: stacked [W, n] buckets between canaries, arenas sized exactly
    W = 4
    n = lion_ops.fused_tile(dtype) * 37
    Pb, Pf = _guarded(W * n, dtype, 7.0)
    Mb, Mf = _guarded(W * n, dtype, 7.0)
    P, M = Pf.view(W, n), Mf.view(W, n)
    P.copy_(torch.randn(n, device="cuda").to(dtype).repeat(W, 1))
    M.zero_()
    G = torch.randn(W, n, device="cuda").to(dtype)
    ex = EmulatedExchange(W, n, dtype, "cuda")
    ex.step(P, G, M, 1e-3, 0.1, 0.9, 0.99, spin_limit=1 << 22)
    torch.cuda.synchronize()
    ex.check_status()
    assert _canaries_intact(Pb, W * n, 7.0) and _canaries_intact(Mb, W * n, 7.0)
    # clip / sumsq on a ragged tail
    xb, x = _guarded(100003, dtype, 7.0)
    x.copy_(torch.randn(100003, device="cuda"))
    bucket_ops.clip_scale_(x, bucket_ops.sumsq(x), 0.5)
    torch.cuda.synchronize()
    assert _canaries_intact(xb, 100003, 7.0)

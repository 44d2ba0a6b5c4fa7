"""Oracle tier (CPU): the frozen semantics against the reference's own functions."""
import pytest
import torch
from hypothesis import given, settings, strategies as st

from distributed_lion_pytorch_b200.optim import semantics as sem


@pytest.mark.parametrize("n", [1, 7, 8, 9, 31, 32, 33, 255, 4097, 2 ** 16 + 5])
def test_pack_unpack_roundtrip(n):
    g = torch.Generator().manual_seed(n)
    bits = torch.rand(n, generator=g) > 0.5
    packed = sem.pack_bits(bits)
    assert packed.dtype == torch.uint8 and packed.numel() == (n + 7) // 8
    assert torch.equal(sem.unpack_bits(packed, n), bits)
    # pad bits are zero
    full = sem.unpack_bits(packed, packed.numel() * 8)
    assert not full[n:].any()


def test_lsb_first_bit_order(ref_mod):
    bits = torch.zeros(16, dtype=torch.bool)
    bits[0] = True      # element 0 -> byte 0 bit 0
    bits[9] = True      # element 9 -> byte 1 bit 1
    packed = sem.pack_bits(bits)
    assert packed.tolist() == [1, 2]
    # the reference's encoder (distributed_lion.py:75-77) agrees, modulo its int64 word
    padded, _ = ref_mod.flatten_and_pad(bits)
    idx = torch.arange(8).unsqueeze(0).unsqueeze(0)
    ref_words = (padded.view(-1, 8).unsqueeze(0).byte() << idx).sum(dim=-1)
    assert ref_words.dtype == torch.int64          # 8 bits/param upstream
    assert ref_words.flatten().tolist() == packed.tolist()


@settings(max_examples=60, deadline=None)
@given(W=st.integers(1, 8), n=st.integers(1, 70), seed=st.integers(0, 10 ** 6))
def test_majority_matches_reference_mode(W, n, seed):
    from conftest import load_reference_module
    ref = load_reference_module()
    if ref is None:
        pytest.skip("no reference")
    g = torch.Generator().manual_seed(seed)
    votes = [torch.rand(n, generator=g) > 0.5 for _ in range(W)]
    expect = ref.majority_vote(votes)                    # torch.mode; tie -> False
    gathered = torch.stack([sem.pack_bits(v) for v in votes])
    cnt = sem.count_votes(gathered, n)
    u = sem.vote_direction(cnt, W, "majority", "neg")
    assert torch.equal(u > 0, expect)
    assert set(u.unique().tolist()) <= {-1.0, 1.0}      # never 0 in reference mode


@pytest.mark.parametrize("W", [2, 4, 6, 8])
def test_tie_rules_and_average(W):
    cnt = torch.arange(0, W + 1, dtype=torch.int16)
    neg = sem.vote_direction(cnt, W, "majority", "neg")
    zero = sem.vote_direction(cnt, W, "majority", "zero")
    avg = sem.vote_direction(cnt, W, "average")
    assert neg[W // 2] == -1 and zero[W // 2] == 0
    assert torch.equal(neg[: W // 2], -torch.ones(W // 2)) and torch.equal(neg[W // 2 + 1:], torch.ones(W // 2))
    assert torch.allclose(avg, (2 * cnt.float() - W) / W)


@pytest.mark.parametrize("mask,W", [(0b0101, 4), (0b1110, 4), (0b10000001, 8), (0b1, 3)])
def test_live_mask_counts(mask, W):
    n = 100
    g = torch.Generator().manual_seed(mask)
    votes = [torch.rand(n, generator=g) > 0.5 for _ in range(W)]
    gathered = torch.stack([sem.pack_bits(v) for v in votes])
    cnt = sem.count_votes(gathered, n, mask)
    live = sem.live_ranks(W, mask)
    expect = sum(votes[r].to(torch.int16) for r in live)
    assert torch.equal(cnt, expect)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_local_step_bitexact_vs_reference(ref_mod, dtype):
    torch.manual_seed(0)
    p = torch.randn(1000).to(dtype)
    g = torch.randn(1000).to(dtype)
    m = (torch.randn(1000) * 0.1).to(dtype)
    g[::17] = 0
    m[::17] = 0      # sign(0) = 0 coordinates
    p2, m2 = p.clone(), m.clone()
    for _ in range(3):
        ref_mod.update_fn(p, g, m, 1e-3, 0.1, 0.9, 0.99)
        sem.local_lion_step_(p2, g, m2, 1e-3, 0.1, 0.9, 0.99)
    assert torch.equal(p, p2) and torch.equal(m, m2)


def test_stochastic_bits_probability():
    n = 200000
    m = torch.zeros(n)
    g = torch.full((n,), 0.5)
    gen = torch.Generator().manual_seed(1)
    beta1, mgn = 0.9, 1.0
    r = (1 + 1 / beta1) * mgn
    bits = sem.local_bits(m, g, beta1, max_grad_norm=mgn, generator=gen)
    c = 0.1 * 0.5
    assert abs(bits.float().mean().item() - (c + r) / (2 * r)) < 5e-3


def test_simulated_workers_stay_identical():
    torch.manual_seed(3)
    W, n = 4, 513
    base = torch.randn(n)
    ps = [base.clone() for _ in range(W)]
    ms = [torch.zeros(n) for _ in range(W)]
    for step in range(5):
        gs = [torch.randn(n) for _ in range(W)]
        sem.simulate_workers_step_(ps, gs, ms, 1e-2, 0.1, 0.9, 0.99)
    for r in range(1, W):
        assert torch.equal(ps[0], ps[r])
    assert not torch.equal(ms[0], ms[1])     # momentum is per worker, never communicated


@pytest.mark.parametrize("n", [1, 7, 8, 9, 63, 64, 65, 100003])
def test_native_cpu_codec_matches_oracle(n):
    from distributed_lion_pytorch_b200 import build
    from distributed_lion_pytorch_b200.ops import cpu_ops
    build.build_cpu()
    assert cpu_ops.lib() is not None
    g = torch.Generator().manual_seed(n)
    for W, mask in ((1, None), (4, 0b1011), (8, None), (16, 0xF0F1)):
        votes = [torch.rand(n, generator=g) > 0.5 for _ in range(W)]
        packed = [cpu_ops.pack_bits(v) for v in votes]
        for v, p in zip(votes, packed):
            assert torch.equal(p, sem.pack_bits(v))
        gathered = torch.stack(packed)
        assert torch.equal(cpu_ops.count_votes(gathered, n, mask), sem.count_votes(gathered, n, mask))


def test_stochastic_rounding_is_unbiased_exact_on_representables_and_deterministic():
    from distributed_lion_pytorch_b200.optim import semantics as sem
    key = sem.sr_key(seed=11, step=4)
    assert key != 0 and key == sem.sr_key(11, 4) and key != sem.sr_key(11, 5) and key != sem.sr_key(12, 4)
    n = 1 << 18
    for frac in (1 / 8, 1 / 2, 15 / 16):
        for sign in (1.0, -1.0):
            x = torch.full((n,), sign * (1.0 + frac * 2.0 ** -7))          # between the bf16 neighbours 1 and 1 + 2^-7
            y = sem.stochastic_round_bf16(x, key).float()
            assert set(y.abs().unique().tolist()) <= {1.0, 1.0 + 2.0 ** -7}
            up = float((y.abs() > 1.0).float().mean())
            assert abs(up - frac) < 0.01, (frac, up)                        # P(round away from zero) = discarded fraction
            assert abs(float(y.mean()) - float(x.mean())) < 2.0 ** -7 * 0.01
    exact = torch.tensor([0.0, 1.0, -2.5, 2.0 ** 127, float("inf"), float("-inf")]).repeat(100)
    assert torch.equal(sem.stochastic_round_bf16(exact, key).float(), exact.bfloat16().float())
    assert torch.isnan(sem.stochastic_round_bf16(torch.tensor([float("nan")]), key)).all()
    # layout invariance: element e of the bucket gets the same noise whether it is rounded in one call or in pieces
    x = torch.randn(4096) * 0.02
    whole = sem.stochastic_round_bf16(x, key)
    parts = torch.cat([sem.stochastic_round_bf16(x[:1024], key, 0), sem.stochastic_round_bf16(x[1024:], key, 1024)])
    assert torch.equal(whole, parts)


def test_stochastic_rounding_lets_sub_ulp_updates_accumulate():
    """lr = 1e-4 on weights of magnitude 1: round-to-nearest never moves them, stochastic rounding moves their mean by
    lr per step — the reason the option exists."""
    from distributed_lion_pytorch_b200.optim import semantics as sem
    n, steps, lr = 1 << 14, 64, 1e-4
    g = torch.ones(n).bfloat16()
    for sr in (False, True):
        p = torch.ones(n).bfloat16()
        m = torch.zeros(n).bfloat16()
        for s in range(steps):
            sem.local_lion_step_(p, g, m, lr, 0.0, 0.9, 0.99, sr_key=sem.sr_key(0, s) if sr else 0)
        drift = 1.0 - float(p.float().mean())
        if sr:
            assert abs(drift - steps * lr) < 0.15 * steps * lr, drift
        else:
            assert drift == 0.0

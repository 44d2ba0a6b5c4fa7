"""Oracle tier (CPU / gloo, W = 2, 3, 4): our Lion against the reference's
update_fn_distributed running in the same processes (reference: distributed_lion.py:61-96)."""
import pytest
import torch

from distutil import run_distributed


def _make(seed, dtype):
    torch.manual_seed(seed)
    m = torch.nn.Sequential(torch.nn.Linear(33, 17), torch.nn.Tanh(), torch.nn.Linear(17, 5))
    return m.to(dtype)


def _worker_vs_reference(rank, world, dtype_name, steps):
    import torch.distributed as dist
    from conftest import load_reference_module
    from distributed_lion import Lion
    dtype = getattr(torch, dtype_name)
    ref = load_reference_module()
    m_ref, m_new = _make(0, dtype), _make(0, dtype)       # identical replicas (X4 broadcast)
    o_ref = ref.Lion(m_ref.parameters(), lr=1e-2, weight_decay=0.1)
    o_new = Lion(m_new.parameters(), lr=1e-2, weight_decay=0.1)
    assert o_ref.update_fn is ref.update_fn_distributed
    s_ref = torch.optim.lr_scheduler.LambdaLR(o_ref, lambda s: 0.9 ** s)
    s_new = torch.optim.lr_scheduler.LambdaLR(o_new, lambda s: 0.9 ** s)
    sent = []
    orig = dist.all_gather_into_tensor

    def spy(out, inp, *a, **k):
        sent.append((inp.dtype, inp.numel() * inp.element_size()))
        return orig(out, inp, *a, **k)
    dist.all_gather_into_tensor = spy
    g = torch.Generator().manual_seed(100 + rank)          # every rank sees different data
    for _ in range(steps):
        x = torch.randn(6, 33, generator=g).to(dtype)
        for m, o, s in ((m_ref, o_ref, s_ref), (m_new, o_new, s_new)):
            o.zero_grad()
            m(x).float().pow(2).mean().backward()
            o.step()
            s.step()
    dist.all_gather_into_tensor = orig
    assert o_new.backend_name == "nccl_1bit"
    exact = all(torch.equal(a, b) for a, b in zip(m_ref.parameters(), m_new.parameters()))
    close = all(torch.allclose(a.float(), b.float(), rtol=2 ** -6, atol=3e-5)
                for a, b in zip(m_ref.parameters(), m_new.parameters()))
    mom_exact = all(torch.equal(o_ref.state[a]["exp_avg"], o_new.state[b]["exp_avg"])
                    for a, b in zip(m_ref.parameters(), m_new.parameters()))
    flat = torch.cat([p.detach().float().flatten() for p in m_new.parameters()])
    mom = torch.cat([o_new.state[p]["exp_avg"].float().flatten() for p in m_new.parameters()])
    return dict(exact=exact, close=close, mom_exact=mom_exact, params=flat, mom=mom, sent=sent,
                payload=sum(p.numel() for p in m_new.parameters()))


@pytest.mark.parametrize("world", [2, 3, 4])
def test_fp32_bitexact_vs_reference(world):
    outs = run_distributed(_worker_vs_reference, world, "float32", 5)
    for o in outs:
        assert o["exact"] and o["mom_exact"]
    for o in outs[1:]:
        assert torch.equal(outs[0]["params"], o["params"])         # replicas bit-identical
    assert not torch.equal(outs[0]["mom"], outs[1]["mom"])          # momenta are per worker
    # wire accounting: one uint8 tensor per bucket per step, 1 bit per (padded) parameter
    dt, nbytes = outs[0]["sent"][0]
    assert dt == torch.uint8 and nbytes == 8192 // 8 and len(outs[0]["sent"]) == 5


def test_bf16_matches_reference_to_one_ulp():
    outs = run_distributed(_worker_vs_reference, 2, "bfloat16", 4)
    for o in outs:
        assert o["close"]
    assert torch.equal(outs[0]["params"], outs[1]["params"])


def _worker_modes(rank, world, vote, tie, live_mask, max_grad_norm):
    from distributed_lion import Lion
    from distributed_lion_pytorch_b200.optim import semantics as sem
    dtype = torch.float32
    model = _make(1, dtype)
    opt = Lion(model.parameters(), lr=1e-2, weight_decay=0.05, max_grad_norm=max_grad_norm, vote=vote, tie=tie,
               live_mask=live_mask, seed=5)
    # shadow: all W workers simulated in this process with the oracle
    shadows = [_make(1, dtype) for _ in range(world)]
    ms = [[torch.zeros_like(p) for p in s.parameters()] for s in shadows]
    gens = [torch.Generator().manual_seed(200 + r) for r in range(world)]
    for step in range(4):
        xs = [torch.randn(6, 33, generator=gens[r]) for r in range(world)]
        opt.zero_grad()
        model(xs[rank]).pow(2).mean().backward()
        opt.step()
        if max_grad_norm is not None:
            continue
        grads = []
        for r in range(world):
            shadows[r].zero_grad()
            shadows[r](xs[r]).pow(2).mean().backward()
            grads.append([p.grad for p in shadows[r].parameters()])
        with torch.no_grad():
            for i in range(len(ms[0])):
                sem.simulate_workers_step_([list(s.parameters())[i].data for s in shadows],
                                           [grads[r][i] for r in range(world)], [ms[r][i] for r in range(world)],
                                           1e-2, 0.05, 0.9, 0.99, vote, tie, live_mask)
    mine = torch.cat([p.detach().flatten() for p in model.parameters()])
    want = torch.cat([p.detach().flatten() for p in shadows[rank].parameters()])
    return dict(mine=mine, want=want)


@pytest.mark.parametrize("vote,tie,mask", [("majority", "zero", None), ("average", "neg", None),
                                           ("majority", "neg", 0b0110), ("average", "neg", 0b1011)])
def test_modes_match_oracle(vote, tie, mask):
    outs = run_distributed(_worker_modes, 4, vote, tie, mask, None)
    for o in outs:
        assert torch.allclose(o["mine"], o["want"], rtol=0, atol=1e-7), (o["mine"] - o["want"]).abs().max()
    for o in outs[1:]:
        assert torch.equal(outs[0]["mine"], o["mine"])


def test_stochastic_mode_runs_and_keeps_replicas_identical():
    outs = run_distributed(_worker_modes, 2, "majority", "neg", None, 1.0)
    assert torch.equal(outs[0]["mine"], outs[1]["mine"])


def _worker_grad_none(rank, world):
    from distributed_lion import Lion
    model = _make(2, torch.float32)
    opt = Lion(model.parameters(), lr=1e-2)
    g = torch.Generator().manual_seed(rank)
    for step in range(3):
        for p in model.parameters():
            p.grad = None
        x = torch.randn(4, 33, generator=g)
        h = model[0](x)
        # rank 1 never touches the second layer: upstream this desynchronises all_gather (D6)
        loss = h.pow(2).mean() if rank == 1 else model[2](torch.tanh(h)).pow(2).mean()
        loss.backward()
        opt.step()
    return torch.cat([p.detach().flatten() for p in model.parameters()])


def test_grad_none_on_one_rank_does_not_hang():
    outs = run_distributed(_worker_grad_none, 2)
    assert torch.equal(outs[0], outs[1])


def _worker_functional(rank, world):
    import distributed_lion as ours
    from conftest import load_reference_module
    ref = load_reference_module()
    torch.manual_seed(0)
    p1 = torch.randn(37, 3)
    p2 = p1.clone()
    g = torch.Generator().manual_seed(50 + rank)
    m1 = torch.zeros_like(p1)
    m2 = torch.zeros_like(p1)
    for _ in range(3):
        gr = torch.randn(37, 3, generator=g)
        ref.update_fn_distributed(p1, gr, m1, 1e-2, 0.1, 0.9, 0.99)
        ours.update_fn_distributed(p2, gr, m2, 1e-2, 0.1, 0.9, 0.99)
    same = torch.equal(p1, p2) and torch.equal(m1, m2)
    ours.update_fn_distributed_stoc(p2, gr, m2, 1e-2, 0.1, 0.9, 0.99, 1.0)     # runs (dead code upstream)
    return dict(same=same, p_ref=p1, p_stoc=p2)


def test_functional_update_fn_distributed_matches_reference():
    outs = run_distributed(_worker_functional, 3)
    assert all(o["same"] for o in outs)                                # bit-exact vs the reference, per tensor
    assert torch.equal(outs[0]["p_ref"], outs[1]["p_ref"])
    assert torch.equal(outs[0]["p_stoc"], outs[2]["p_stoc"])           # stochastic ballots still give one common update


def _worker_vote_stats(rank, world):
    from distributed_lion import Lion
    p = torch.nn.Parameter(torch.zeros(1000))
    opt = Lion([p], lr=1e-2)
    # element i: ranks < (i % 5) vote +1, the others -1  ->  known counts
    g = torch.where(torch.arange(1000) % 5 > rank, 1.0, -1.0)
    p.grad = g.clone()
    before = p.detach().clone()
    st = opt.vote_statistics()
    assert torch.equal(p.detach(), before)                      # a diagnostic: nothing moves
    return st


def test_vote_statistics_known_pattern():
    outs = run_distributed(_worker_vote_stats, 4)
    st = outs[0]
    assert all(o == st for o in outs)
    # counts cycle through 0,1,2,3,4 (of 4): ties at cnt=2 (1/5), unanimous at 0 and 4 (2/5)
    assert abs(st["tie_fraction"] - 0.2) < 0.01 and abs(st["unanimous_fraction"] - 0.4) < 0.01
    assert abs(st["mean_agreement"] - (1 + 0.5 + 0 + 0.5 + 1) / 5) < 0.01


def _worker_resync(rank, world):
    from distributed_lion import Lion
    torch.manual_seed(rank)                                   # replicas deliberately different
    p = torch.nn.Parameter(torch.randn(300))
    opt = Lion([p], lr=1e-2)
    p.grad = torch.randn(300)
    opt.step()
    opt.resync_replicas(src=1)
    return p.detach().clone()


def test_resync_replicas_rebroadcasts_parameters():
    outs = run_distributed(_worker_resync, 3)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[2], outs[1])


def _worker_replica_check(rank, world):
    from distributed_lion import Lion
    torch.manual_seed(0)                                      # identical replicas, different gradients
    ps = [torch.nn.Parameter(torch.randn(300)), torch.nn.Parameter(torch.randn(40, 7).bfloat16())]
    opt = Lion(ps, lr=1e-2)
    g = torch.Generator().manual_seed(100 + rank)
    for _ in range(3):
        for p in ps:
            p.grad = torch.randn(p.shape, generator=g).to(p.dtype)
        opt.step()
    same_before = opt.replicas_identical()
    cs = opt.replica_checksum().tolist()
    if rank == 1:                                             # one worker silently corrupts one element
        with torch.no_grad():
            ps[1].view(-1)[5] += 0.5
    same_after = opt.replicas_identical()
    opt.resync_replicas(0)
    return same_before, same_after, opt.replicas_identical(), cs


def test_replica_checksums_detect_a_diverged_worker():
    outs = run_distributed(_worker_replica_check, 3)
    assert all(o[0] is True and o[1] is False and o[2] is True for o in outs)
    assert outs[0][3] == outs[1][3] == outs[2][3] and len(outs[0][3]) == 2      # one checksum per bucket


def _worker_sr(rank, world, comm):
    from distributed_lion import Lion
    torch.manual_seed(0)
    ps = [torch.nn.Parameter((torch.randn(1000) * 0.5).bfloat16()), torch.nn.Parameter(torch.randn(64, 9).bfloat16())]
    ref = [p.detach().clone() for p in ps]
    opt = Lion(ps, lr=1e-4, weight_decay=0.1, stochastic_rounding=True, seed=5, comm=comm)
    g = torch.Generator().manual_seed(100 + rank)
    for _ in range(4):
        for p in ps:
            p.grad = torch.randn(p.shape, generator=g).to(p.dtype)
        opt.step()
    moved = sum(int((p.detach() != r).sum()) for p, r in zip(ps, ref))
    return [p.detach().clone() for p in ps], opt.replicas_identical(), moved


@pytest.mark.parametrize("comm", ["nccl_1bit"])
def test_stochastic_rounding_keeps_replicas_bit_identical(comm):
    outs = run_distributed(_worker_sr, 3, comm)
    for o in outs[1:]:
        assert all(torch.equal(a, b) for a, b in zip(o[0], outs[0][0]))      # same noise on every rank
    assert all(o[1] for o in outs) and outs[0][2] > 0


def _worker_crosscheck(rank, world):
    """step_crosscheck / clip_grad_norm_ / zero_grads_in_step plumbing on the gloo path (the kernels' folded forms
    are covered by the GPU tests): an optimizer step taken through the active back-end equals the split path on
    clones, replicas stay identical, and the eager clip keeps twin optimizers in lock-step."""
    from distributed_lion import Lion
    m = _make(0, torch.float32)
    o = Lion(m.parameters(), lr=1e-2, weight_decay=0.1, zero_grads_in_step=True)
    g = torch.Generator().manual_seed(100 + rank)
    res = []
    for _ in range(3):
        x = torch.randn(6, 33, generator=g)
        (m(x).pow(2).mean() * 30).backward()
        norm = o.clip_grad_norm_(1.0)
        chk = o.step_crosscheck()
        o.zero_grad()
        res.append((float(norm) > 0, chk["equal"], chk["replicas_identical"]))
    return dict(res=res, steps=o.steps_done, params=torch.cat([p.detach().flatten() for p in m.parameters()]))


def test_step_crosscheck_and_clip_on_gloo():
    outs = run_distributed(_worker_crosscheck, 2)
    for o in outs:
        assert o["steps"] == 3 and all(all(t) for t in o["res"])
    assert torch.equal(outs[0]["params"], outs[1]["params"])

"""GPU-multi tier (needs >= 2 GPUs; one process per GPU over NCCL): the fused NVLink kernel
against the nccl_1bit split path and against the CPU oracle; replicas bit-identical; live
masks; skewed arrival."""
import pytest
import torch

from distutil import run_distributed

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _net(seed, dtype):
    torch.manual_seed(seed)
    m = torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.Tanh(), torch.nn.Linear(512, 512), torch.nn.Tanh(),
                            torch.nn.Linear(512, 64))
    return m.to(dtype)


def _worker(rank, world, dtype_name, vote, tie, mask, steps, skew, multicast=False, sr=False):
    import os
    import time
    os.environ["DLION_MULTICAST"] = "1" if multicast else "0"
    from distributed_lion import Lion
    dtype = getattr(torch, dtype_name)
    dev = torch.device("cuda", rank)
    nets, opts = {}, {}
    for comm in ("fused", "nccl_1bit"):
        nets[comm] = _net(0, dtype).to(dev)
        opts[comm] = Lion(nets[comm].parameters(), lr=1e-3, weight_decay=0.1, vote=vote, tie=tie, comm=comm,
                          live_mask=mask, stochastic_rounding=sr, seed=9)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    for s in range(steps):
        x = torch.randn(64, 256, device=dev, generator=g).to(dtype)
        if skew and rank == (s % world):
            torch.cuda.synchronize()
            time.sleep(0.05)                       # this rank arrives late: peers must wait on its flags
        for comm in ("fused", "nccl_1bit"):
            opts[comm].zero_grad()
            nets[comm](x).float().pow(2).mean().backward()
            opts[comm].step()
    torch.cuda.synchronize()
    opts["fused"]._backend._ctx.check_status()
    flat = {c: torch.cat([p.detach().float().flatten() for p in nets[c].parameters()]).cpu() for c in nets}
    mom = torch.cat([opts["fused"].state[p]["exp_avg"].float().flatten() for p in nets["fused"].parameters()]).cpu()
    same = opts["fused"].replicas_identical()
    return dict(fused=flat["fused"], split=flat["nccl_1bit"], mom=mom, backend=opts["fused"].backend_name, same=same,
                how=opts["fused"]._backend._ctx.flags.how, mc=opts["fused"]._backend._ctx.uses_multicast)


@pytest.mark.parametrize("dtype", ["bfloat16", "float32"])
@pytest.mark.parametrize("vote,tie,mask", [("majority", "neg", None), ("average", "neg", None),
                                           ("majority", "zero", 0b01)])
def test_fused_equals_split_path_and_replicas_identical(dtype, vote, tie, mask):
    world = min(torch.cuda.device_count(), 8)
    steps = 60 if (dtype == "bfloat16" and vote == "majority" and mask is None) else 6     # one long run: many flag epochs
    outs = run_distributed(_worker, world, dtype, vote, tie, mask, steps, True, backend="nccl", timeout=300)
    assert outs[0]["backend"] == "fused"
    for o in outs:
        assert torch.equal(o["fused"], o["split"])                 # same bits, same arithmetic
        assert torch.equal(o["fused"], outs[0]["fused"])           # replicas bit-identical
    assert not torch.equal(outs[0]["mom"], outs[1]["mom"])         # momenta stay per worker
    print("peer mapping:", outs[0]["how"])


def _worker_auto(rank, world):
    from distributed_lion import Lion
    dev = torch.device("cuda", rank)
    net = _net(0, torch.bfloat16).to(dev)
    opt = Lion(net.parameters(), lr=1e-3)
    net(torch.randn(8, 256, device=dev).bfloat16()).float().pow(2).mean().backward()
    opt.step()
    torch.cuda.synchronize()
    return opt.backend_name


def test_auto_picks_fused_on_one_node():
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_worker_auto, world, backend="nccl", timeout=300)
    assert all(o == "fused" for o in outs)


def test_fused_nvls_multicast_send_path():
    """DLION_MULTICAST=1: the packed words leave the GPU once (multimem.st) and NVSwitch replicates
    them into every arena.  Must give exactly the bits of the split path; skipped where the fabric
    exposes no multicast pointer."""
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_worker, world, "bfloat16", "majority", "neg", None, 8, True, True, backend="nccl", timeout=300)
    if not outs[0]["mc"]:
        pytest.skip("no NVLS multicast pointer on this fabric")
    for o in outs:
        assert torch.equal(o["fused"], o["split"]) and torch.equal(o["fused"], outs[0]["fused"])



def test_stochastic_rounding_fused_equals_split_and_replicas_identical():
    """Opt-in stochastic rounding on real GPUs: the rounding noise is keyed by (seed, step, element), so the fused
    kernel, the split path and every rank draw the same bits."""
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_worker, world, "bfloat16", "majority", "neg", None, 10, True, False, True, backend="nccl",
                           timeout=300)
    base = run_distributed(_worker, world, "bfloat16", "majority", "neg", None, 10, False, False, False, backend="nccl",
                           timeout=300)
    for o in outs:
        assert o["same"] and torch.equal(o["fused"], o["split"]) and torch.equal(o["fused"], outs[0]["fused"])
    assert not torch.equal(outs[0]["fused"], base[0]["fused"])          # and it is not round-to-nearest


# --------------------------------------------------------------------------- real size, real NVLink
def _big_params(dev, seed, small=False):
    """Two param groups; each holds bf16 AND fp32 tensors, so the optimizer owns four buckets that share one flag
    table and one arena (which has to re-grow when the larger bucket comes second).  The bf16 bucket of group 0
    has > 64 M elements with odd shapes (padding inside the bucket)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    mk = lambda shape, dt: torch.nn.Parameter((torch.randn(*shape, device=dev, generator=g) * 0.05).to(dt))  # noqa: E731
    if small:      # same structure, ~8 M elements: for grids of a few CTAs
        g0 = [mk((501, 1003), torch.float32), mk((1024, 4096), torch.bfloat16), mk((501, 7001), torch.bfloat16),
              mk((77,), torch.bfloat16)]
        g1 = [mk((256, 4096), torch.bfloat16), mk((555,), torch.float32)]
        return g0, g1
    g0 = [mk((1001, 1003), torch.float32), mk((4096, 8192), torch.bfloat16), mk((3001, 7001), torch.bfloat16),
          mk((2048, 2048), torch.float32), mk((1024, 10001), torch.bfloat16), mk((77,), torch.bfloat16)]
    g1 = [mk((1024, 4096), torch.bfloat16), mk((555,), torch.float32), mk((2000, 999), torch.bfloat16)]
    return g0, g1


def _worker_big(rank, world, chunk_elems, steps, multicast, grid, skew):
    import os
    import random
    import time
    os.environ["DLION_MULTICAST"] = "1" if multicast else "0"
    os.environ["DLION_CHUNK_ELEMS"] = str(chunk_elems)
    if grid:
        os.environ["DLION_FUSED_GRID"] = str(grid)
    from distributed_lion import Lion
    dev = torch.device("cuda", rank)
    opts, params = {}, {}
    for comm in ("fused", "nccl_1bit"):
        g0, g1 = _big_params(dev, seed=0, small=bool(grid))
        params[comm] = g0 + g1
        opts[comm] = Lion([dict(params=g0, lr=1e-3, weight_decay=0.1), dict(params=g1, lr=3e-4, weight_decay=0.0)],
                          comm=comm, zero_grads_in_step=True)
    gen = torch.Generator(device=dev).manual_seed(1000 + rank)
    rnd = random.Random(5 + rank)
    bad_step = -1
    for s in range(steps):
        for i, p in enumerate(params["fused"]):
            # a norm well above the clip threshold in some steps, below it in others
            gr = (torch.randn(p.shape, device=dev, generator=gen) * (1e-3 if s % 3 else 1e-5)).to(p.dtype)
            for comm in ("fused", "nccl_1bit"):
                q = params[comm][i]
                if q.grad is None:
                    q.grad = gr.clone()
                else:
                    q.grad.copy_(gr)
        if skew:
            torch.cuda.synchronize()
            time.sleep(rnd.random() * 0.004)             # ranks enter the kernel at different times
        for comm in ("fused", "nccl_1bit"):
            opts[comm].clip_grad_norm_(1.0)
            opts[comm].step()
            opts[comm].zero_grad()
        if s % 10 == 9 or s == steps - 1:
            same = all(torch.equal(bf.flat_p, bs.flat_p) and torch.equal(bf.flat_m, bs.flat_m)
                       for bf, bs in zip(opts["fused"].buckets, opts["nccl_1bit"].buckets))
            if not same and bad_step < 0:
                bad_step = s
    torch.cuda.synchronize()
    ctx = opts["fused"]._backend._ctx
    ctx.check_status()
    nb = [(str(b.dtype), b.numel) for b in opts["fused"].buckets]
    return dict(bad_step=bad_step, identical=opts["fused"].replicas_identical(), backend=opts["fused"].backend_name,
                buckets=nb, launches=ctx.epoch, mc=ctx.uses_multicast, grid=ctx.grid_for(torch.bfloat16),
                clean=all(int(b.flat_g.count_nonzero()) == 0 for b in opts["fused"].buckets))


@pytest.mark.parametrize("multicast", [False, True])
def test_fused_real_size_multi_chunk_multi_bucket(multicast):
    """The fused kernel at real size on real NVLink: > 64 M-element bf16 bucket + fp32 buckets + a second param
    group (four buckets, one flag table, the arena re-grown once), DLION_CHUNK_ELEMS = 4 Mi (about 20 launches and
    epochs per step, every ring wraps many times), gradient clip and zero-fill folded into the kernel, 50 steps —
    compared bit for bit with the split pack / all_gather / vote path every 10 steps on every rank."""
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_worker_big, world, 1 << 22, 50, multicast, 0, False, backend="nccl", timeout=900)
    if multicast and not outs[0]["mc"]:
        pytest.skip("no NVLS multicast pointer on this fabric")
    assert outs[0]["backend"] == "fused"
    assert max(n for _, n in outs[0]["buckets"]) >= 64 * 1000 * 1000 and len(outs[0]["buckets"]) == 4
    assert outs[0]["launches"] >= 50 * 18
    for o in outs:
        assert o["bad_step"] < 0, f"fused != split from step {o['bad_step']}"
        assert o["identical"] and o["clean"]
    print(f"W={world} buckets={outs[0]['buckets']} launches={outs[0]['launches']} multicast={outs[0]['mc']}")


@pytest.mark.parametrize("grid", [3, 37])
def test_fused_small_grid_skewed_arrival_real_gpus(grid):
    """Few CTAs own many tiles each (every smem ring and the 32-tile sender ring wrap and fill), ranks arrive up to
    4 ms apart, 1 Mi-element chunks: the protocol's back-pressure paths on real peer memory."""
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_worker_big, world, 1 << 20, 6, grid == 37, grid, True, backend="nccl", timeout=900)
    for o in outs:
        assert o["grid"] == grid and o["bad_step"] < 0 and o["identical"] and o["clean"]


# --------------------------------------------------------------------------- a peer that stops stepping
def _worker_stall(rank, world, tmp):
    """Rank 1 stops taking optimizer steps after step 3 (its process stays alive, as a hung or very late worker
    would).  Rank 0 must not hang: its kernel gives up after the wait time-out, leaves the affected parameter tiles
    untouched, the NEXT step() raises without a device sync, a checkpoint attempt raises instead of writing, and
    after declare_dead(1) rank 0 keeps training on its own ballots."""
    import os
    import time
    os.environ["DLION_WAIT_TIMEOUT_S"] = "1.0"
    from distributed_lion import Lion
    dev = torch.device("cuda", rank)
    net = _net(0, torch.bfloat16).to(dev)
    opt = Lion(net.parameters(), lr=1e-3, weight_decay=0.0, comm="fused")
    g = torch.Generator(device=dev).manual_seed(100 + rank)

    def one_step():
        x = torch.randn(64, 256, device=dev, generator=g).bfloat16()
        opt.zero_grad()
        net(x).float().pow(2).mean().backward()
        opt.step()

    for _ in range(3):
        one_step()
    torch.cuda.synchronize()
    out = {"rank": rank}
    if rank == 1:
        time.sleep(12)                     # "dead": never steps again, but its memory stays mapped
        return out
    before = [p.detach().clone() for p in net.parameters()]
    t0 = time.time()
    one_step()                             # peers' flags never come: bounded wait inside the kernel
    raised_at = None
    for i in range(50):
        try:
            one_step()
        except RuntimeError as e:
            raised_at, out["msg"] = i, str(e)
            break
        time.sleep(0.05)
    out["seconds_to_raise"] = time.time() - t0
    out["raised_at"] = raised_at
    torch.cuda.synchronize()
    # the stalled step applied nothing to the parameters (every tile's ballots were incomplete)
    out["untouched"] = all(torch.equal(a, p.detach()) for a, p in zip(before, net.parameters()))
    try:
        opt.check_health()
        out["health_raises"] = False
    except RuntimeError:
        out["health_raises"] = True
    opt.declare_dead(1)
    for _ in range(3):
        one_step()
    torch.cuda.synchronize()
    opt.check_health()
    out["moved_after_declare_dead"] = not all(torch.equal(a, p.detach()) for a, p in zip(before, net.parameters()))
    return out


def test_stalled_peer_raises_quickly_and_can_be_declared_dead(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    outs = run_distributed(_worker_stall, 2, str(tmp_path), backend="nccl", timeout=300)
    o = outs[0]
    assert o["raised_at"] is not None and "stalled exchange" in o["msg"], o
    assert o["seconds_to_raise"] < 5.0, o
    assert o["untouched"] and o["health_raises"] and o["moved_after_declare_dead"], o

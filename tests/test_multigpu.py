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

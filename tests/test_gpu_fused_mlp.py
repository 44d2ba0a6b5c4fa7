"""tcgen05 dgrad GEMM with the GELU backward and the bias gradient in its epilogue (csrc/gemm_dgelu_sm100.cu,
ops/mlp.py) against a plain fp32 PyTorch reference of the same op, and the fused MLP node against autograd.
reference: HF GPT2MLP (Conv1D -> NewGELUActivation -> Conv1D) backward under autograd, async_trainer.py:33."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ref_dgrad_dgelu(dy, w, pre):
    x = pre.float().requires_grad_(True)
    F.gelu(x, approximate="tanh").backward(dy.float() @ w.float().t())
    return x.grad


@pytest.mark.parametrize("m,n,k", [(128, 256, 64), (300, 520, 72), (2048, 3072, 768), (20480, 3072, 768), (1000, 16384, 4096)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_dgrad_dgelu_matches_fp32_reference(m, n, k, dtype):
    from distributed_lion_pytorch_b200.ops.mlp import dgrad_dgelu
    if dtype == torch.float16 and m > 2048:
        pytest.skip("large shapes in bf16 only")
    g = torch.Generator(device="cuda").manual_seed(m + n)
    dy = (torch.randn(m, k, device="cuda", generator=g) * 0.5).to(dtype)
    w = (torch.randn(n, k, device="cuda", generator=g) * k ** -0.5).to(dtype)
    pre = (torch.randn(m, n, device="cuda", generator=g) * 1.5).to(dtype)
    dpre, db = dgrad_dgelu(dy, w, pre)
    torch.cuda.synchronize()
    ref = _ref_dgrad_dgelu(dy, w, pre)
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    torch.testing.assert_close(dpre.float(), ref, atol=tol, rtol=tol)
    # the bias gradient is the column sum of what was stored
    torch.testing.assert_close(db.float(), dpre.float().sum(0), atol=2e-2 * max(1.0, m ** 0.5 / 8), rtol=1e-2)
    # accumulate into an existing gradient
    acc = torch.ones(n, device="cuda", dtype=dtype)
    dgrad_dgelu(dy, w, pre, bias_grad_out=acc, accumulate=True)
    torch.testing.assert_close(acc.float(), db.float() + 1.0, atol=2e-2 * max(1.0, m ** 0.5 / 8), rtol=1e-2)
    # deterministic
    dpre2, db2 = dgrad_dgelu(dy, w, pre)
    assert torch.equal(dpre, dpre2) and torch.equal(db, db2)


def test_fused_mlp_node_matches_autograd():
    from distributed_lion_pytorch_b200.ops.mlp import gelu_mlp
    torch.manual_seed(0)
    B, S, E, I = 3, 130, 256, 1024
    x = torch.randn(B, S, E, device="cuda").to(torch.bfloat16).requires_grad_(True)
    w_fc = (torch.randn(E, I, device="cuda") * E ** -0.5).to(torch.bfloat16).requires_grad_(True)
    b_fc = (torch.randn(I, device="cuda") * 0.1).to(torch.bfloat16).requires_grad_(True)
    w_pr = (torch.randn(I, E, device="cuda") * I ** -0.5).to(torch.bfloat16).requires_grad_(True)
    b_pr = (torch.randn(E, device="cuda") * 0.1).to(torch.bfloat16).requires_grad_(True)
    dy = torch.randn(B, S, E, device="cuda").to(torch.bfloat16)
    y = gelu_mlp(x, w_fc, b_fc, w_pr, b_pr)
    assert y is not None
    y.backward(dy)
    got = [t.grad.float().clone() for t in (x, w_fc, b_fc, w_pr, b_pr)]
    ps = [t.detach().float().requires_grad_(True) for t in (x, w_fc, b_fc, w_pr, b_pr)]
    yr = F.gelu(ps[0] @ ps[1] + ps[2], approximate="tanh") @ ps[3] + ps[4]
    yr.backward(dy.float())
    torch.testing.assert_close(y.float(), yr, atol=3e-2, rtol=3e-2)
    for a, p_, name in zip(got, ps, ("dx", "dw_fc", "db_fc", "dw_proj", "db_proj")):
        scale = p_.grad.abs().max().item()
        assert (a - p_.grad).abs().max().item() <= 2e-2 * scale + 1e-3, name
    # second backward pass accumulates into the live gradients (the steady state of training)
    y2 = gelu_mlp(x, w_fc, b_fc, w_pr, b_pr)
    y2.backward(dy)
    for t, a in zip((x, w_fc, b_fc, w_pr, b_pr), got):
        torch.testing.assert_close(t.grad.float(), 2 * a, atol=3e-2 * a.abs().max().item(), rtol=2e-2)


def test_gpt2_block_uses_the_fused_mlp_and_matches_the_unfused_path(monkeypatch):
    from distributed_lion_pytorch_b200.models.gpt2 import GPT2Config, GPT2LMHeadModel
    from distributed_lion_pytorch_b200.ops import mlp
    cfg = GPT2Config(n_embd=128, n_layer=2, n_head=4, vocab_size=512, n_positions=64, resid_pdrop=0.0, embd_pdrop=0.0,
                     attn_pdrop=0.0)
    torch.manual_seed(1)
    model = GPT2LMHeadModel(cfg).cuda().to(torch.bfloat16)
    ids = torch.randint(0, 512, (4, 64), device="cuda")
    calls = []
    real = mlp.dgrad_dgelu
    monkeypatch.setattr(mlp, "dgrad_dgelu", lambda *a, **k: calls.append(1) or real(*a, **k))
    model(ids, labels=ids).loss.backward()
    assert len(calls) == cfg.n_layer
    fused = {n: p.grad.float().clone() for n, p in model.named_parameters()}
    model.zero_grad(set_to_none=True)
    monkeypatch.setattr(mlp, "_ENABLED", False)
    model(ids, labels=ids).loss.backward()
    assert len(calls) == cfg.n_layer
    for n, p in model.named_parameters():
        ref = p.grad.float()
        assert (fused[n] - ref).abs().max().item() <= 3e-2 * ref.abs().max().item() + 1e-4, n


def test_cta_pair_variant_matches_reference_in_a_subprocess():
    """The cta_group::2 build of the same kernel (two CTAs of a cluster share a [256 x 256] tile; opt-in with
    DLION_DGELU_VARIANT=pair, read once per process) gives the same numbers as the fp32 reference."""
    import os
    import subprocess
    import sys
    code = r'''
import torch, torch.nn.functional as F
from distributed_lion_pytorch_b200.ops.mlp import dgrad_dgelu
for m, n, k in ((300, 520, 72), (4096, 3072, 768)):
    g = torch.Generator(device="cuda").manual_seed(m)
    dy = (torch.randn(m, k, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
    w = (torch.randn(n, k, device="cuda", generator=g) * k ** -0.5).to(torch.bfloat16)
    pre = (torch.randn(m, n, device="cuda", generator=g) * 1.5).to(torch.bfloat16)
    dpre, db = dgrad_dgelu(dy, w, pre)
    x = pre.float().requires_grad_(True)
    F.gelu(x, approximate="tanh").backward(dy.float() @ w.float().t())
    torch.testing.assert_close(dpre.float(), x.grad, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(db.float(), dpre.float().sum(0), atol=0.2, rtol=1e-2)
print("pair ok")
'''
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, DLION_DGELU_VARIANT="pair", PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0 and "pair ok" in r.stdout, r.stdout + r.stderr

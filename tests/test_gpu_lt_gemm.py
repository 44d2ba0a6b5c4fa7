"""GPU tier: cuBLASLt epilogue GEMMs (csrc/lt_gemm.cu) against fp32 PyTorch math, and the fused
linear+GELU autograd function the GPT-2 MLP uses."""
import pytest
import torch
import torch.nn.functional as F

from distributed_lion_pytorch_b200.ops import lt_gemm

pytestmark = pytest.mark.gpu


def _need_lt():
    if not lt_gemm.available():
        pytest.skip("cuBLASLt not loadable in this process")


def _close(a, ref, tol=2e-2):
    scale = ref.abs().max().clamp_min(1e-6)
    assert float((a.float() - ref).abs().max() / scale) < tol


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("trans_a,trans_b", [(False, False), (True, False), (False, True), (True, True)])
def test_plain_and_bias_gemm_all_layouts(dtype, trans_a, trans_b):
    _need_lt()
    torch.manual_seed(0)
    m, n, k = 384, 256, 512
    a = torch.randn((k, m) if trans_a else (m, k), device="cuda").to(dtype)
    b = torch.randn((n, k) if trans_b else (k, n), device="cuda").to(dtype)
    bias = torch.randn(n, device="cuda").to(dtype)
    ref = (a.float().t() if trans_a else a.float()) @ (b.float().t() if trans_b else b.float())
    d = torch.empty(m, n, device="cuda", dtype=dtype)
    assert lt_gemm.gemm(a, b, d, trans_a=trans_a, trans_b=trans_b)
    _close(d, ref)
    assert lt_gemm.gemm(a, b, d, trans_a=trans_a, trans_b=trans_b, epilogue=lt_gemm.BIAS, bias=bias)
    _close(d, ref + bias.float())
    d0 = torch.randn(m, n, device="cuda").to(dtype)
    d = d0.clone()
    assert lt_gemm.gemm(a, b, d, trans_a=trans_a, trans_b=trans_b, alpha=0.5, beta=1.0)     # accumulate in place
    _close(d, 0.5 * ref + d0.float())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_training_epilogues(dtype):
    _need_lt()
    torch.manual_seed(1)
    M, E, Fd = 1024, 256, 768
    x = torch.randn(M, E, device="cuda").to(dtype)
    w = (torch.randn(E, Fd, device="cuda") * 0.05).to(dtype)
    bias = (torch.randn(Fd, device="cuda") * 0.1).to(dtype)
    pre_ref = x.float() @ w.float() + bias.float()
    act = torch.empty(M, Fd, device="cuda", dtype=dtype)
    pre = torch.empty_like(act)
    assert lt_gemm.gemm(x, w, act, epilogue=lt_gemm.GELU_AUX_BIAS, bias=bias, aux=pre)
    _close(pre, pre_ref)
    _close(act, F.gelu(pre_ref, approximate="tanh"))
    # dgrad with the GELU derivative (and the bias gradient) applied in the epilogue
    w2 = (torch.randn(Fd, E, device="cuda") * 0.05).to(dtype)
    dy = torch.randn(M, E, device="cuda").to(dtype)
    p32 = pre.float().requires_grad_()
    (F.gelu(p32, approximate="tanh") * (dy.float() @ w2.float().t())).sum().backward()
    dpre = torch.empty_like(pre)
    if lt_gemm.gemm(dy, w2, dpre, trans_b=True, epilogue=lt_gemm.DGELU, aux=pre):
        _close(dpre, p32.grad)
    bg = torch.empty(Fd, device="cuda", dtype=dtype)
    if lt_gemm.gemm(dy, w2, dpre, trans_b=True, epilogue=lt_gemm.DGELU_BGRAD, aux=pre, bias_grad=bg):
        _close(dpre, p32.grad)
        _close(bg, p32.grad.sum(0))
    # weight gradient + bias gradient in one GEMM, accumulating into an existing buffer
    g0 = torch.randn(Fd, E, device="cuda").to(dtype)
    g = g0.clone()
    bg2 = torch.empty(E, device="cuda", dtype=dtype)
    if lt_gemm.gemm(act, dy, g, trans_a=True, beta=1.0, epilogue=lt_gemm.BGRAD_B, bias_grad=bg2):
        _close(g, g0.float() + act.float().t() @ dy.float())
        _close(bg2, dy.float().sum(0))


def test_argument_errors_and_refusals_do_not_raise_spuriously():
    _need_lt()
    a = torch.randn(64, 64, device="cuda", dtype=torch.bfloat16)
    d = torch.empty(64, 64, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(ValueError):
        lt_gemm.gemm(a, torch.randn(32, 64, device="cuda", dtype=torch.bfloat16), d)
    assert lt_gemm.gemm(a.t(), a, d) is False                  # not row-major: the caller falls back
    assert lt_gemm.gemm(a, a.float(), d) is False               # mixed dtypes
    with pytest.raises(RuntimeError):                            # missing bias is an argument error, not a refusal
        lt_gemm.gemm(a, a, d, epilogue=lt_gemm.BIAS)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("grads_exist", [True, False])
def test_linear_gelu_forward_backward(dtype, grads_exist):
    _need_lt()
    from distributed_lion_pytorch_b200.ops import _cabi
    torch.manual_seed(2)
    B, S, E, Fd = 3, 128, 256, 1024
    x = torch.randn(B, S, E, device="cuda").to(dtype).requires_grad_()
    w = (torch.randn(E, Fd, device="cuda") * 0.05).to(dtype).requires_grad_()
    b = (torch.randn(Fd, device="cuda") * 0.1).to(dtype).requires_grad_()
    w0, b0 = torch.randn_like(w), torch.randn_like(b)
    if grads_exist:
        w.grad, b.grad = w0.detach().clone(), b0.detach().clone()
    dy = torch.randn(B, S, Fd, device="cuda").to(dtype)
    n0 = _cabi.LAUNCHES
    y = lt_gemm.linear_gelu(x, w, b)
    assert y is not None and y.shape == (B, S, Fd) and y._base is None
    y.backward(dy)
    assert _cabi.LAUNCHES > n0                       # gelu_bwd + column-sum kernels of this package ran in backward
    xs, ws, bs = (t.detach().float().requires_grad_() for t in (x, w, b))
    ref = F.gelu(xs @ ws + bs, approximate="tanh")
    ref.backward(dy.float())
    _close(y, ref)
    _close(x.grad, xs.grad)
    _close(w.grad, ws.grad + (w0.float() if grads_exist else 0))
    _close(b.grad, bs.grad + (b0.float() if grads_exist else 0))
    # fp32 / CPU / no-grad requests are declined, the model then runs linear + GELU
    assert lt_gemm.linear_gelu(x.float(), w.float(), b.float()) is None
    with torch.no_grad():
        assert lt_gemm.linear_gelu(x, w, b) is None


def test_gpt2_model_uses_the_fused_mlp_and_agrees_with_the_unfused_one(monkeypatch):
    _need_lt()
    from distributed_lion_pytorch_b200.models import build_model
    torch.manual_seed(0)
    model = build_model("gpt2-tiny", dtype=torch.bfloat16, device="cuda").eval()      # eval: no dropout noise
    ids = torch.randint(0, 512, (4, 64), device="cuda")
    from distributed_lion_pytorch_b200.ops import mlp
    calls = {"n": 0}
    real = lt_gemm._LinearGelu.apply

    def counting(*a):
        calls["n"] += 1
        return real(*a)
    monkeypatch.setattr(lt_gemm._LinearGelu, "apply", staticmethod(counting))
    monkeypatch.setattr(mlp, "_ENABLED", False)        # the one-node MLP (ops/mlp.py, tested in test_gpu_fused_mlp.py) off:
    loss_f = model(ids, labels=ids).loss               # this test is about the GELU-epilogue GEMM underneath it
    loss_f.backward()
    assert calls["n"] == model.config.n_layer
    g_f = {n: p.grad.float().clone() for n, p in model.named_parameters()}
    model.zero_grad()
    monkeypatch.setattr(lt_gemm, "_ENABLED", False)
    loss_u = model(ids, labels=ids).loss
    loss_u.backward()
    assert calls["n"] == model.config.n_layer          # not called again
    assert abs(float(loss_f) - float(loss_u)) < 2e-2
    for n, p in model.named_parameters():
        ref = p.grad.float()
        assert float((g_f[n] - ref).abs().max()) <= 0.05 * float(ref.abs().max()) + 1e-4, n

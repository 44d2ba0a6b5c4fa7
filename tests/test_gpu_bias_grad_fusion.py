"""GPU tier: bias gradients produced by the kernel that already has the tensor in registers — the GELU backward
(c_fc), the norm backward (attn.c_proj / mlp.c_proj) and the qkv packer (c_attn) — against plain PyTorch.
reference: `grad_output.sum(0)` in HF Conv1D's backward, one more pass over each of those tensors upstream."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _tol(dtype):
    return dict(rtol=2e-2, atol=3e-2) if dtype != torch.float32 else dict(rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("rows,cols", [(4096, 3072), (517, 264), (9, 8)])
@pytest.mark.parametrize("accumulate", [False, True])
def test_gelu_backward_with_column_sums(dtype, rows, cols, accumulate):
    from distributed_lion_pytorch_b200.ops import _cabi
    from distributed_lion_pytorch_b200.ops import linear as lin
    L = _cabi.lib()
    g = torch.Generator(device="cuda").manual_seed(rows)
    pre = torch.randn(rows, cols, device="cuda", generator=g).to(dtype)
    dy = torch.randn(rows, cols, device="cuda", generator=g).to(dtype)
    dpre = torch.empty_like(pre)
    ref_dpre = torch.empty_like(pre)
    _cabi.check(L.dlion_gelu_tanh_bwd(pre.data_ptr(), dy.data_ptr(), ref_dpre.data_ptr(), pre.numel(),
                                      _cabi.dtype_code(dtype), _cabi.stream_ptr(pre.device)), "gelu_bwd")
    out = torch.full((cols,), 0.5, device="cuda", dtype=dtype)
    partial = torch.empty(int(L.dlion_colsum_parts(rows, cols)), cols, dtype=torch.float32, device="cuda")
    for _ in range(2):          # the tickets must come back zeroed
        out.fill_(0.5)
        _cabi.check(L.dlion_gelu_tanh_bwd_colsum(pre.data_ptr(), dy.data_ptr(), dpre.data_ptr(), rows, cols,
                                                 partial.data_ptr(), lin._tickets(pre.device).data_ptr(), out.data_ptr(),
                                                 int(accumulate), _cabi.dtype_code(dtype), _cabi.stream_ptr(pre.device)),
                    "gelu_bwd_colsum")
        torch.cuda.synchronize()
        assert torch.equal(dpre, ref_dpre)
        want = ref_dpre.float().sum(0) + (0.5 if accumulate else 0.0)
        assert torch.allclose(out.float(), want, **_tol(dtype)), float((out.float() - want).abs().max())
    # against autograd on the fp32 formula
    x32 = pre.float().requires_grad_()
    F.gelu(x32, approximate="tanh").backward(dy.float())
    assert torch.allclose(dpre.float(), x32.grad, **_tol(dtype))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("layout", ["bhsd", "bshd", "mixed"])
def test_qkv_pack_with_column_sums(dtype, layout):
    from distributed_lion_pytorch_b200.ops.qkv import split_qkv
    from distributed_lion_pytorch_b200.ops import linear as lin
    B, S, H, D = 3, 37, 4, 16
    torch.manual_seed(0)
    qkv = torch.randn(B, S, 3 * H * D, device="cuda", dtype=dtype, requires_grad=True)
    bias = torch.nn.Parameter(torch.zeros(3 * H * D, device="cuda", dtype=dtype))
    q, k, v = split_qkv(qkv, H, bias)
    assert q.shape == (B, H, S, D) and q.data_ptr() == qkv.data_ptr()          # views, no copy
    ref = qkv.detach().view(B, S, 3, H, D)
    assert torch.equal(k, ref[:, :, 1].transpose(1, 2))

    def grad_like(i):
        g = torch.randn(B, H, S, D, device="cuda", dtype=dtype)
        if layout == "bshd" or (layout == "mixed" and i == 1):
            g = torch.randn(B, S, H, D, device="cuda", dtype=dtype).transpose(1, 2)     # strided [B, H, S, D]
        return g
    gs = [grad_like(i) for i in range(3)]
    seen = {}
    qkv.register_hook(lambda g: seen.setdefault("g", g))
    torch.autograd.backward([q, k, v], gs)
    want = torch.stack([g.transpose(1, 2) for g in gs], dim=2).reshape(B, S, -1)
    assert torch.equal(qkv.grad, want)
    tag = getattr(seen["g"], "_dlion_colsum", None)
    assert tag is not None and tag[0] == id(bias)
    assert torch.allclose(tag[1].float(), want.float().sum((0, 1)), **_tol(dtype))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("cols,rows,drop", [(768, 2051, 0.1), (768, 300, 0.0), (64, 4100, 0.0), (2048, 700, 0.1)])
def test_norm_backward_hands_over_the_producing_linears_bias_gradient(dtype, cols, rows, drop):
    """x = linear(h) + bias feeds add_layer_norm; the norm backward's dx column sums must equal the bias gradient
    the linear would have computed, and the linear must use them (no colsum launch of its own)."""
    from distributed_lion_pytorch_b200.ops import linear as lin
    from distributed_lion_pytorch_b200.ops.linear import fused_linear
    from distributed_lion_pytorch_b200.ops.norm import add_layer_norm
    torch.manual_seed(1)
    h = torch.randn(rows, 96, device="cuda", dtype=dtype)
    w = torch.nn.Parameter(torch.randn(96, cols, device="cuda", dtype=dtype) * 0.1)
    b = torch.nn.Parameter(torch.randn(cols, device="cuda", dtype=dtype) * 0.1)
    lw = torch.nn.Parameter(torch.ones(cols, device="cuda", dtype=dtype))
    lb = torch.nn.Parameter(torch.zeros(cols, device="cuda", dtype=dtype))
    res = torch.randn(rows, cols, device="cuda", dtype=dtype)
    up = torch.randn(rows, cols, device="cuda", dtype=dtype)
    results = {}
    for mode in ("handoff", "plain"):
        for p in (w, b, lw, lb):
            p.grad = None
        torch.manual_seed(7)                 # same dropout seed draw
        x = fused_linear(h, w, b, in_out=True)
        y, s = add_layer_norm(x, res, lw, lb, 1e-5, drop, True, x_bias=b if mode == "handoff" else None)
        before = lin.HANDOFFS
        (y * up).sum().backward()
        assert lin.HANDOFFS - before == (1 if mode == "handoff" else 0)
        results[mode] = (b.grad.clone(), w.grad.clone(), lw.grad.clone())
    for a, c in zip(results["handoff"], results["plain"]):
        scale = float(c.float().abs().max()) + 1e-6
        assert float((a.float() - c.float()).abs().max()) / scale < (2e-2 if dtype != torch.float32 else 1e-4)


def test_gpt2_block_uses_all_three_handoffs():
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.ops import linear as lin
    torch.manual_seed(0)
    m = build_model("gpt2-tiny", dtype=torch.bfloat16, device=torch.device("cuda"), pretrained=False)
    m.train()
    ids = torch.randint(0, 512, (2, 64), device="cuda")
    before = lin.HANDOFFS
    m(ids, labels=ids).loss.backward()
    n_layer = m.config.n_layer
    assert lin.HANDOFFS - before == 3 * n_layer, (lin.HANDOFFS - before, n_layer)
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("with_pos", [True, False])
@pytest.mark.parametrize("grads_exist", [True, False])
def test_embedding_forward_and_sorted_backward(dtype, with_pos, grads_exist):
    """One-pass wte[ids] + wpe[pos] and the deterministic backward (sorted ids, one CTA per run of equal ids, adds
    into the existing .grad) against torch.nn.functional.embedding + autograd — with many repeated tokens."""
    from distributed_lion_pytorch_b200.ops.embedding import embed
    torch.manual_seed(0)
    V, H, B, S = 1000, 264, 7, 53
    wte = torch.nn.Parameter(torch.randn(V, H, device="cuda", dtype=dtype))
    wpe = torch.nn.Parameter(torch.randn(64, H, device="cuda", dtype=dtype)) if with_pos else None
    ids = torch.randint(0, 40, (B, S), device="cuda")            # 371 tokens over 40 ids: long runs
    ids[0, :5] = 999
    up = torch.randn(B, S, H, device="cuda", dtype=dtype)
    wte2 = torch.nn.Parameter(wte.detach().clone())
    wpe2 = torch.nn.Parameter(wpe.detach().clone()) if with_pos else None
    ref = F.embedding(ids, wte2)
    if with_pos:
        ref = ref + wpe2[:S]
    (ref * up).sum().backward()
    if grads_exist:
        wte.grad = torch.full_like(wte, 0.25)
        if with_pos:
            wpe.grad = torch.full_like(wpe, -0.5)
    out = embed(ids, wte, wpe)
    assert torch.equal(out, ref)
    (out * up).sum().backward()
    tol = dict(rtol=2e-2, atol=6e-2) if dtype == torch.bfloat16 else dict(rtol=1e-5, atol=1e-5)
    assert torch.allclose(wte.grad.float(), wte2.grad.float() + (0.25 if grads_exist else 0.0), **tol)
    if with_pos:
        assert torch.allclose(wpe.grad.float(), wpe2.grad.float() + (-0.5 if grads_exist else 0.0), **tol)
    # deterministic: a second backward on fresh buffers gives identical bits
    g1 = wte.grad.clone()
    wte.grad = torch.full_like(wte, 0.25) if grads_exist else None
    if with_pos:
        wpe.grad = None
    (embed(ids, wte, wpe) * up).sum().backward()
    assert torch.equal(wte.grad, g1)

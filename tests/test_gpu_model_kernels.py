"""GPU tier: N3 (add+LayerNorm/RMSNorm), N4 (RoPE), N5 (chunked cross-entropy) against
plain PyTorch fp32 references, forward and backward, plus the models built on them."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _tol(dtype):
    return dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("cols", [64, 768, 1024, 4096, 8192])
@pytest.mark.parametrize("rms", [False, True])
@pytest.mark.parametrize("with_res", [True, False])
def test_add_norm_fwd_bwd(dtype, cols, rms, with_res):
    from distributed_lion_pytorch_b200.ops.norm import add_layer_norm, add_rms_norm
    torch.manual_seed(cols)
    rows = 1000 if cols <= 1024 else 301
    x = torch.randn(rows, cols, device="cuda").to(dtype).requires_grad_()
    r = torch.randn(rows, cols, device="cuda").to(dtype).requires_grad_() if with_res else None
    w = (1 + 0.1 * torch.randn(cols, device="cuda")).to(dtype).requires_grad_()
    b = (0.1 * torch.randn(cols, device="cuda")).to(dtype).requires_grad_() if not rms else None
    eps = 1e-5
    y, s = (add_rms_norm(x, r, w, eps) if rms else add_layer_norm(x, r, w, b, eps))
    dy = torch.randn_like(y)
    ds = torch.randn_like(y)
    (y.float() * dy.float()).sum().add((s.float() * ds.float()).sum()).backward()
    # fp32 reference of the same op
    xf, wf = x.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    rf = r.detach().float().requires_grad_() if with_res else None
    bf = b.detach().float().requires_grad_() if b is not None else None
    sf = xf + rf if with_res else xf
    if dtype == torch.bfloat16 and with_res:
        sf = sf + (sf.detach().to(dtype).float() - sf.detach())       # the stream is stored in bf16
    if rms:
        yf = wf * (sf * torch.rsqrt(sf.pow(2).mean(-1, keepdim=True) + eps))
    else:
        yf = F.layer_norm(sf, (cols,), wf, bf, eps)
    (yf * dy.float()).sum().add((sf * ds.float()).sum()).backward()
    t = _tol(dtype)
    assert torch.allclose(y.float(), yf, **t)
    assert torch.allclose(s.float(), sf.detach(), **t)
    assert torch.allclose(x.grad.float(), xf.grad, **t)
    if with_res:
        assert torch.allclose(r.grad.float(), rf.grad, **t)
    wt = dict(rtol=3e-2, atol=0.3) if dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-2)
    assert torch.allclose(w.grad.float(), wf.grad, **wt)
    if b is not None:
        assert torch.allclose(b.grad.float(), bf.grad, **wt)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("hq,hk,d", [(4, 4, 64), (8, 2, 128)])
def test_rope_fwd_bwd(dtype, hq, hk, d):
    from distributed_lion_pytorch_b200.ops.rope import apply_rope, rope_tables, _rope_functional
    torch.manual_seed(0)
    B, S = 3, 37
    cos, sin = rope_tables(64, d, device="cuda")
    q0 = torch.randn(B, S, hq * d, device="cuda").to(dtype)
    k0 = torch.randn(B, S, hk * d, device="cuda").to(dtype)
    ql = q0.clone().requires_grad_()
    kl = k0.clone().requires_grad_()
    q, k = apply_rope(ql * 1.0, kl * 1.0, cos, sin, d)      # the op owns (and overwrites) its inputs
    dq, dk = torch.randn_like(q), torch.randn_like(k)
    ((q.float() * dq.float()).sum() + (k.float() * dk.float()).sum()).backward()
    qf, kf = q0.float().requires_grad_(), k0.float().requires_grad_()
    qr = _rope_functional(qf, cos, sin, S, d)
    kr = _rope_functional(kf, cos, sin, S, d)
    ((qr * dq.float()).sum() + (kr * dk.float()).sum()).backward()
    t = _tol(dtype)
    assert torch.allclose(q.float(), qr, **t) and torch.allclose(k.float(), kr, **t)
    assert torch.allclose(ql.grad.float(), qf.grad, **t) and torch.allclose(kl.grad.float(), kf.grad, **t)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("V", [512, 50257, 32000])
def test_linear_cross_entropy_fwd_bwd(dtype, V):
    from distributed_lion_pytorch_b200.ops.cross_entropy import linear_cross_entropy, linear_log_probs
    torch.manual_seed(V)
    N, H = 700, 96
    h = (torch.randn(N, H, device="cuda") * 0.5).to(dtype).requires_grad_()
    w = (torch.randn(V, H, device="cuda") * 0.2).to(dtype).requires_grad_()
    labels = torch.randint(0, V, (N,), device="cuda")
    labels[::7] = -100
    loss = linear_cross_entropy(h, w, labels, chunk=256)
    (loss * 0.5).backward()
    hf, wf = h.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    lf = F.cross_entropy(hf @ wf.t(), labels, ignore_index=-100)
    (lf * 0.5).backward()
    t = dict(rtol=3e-2, atol=3e-3) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-5)
    assert torch.allclose(loss.float(), lf, rtol=t["rtol"], atol=2e-2 if dtype == torch.bfloat16 else 1e-4)
    assert torch.allclose(h.grad.float(), hf.grad, **t)
    assert torch.allclose(w.grad.float(), wf.grad, **t)
    # per-token log-probs (DPO building block) with a non-uniform upstream gradient
    h2 = h.detach().clone().requires_grad_()
    w2 = w.detach().clone().requires_grad_()
    lp = linear_log_probs(h2, w2, labels, chunk=300)
    coef = torch.randn(N, device="cuda")
    (lp * coef).sum().backward()
    hf2, wf2 = h.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    lpf = -F.cross_entropy(hf2 @ wf2.t(), labels, ignore_index=-100, reduction="none")
    (lpf * coef).sum().backward()
    assert torch.allclose(lp, lpf, rtol=t["rtol"], atol=5e-2 if dtype == torch.bfloat16 else 1e-4)
    assert torch.allclose(h2.grad.float(), hf2.grad, rtol=5e-2, atol=2e-2 if dtype == torch.bfloat16 else 1e-4)
    assert torch.allclose(w2.grad.float(), wf2.grad, rtol=5e-2, atol=5e-2 if dtype == torch.bfloat16 else 1e-4)


@pytest.mark.parametrize("family", ["gpt2-tiny", "llama-tiny"])
def test_model_gpu_matches_cpu_reference_math(family):
    """The kernel-backed model on the GPU against the same weights run through the plain
    PyTorch fallbacks on the CPU (fp32)."""
    from distributed_lion_pytorch_b200.models import build_model
    torch.manual_seed(0)
    m_cpu = build_model(family)
    m_cpu.eval()
    m_gpu = build_model(family)
    m_gpu.load_state_dict(m_cpu.state_dict())
    m_gpu = m_gpu.cuda().eval()
    ids = torch.randint(0, 512, (2, 48))
    out_c = m_cpu(ids, labels=ids)
    out_g = m_gpu(ids.cuda(), labels=ids.cuda())
    assert abs(float(out_c.loss) - float(out_g.loss)) < 2e-3
    out_c.loss.backward()
    out_g.loss.backward()
    for (n1, p1), (_, p2) in zip(m_cpu.named_parameters(), m_gpu.named_parameters()):
        assert torch.allclose(p1.grad, p2.grad.cpu(), rtol=2e-2, atol=2e-4), n1


def test_lion_optimizer_on_gpu_uses_kernels_and_matches_reference(ref_mod):
    from distributed_lion import Lion
    from distributed_lion_pytorch_b200 import ops
    torch.manual_seed(0)
    net_a = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.Tanh(), torch.nn.Linear(256, 10)).cuda()
    net_b = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.Tanh(), torch.nn.Linear(256, 10)).cuda()
    net_b.load_state_dict(net_a.state_dict())
    oa = ref_mod.Lion(net_a.parameters(), lr=1e-3, weight_decay=0.1)
    ob = Lion(net_b.parameters(), lr=1e-3, weight_decay=0.1)
    x = torch.randn(32, 64, device="cuda")
    for _ in range(5):
        for net, o in ((net_a, oa), (net_b, ob)):
            o.zero_grad()
            net(x).pow(2).mean().backward()
            o.step()
    assert ops.library_loaded()
    for pa, pb in zip(net_a.parameters(), net_b.parameters()):
        assert torch.allclose(pa, pb, rtol=0, atol=1e-7)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("in_out", [True, False])
@pytest.mark.parametrize("with_bias", [True, False])
def test_fused_linear_accumulates_into_existing_grads(dtype, in_out, with_bias):
    """Weight gradient accumulated by the GEMM epilogue into a pre-existing .grad, bias gradient
    from the column-sum kernel — against plain autograd in fp32."""
    from distributed_lion_pytorch_b200.ops.linear import column_sum, fused_linear
    torch.manual_seed(0)
    rows, cin, cout = 3001, 256, 776
    x = torch.randn(7, rows // 7 + 1, cin, device="cuda").to(dtype).requires_grad_()
    w = (torch.randn(cin, cout, device="cuda") * 0.05).to(dtype)
    w = (w if in_out else w.t().contiguous()).requires_grad_()
    b = torch.randn(cout, device="cuda").to(dtype).requires_grad_() if with_bias else None
    g0 = torch.randn_like(w)
    w.grad = g0.clone()                               # what Lion's flat gradient bucket provides
    if with_bias:
        b.grad = torch.ones_like(b)
    dy = torch.randn(*x.shape[:-1], cout, device="cuda").to(dtype)
    for _ in range(2):                                # two micro-batches accumulate
        y = fused_linear(x, w, b, in_out)
        y.backward(dy)
    xf, wf = x.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    bf = b.detach().float().requires_grad_() if with_bias else None
    yf = xf @ (wf if in_out else wf.t())
    if with_bias:
        yf = yf + bf
    yf.backward(dy.float())
    t = dict(rtol=3e-2, atol=0.5) if dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-2)
    assert torch.allclose(y.float(), yf, rtol=3e-2 if dtype == torch.bfloat16 else 1e-4, atol=3e-2)
    assert torch.allclose(x.grad.float(), 2 * xf.grad, **t)
    assert torch.allclose(w.grad.float(), g0.float() + 2 * wf.grad, **t)
    if with_bias:
        assert torch.allclose(b.grad.float(), 1 + 2 * bf.grad, **t)
    cs = column_sum(dy.reshape(-1, cout), torch.float32)
    assert torch.allclose(cs, dy.reshape(-1, cout).float().sum(0), rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rms,with_res,cols,rows", [(False, True, 768, 512), (False, False, 768, 512), (True, True, 4096, 512),
                                                    (False, True, 768, 2 * 3108 + 5)])      # last: several rows per CTA
def test_add_norm_with_fused_dropout(dtype, rms, with_res, cols, rows):
    """dropout(x) + residual + norm in one kernel: the keep-mask is regenerated from the seed in
    the backward.  The mask is recovered from the kernel's own output and the op is then compared
    with plain autograd using that fixed mask."""
    from distributed_lion_pytorch_b200.ops.norm import _FusedNorm
    torch.manual_seed(1)
    p, seed = 0.3, 12345
    x = (torch.randn(rows, cols, device="cuda") + 8.0).to(dtype).requires_grad_()       # kept values stay far from zero
    r = torch.randn(rows, cols, device="cuda").to(dtype).requires_grad_() if with_res else None
    w = (1 + 0.1 * torch.randn(cols, device="cuda")).to(dtype).requires_grad_()
    b = None if rms else (0.1 * torch.randn(cols, device="cuda")).to(dtype).requires_grad_()
    y, s = _FusedNorm.apply(x, r, w, b, 1e-5, rms, p, seed)
    y2, s2 = _FusedNorm.apply(x.detach(), None if r is None else r.detach(), w.detach(), None if b is None else b.detach(),
                              1e-5, rms, p, seed)
    assert torch.equal(y, y2) and torch.equal(s, s2)                                     # same seed, same mask
    y3, _ = _FusedNorm.apply(x.detach(), None if r is None else r.detach(), w.detach(), None if b is None else b.detach(),
                             1e-5, rms, p, seed + 1)
    assert not torch.equal(y, y3)
    base = s.detach().float() - (r.detach().float() if with_res else 0)
    keep = base.abs() > (0.5 if dtype == torch.bfloat16 else 1e-3)                        # x ~ N(8,1)/0.7: kept values are large
    frac = keep.float().mean().item()
    assert abs(frac - (1 - p)) < 0.01, frac
    dy, ds = torch.randn_like(y), torch.randn_like(y)
    (y.float() * dy.float()).sum().add((s.float() * ds.float()).sum()).backward()
    xf, wf = x.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    rf = r.detach().float().requires_grad_() if with_res else None
    bf = b.detach().float().requires_grad_() if b is not None else None
    xd = xf * keep.float() / (1 - p)
    sf = xd + rf if with_res else xd
    if rms:
        yf = wf * (sf * torch.rsqrt(sf.pow(2).mean(-1, keepdim=True) + 1e-5))
    else:
        yf = F.layer_norm(sf, (cols,), wf, bf, 1e-5)
    (yf * dy.float()).sum().add((sf * ds.float()).sum()).backward()
    t = dict(rtol=3e-2, atol=6e-2) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)
    assert torch.allclose(y.float(), yf, **t)
    assert torch.allclose(x.grad.float(), xf.grad, **t)
    assert (x.grad[~keep] == 0).all()                                                     # dropped inputs get no gradient
    if with_res:
        assert torch.allclose(r.grad.float(), rf.grad, **t)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_activation_kernels(dtype):
    from distributed_lion_pytorch_b200.ops.activation import gelu_tanh, swiglu
    torch.manual_seed(0)
    x = (torch.randn(333, 3072, device="cuda") * 2).to(dtype).requires_grad_()
    dy = torch.randn_like(x)
    y = gelu_tanh(x)
    y.backward(dy)
    xf = x.detach().float().requires_grad_()
    yf = F.gelu(xf, approximate="tanh")
    yf.backward(dy.float())
    t = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=1e-5, atol=1e-5)
    assert torch.allclose(y.float(), yf, **t) and torch.allclose(x.grad.float(), xf.grad, **t)
    g = (torch.randn(129, 11008, device="cuda") * 2).to(dtype).requires_grad_()
    u = torch.randn(129, 11008, device="cuda").to(dtype).requires_grad_()
    do = torch.randn_like(g)
    o = swiglu(g, u)
    o.backward(do)
    gf, uf = g.detach().float().requires_grad_(), u.detach().float().requires_grad_()
    of = F.silu(gf) * uf
    of.backward(do.float())
    t = dict(rtol=3e-2, atol=3e-2) if dtype == torch.bfloat16 else dict(rtol=1e-5, atol=1e-5)
    assert torch.allclose(o.float(), of, **t)
    assert torch.allclose(g.grad.float(), gf.grad, **t) and torch.allclose(u.grad.float(), uf.grad, **t)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("cols,rows,rms", [(768, 2 * 3108 + 17, False), (4096, 3 * 1184 + 5, True), (64, 3 * 9472 + 1, False)])
def test_add_norm_many_rows_per_cta(dtype, cols, rows, rms):
    """More rows than resident CTAs: every CTA walks several rows, which is the software-pipelined path of the
    forward (next row's loads in flight while this one is reduced) and the partial-row accumulation of the backward."""
    from distributed_lion_pytorch_b200.ops.norm import add_layer_norm, add_rms_norm
    torch.manual_seed(rows)
    x = torch.randn(rows, cols, device="cuda").to(dtype).requires_grad_()
    r = torch.randn(rows, cols, device="cuda").to(dtype).requires_grad_()
    w = (1 + 0.1 * torch.randn(cols, device="cuda")).to(dtype).requires_grad_()
    b = (0.1 * torch.randn(cols, device="cuda")).to(dtype).requires_grad_() if not rms else None
    y, s = add_rms_norm(x, r, w, 1e-5) if rms else add_layer_norm(x, r, w, b, 1e-5)
    dy = torch.randn_like(y)
    y.backward(dy)
    xf, rf, wf = (t.detach().float().requires_grad_() for t in (x, r, w))
    sf = xf + rf
    if dtype == torch.bfloat16:
        sf = sf + (sf.detach().to(dtype).float() - sf.detach())
    if rms:
        yf = wf * (sf * torch.rsqrt(sf.pow(2).mean(-1, keepdim=True) + 1e-5))
    else:
        bf = b.detach().float().requires_grad_()
        yf = F.layer_norm(sf, (cols,), wf, bf, 1e-5)
    yf.backward(dy.float())
    t = _tol(dtype)
    assert torch.allclose(y.float(), yf, **t)
    assert torch.equal(s.float(), sf.detach()) or torch.allclose(s.float(), sf.detach(), **t)
    assert torch.allclose(x.grad.float(), xf.grad, **t)
    assert torch.allclose(r.grad.float(), rf.grad, **t)
    scale = float(wf.grad.abs().max())
    assert float((w.grad.float() - wf.grad).abs().max()) < (3e-2 if dtype == torch.bfloat16 else 1e-3) * scale


@pytest.mark.parametrize("family", ["gpt2-tiny", "llama-tiny"])
def test_gradient_checkpointing_replays_fused_dropout_on_gpu(family):
    """Activation checkpointing recomputes the block: the in-kernel dropout masks (seed drawn from the CPU generator)
    and cuDNN's attention dropout (CUDA generator) must replay, so gradients match the uncheckpointed run."""
    from distributed_lion_pytorch_b200.models import build_model
    torch.manual_seed(0)
    m = build_model(family, dtype=torch.bfloat16, device="cuda").train()
    ids = torch.randint(0, 512, (4, 64), device="cuda")
    runs = []
    for ckpt in (False, True):
        m.zero_grad()
        (m.gradient_checkpointing_enable if ckpt else m.gradient_checkpointing_disable)()
        torch.manual_seed(7)
        loss = m(ids, labels=ids).loss
        loss.backward()
        runs.append((float(loss), {n: p.grad.float().clone() for n, p in m.named_parameters()}))
    assert runs[0][0] == runs[1][0]
    for n, g in runs[0][1].items():
        assert float((g - runs[1][1][n]).abs().max()) <= 1e-2 * float(g.abs().max()) + 1e-6, n


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_one_launch_column_sum_repeated_calls_and_accumulation(dtype):
    """The fused column sum keeps its tile tickets zeroed between calls: many calls of different shapes in a row,
    writing and accumulating, all agree with torch."""
    from distributed_lion_pytorch_b200.ops import _cabi
    from distributed_lion_pytorch_b200.ops.linear import column_sum
    torch.manual_seed(0)
    tol = dict(rtol=2e-2, atol=0.5) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-3)
    for rep in range(3):
        for rows, cols in [(20480, 768), (4096, 3072), (1000, 2304), (7, 8), (20480, 768), (333, 16384)]:
            x = torch.randn(rows, cols, device="cuda").to(dtype)
            n0 = _cabi.LAUNCHES
            out = column_sum(x, dtype)
            assert _cabi.LAUNCHES == n0 + 1                     # one kernel, no separate fold
            assert torch.allclose(out.float(), x.float().sum(0), **tol)
            g0 = torch.randn(cols, device="cuda").to(dtype)
            g = g0.clone()
            n0 = _cabi.LAUNCHES
            assert column_sum(x, dtype, accumulate_into=g) is None
            assert _cabi.LAUNCHES == n0 + 1
            assert torch.allclose(g.float(), g0.float() + x.float().sum(0), **tol)
    # a different output dtype takes the two-launch path and still agrees
    x = torch.randn(512, 256, device="cuda").to(dtype)
    other = torch.float32 if dtype == torch.bfloat16 else torch.bfloat16
    assert torch.allclose(column_sum(x, other).float(), x.float().sum(0), rtol=2e-2, atol=0.5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rms", [False, True])
def test_add_norm_accumulates_into_existing_grads(dtype, rms):
    from distributed_lion_pytorch_b200.ops.norm import add_layer_norm, add_rms_norm
    torch.manual_seed(3)
    rows, cols = 4000, 768
    x = torch.randn(rows, cols, device="cuda").to(dtype).requires_grad_()
    w = (1 + 0.1 * torch.randn(cols, device="cuda")).to(dtype).requires_grad_()
    b = None if rms else (0.1 * torch.randn(cols, device="cuda")).to(dtype).requires_grad_()
    w0 = torch.randn(cols, device="cuda").to(dtype)
    b0 = torch.randn(cols, device="cuda").to(dtype)
    w.grad = w0.clone()
    if b is not None:
        b.grad = b0.clone()
    wg_ptr = w.grad.data_ptr()
    y, _ = add_rms_norm(x, None, w, 1e-5) if rms else add_layer_norm(x, None, w, b, 1e-5)
    dy = torch.randn_like(y)
    y.backward(dy)
    assert w.grad.data_ptr() == wg_ptr                           # written in place, not replaced
    xf, wf = x.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    if rms:
        yf = wf * (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5))
    else:
        bf = b.detach().float().requires_grad_()
        yf = F.layer_norm(xf, (cols,), wf, bf, 1e-5)
    yf.backward(dy.float())
    tol = dict(rtol=3e-2, atol=0.5) if dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-2)
    assert torch.allclose(w.grad.float(), w0.float() + wf.grad, **tol)
    if b is not None:
        assert torch.allclose(b.grad.float(), b0.float() + bf.grad, **tol)

"""CPU tier: the autograd plumbing around the kernels (chunking, padded vocab, in-place RoPE,
gradient-in-forward cross-entropy) with torch stand-ins that honour the kernels' contracts."""
import pytest
import torch
import torch.nn.functional as F


@pytest.fixture()
def emulate(monkeypatch):
    from distributed_lion_pytorch_b200.ops import cross_entropy, rope
    monkeypatch.setattr(cross_entropy, "_EMULATE", True)
    monkeypatch.setattr(rope, "_EMULATE", True)


@pytest.mark.parametrize("V", [97, 512])
def test_linear_cross_entropy_plumbing(emulate, V):
    from distributed_lion_pytorch_b200.ops.cross_entropy import linear_cross_entropy, linear_log_probs
    torch.manual_seed(0)
    N, H = 50, 16
    h = torch.randn(N, H, requires_grad=True)
    w = torch.randn(V, H, requires_grad=True)
    labels = torch.randint(0, V, (N,))
    labels[::5] = -100
    loss = linear_cross_entropy(h, w, labels, chunk=16)
    (loss * 0.25).backward()
    h2, w2 = h.detach().clone().requires_grad_(), w.detach().clone().requires_grad_()
    ref = F.cross_entropy(h2 @ w2.t(), labels, ignore_index=-100)
    (ref * 0.25).backward()
    assert torch.allclose(loss, ref, atol=1e-5)
    assert torch.allclose(h.grad, h2.grad, atol=1e-5) and torch.allclose(w.grad, w2.grad, atol=1e-5)
    h3, w3 = h.detach().clone().requires_grad_(), w.detach().clone().requires_grad_()
    coef = torch.randn(N)
    (linear_log_probs(h3, w3, labels, chunk=16) * coef).sum().backward()
    h4, w4 = h.detach().clone().requires_grad_(), w.detach().clone().requires_grad_()
    (-F.cross_entropy(h4 @ w4.t(), labels, ignore_index=-100, reduction="none") * coef).sum().backward()
    assert torch.allclose(h3.grad, h4.grad, atol=1e-5) and torch.allclose(w3.grad, w4.grad, atol=1e-5)


def test_llama_with_inplace_rope_matches_functional(emulate, monkeypatch):
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.ops import cross_entropy, rope
    torch.manual_seed(0)
    m = build_model("llama-tiny")
    ids = torch.randint(0, 512, (2, 24))
    calls = {"n": 0}
    orig = rope._Rope.apply

    def spy(*a):
        calls["n"] += 1
        return orig(*a)
    monkeypatch.setattr(rope._Rope, "apply", staticmethod(spy))
    m(ids, labels=ids).loss.backward()
    g1 = [p.grad.clone() for p in m.parameters()]
    assert calls["n"] == m.config.num_hidden_layers          # the in-place autograd path really ran
    monkeypatch.setattr(rope, "_EMULATE", False)
    monkeypatch.setattr(cross_entropy, "_EMULATE", False)
    m.zero_grad()
    m(ids, labels=ids).loss.backward()
    for a, p in zip(g1, m.parameters()):
        assert torch.allclose(a, p.grad, atol=1e-5)


def test_gpt2_tied_embedding_gets_both_gradients(emulate):
    from distributed_lion_pytorch_b200.models import build_model
    torch.manual_seed(0)
    m = build_model("gpt2-tiny")
    m.eval()
    assert m.lm_head.weight is m.transformer.wte.weight
    ids = torch.randint(0, 512, (2, 16))
    m(ids, labels=ids).loss.backward()
    g = m.transformer.wte.weight.grad.clone()
    m.zero_grad()
    out = m(ids, return_logits=True)
    F.cross_entropy(out.logits[:, :-1].reshape(-1, 512), ids[:, 1:].reshape(-1)).backward()
    assert torch.allclose(g, m.transformer.wte.weight.grad, atol=1e-5)


def test_linear_cross_entropy_under_autocast_with_fp32_master_weights(emulate):
    """Mixed precision: fp32 parameters under autocast (``Trainer.autocast`` for --bf16 without --torch_dtype).
    The logits GEMMs run in bf16, ``out=`` / in-place GEMMs are not autocast-aware, and the gradients must come
    back in the parameters' fp32 (this path used to raise 'Expected out tensor to have dtype
    BFloat16')."""
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.ops.cross_entropy import linear_cross_entropy, linear_log_probs
    torch.manual_seed(0)
    N, H, V = 40, 16, 97
    h = torch.randn(N, H, requires_grad=True)
    w = torch.randn(V, H, requires_grad=True)
    labels = torch.randint(0, V, (N,))
    with torch.autocast("cpu", dtype=torch.bfloat16):
        loss = linear_cross_entropy(h, w, labels, chunk=16)
        lp = linear_log_probs(h, w, labels, chunk=16)
    (loss + lp.sum() * 0.01).backward()
    assert h.grad.dtype == torch.float32 and w.grad.dtype == torch.float32
    ref = F.cross_entropy(h.detach() @ w.detach().t(), labels)
    assert abs(float(loss) - float(ref)) < 0.15            # bf16 logits
    m = build_model("gpt2-tiny")
    ids = torch.randint(0, 512, (2, 16))
    with torch.autocast("cpu", dtype=torch.bfloat16):
        out = m(ids, labels=ids)
    out.loss.backward()
    assert all(p.grad is not None and p.grad.dtype == torch.float32 for p in m.parameters())


def test_tensor_core_ops_decline_on_cpu_and_callers_fall_back():
    """The tcgen05 kernels (LM head, MLP backward) are CUDA-only: on CPU their entry points say "not applicable" and the
    callers run the plain paths with the same results (ops/lm_head.py, ops/mlp.py)."""
    import torch.nn.functional as F
    from distributed_lion_pytorch_b200.ops import cross_entropy as ce, lm_head, mlp
    torch.manual_seed(0)
    h = torch.randn(37, 32)
    w = torch.randn(101, 32)
    labels = torch.randint(0, 101, (37,))
    labels[::5] = -100
    assert not lm_head.supported(h, w) and not lm_head.supported(h.bfloat16(), w.bfloat16())
    assert ce._fused_no_grad(h, w, labels, -100) is None
    with torch.no_grad():
        loss = ce.linear_cross_entropy(h, w, labels)
        lp = ce.linear_log_probs(h, w, labels)
    ref = F.cross_entropy(h @ w.t(), labels, ignore_index=-100)
    torch.testing.assert_close(loss, ref, atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(-lp.sum() / (labels != -100).sum(), ref, atol=1e-5, rtol=1e-5)
    x = torch.randn(2, 5, 16, requires_grad=True)
    assert mlp.gelu_mlp(x, torch.randn(16, 64), torch.zeros(64), torch.randn(64, 16), torch.zeros(16)) is None


def test_fused_mlp_node_plumbing_on_cpu(monkeypatch):
    """The one-node GPT-2 MLP (ops/mlp.py) with torch stand-ins for its kernels: gradients equal plain autograd's, a
    second backward accumulates into the live gradients and marks them written, and the down-projection's bias gradient
    arrives through the producer hand-off when the incoming gradient carries it (reference: HF GPT2MLP under autograd,
    async_trainer.py:33)."""
    import torch.nn.functional as F
    from distributed_lion_pytorch_b200.ops import linear, mlp
    monkeypatch.setattr(mlp, "_EMULATE", True)
    torch.manual_seed(0)
    E, I = 16, 64
    x = torch.randn(3, 5, E, requires_grad=True)
    w_fc = torch.nn.Parameter(torch.randn(E, I) * E ** -0.5)
    b_fc = torch.nn.Parameter(torch.randn(I) * 0.1)
    w_pr = torch.nn.Parameter(torch.randn(I, E) * I ** -0.5)
    b_pr = torch.nn.Parameter(torch.randn(E) * 0.1)
    dy = torch.randn(3, 5, E)
    y = mlp.gelu_mlp(x, w_fc, b_fc, w_pr, b_pr)
    assert y is not None
    y.backward(dy)
    got = [t.grad.clone() for t in (x, w_fc, b_fc, w_pr, b_pr)]
    ref_in = [t.detach().clone().requires_grad_(True) for t in (x, w_fc, b_fc, w_pr, b_pr)]
    yr = F.gelu(ref_in[0] @ ref_in[1] + ref_in[2], approximate="tanh") @ ref_in[3] + ref_in[4]
    yr.backward(dy)
    torch.testing.assert_close(y, yr, atol=1e-5, rtol=1e-5)
    for a, r in zip(got, ref_in):
        torch.testing.assert_close(a, r.grad, atol=1e-4, rtol=1e-4)
    # steady state: gradients exist, the node adds into them itself (autograd gets None) and marks them
    for t in (w_fc, b_fc, w_pr, b_pr):
        t._dlion_touched = False
    mlp.gelu_mlp(x, w_fc, b_fc, w_pr, b_pr).backward(dy)
    for t, a in zip((x, w_fc, b_fc, w_pr, b_pr), got):
        torch.testing.assert_close(t.grad, 2 * a, atol=2e-4, rtol=1e-4)
    assert all(getattr(t, "_dlion_touched", False) for t in (w_fc, b_fc, w_pr, b_pr))
    # bias hand-off: the producer of dy attached the column sums for b_proj -> no pass over dy, the sums are used as is
    before = linear.HANDOFFS
    y3 = mlp.gelu_mlp(x, w_fc, b_fc, w_pr, b_pr)
    g = dy.clone()
    fake = torch.full((E,), 7.0)
    linear.attach_column_sums(g, b_pr, fake)
    old = b_pr.grad.clone()
    y3.backward(g)
    assert linear.HANDOFFS == before + 1
    torch.testing.assert_close(b_pr.grad, old + fake)
    # switched off / no bias / autograd off: the caller gets None and runs the plain modules
    monkeypatch.setattr(mlp, "_ENABLED", False)
    assert mlp.gelu_mlp(x, w_fc, b_fc, w_pr, b_pr) is None
    monkeypatch.setattr(mlp, "_ENABLED", True)
    assert mlp.gelu_mlp(x, w_fc, None, w_pr, b_pr) is None
    with torch.no_grad():
        assert mlp.gelu_mlp(x, w_fc, b_fc, w_pr, b_pr) is None


def test_gpt2_with_the_fused_mlp_node_equals_the_plain_modules_on_cpu(monkeypatch):
    from distributed_lion_pytorch_b200.models import build_model
    from distributed_lion_pytorch_b200.ops import mlp
    torch.manual_seed(0)
    model = build_model("gpt2-tiny", dtype=torch.float32, device="cpu", pretrained=False).eval()
    ids = torch.randint(0, 512, (2, 32))
    monkeypatch.setattr(mlp, "_EMULATE", True)
    calls = []
    real = mlp.dgrad_dgelu
    monkeypatch.setattr(mlp, "dgrad_dgelu", lambda *a, **k: calls.append(1) or real(*a, **k))
    model(ids, labels=ids).loss.backward()
    assert len(calls) == model.config.n_layer
    fused = {n: p.grad.clone() for n, p in model.named_parameters()}
    model.zero_grad(set_to_none=True)
    monkeypatch.setattr(mlp, "_EMULATE", False)
    model(ids, labels=ids).loss.backward()
    assert len(calls) == model.config.n_layer
    for n, p in model.named_parameters():
        torch.testing.assert_close(fused[n], p.grad, atol=1e-5, rtol=1e-4, msg=n)


def test_lm_head_dispatch_on_cpu(monkeypatch):
    """Who gets the logits-free LM-head kernel (ops/lm_head.py): calls that need no gradient — evaluation loss, frozen
    reference model, no_grad log-probs — and nobody else; with a CPU stand-in for the kernel the dispatch is testable
    here (reference: run_clm.py:626-639 evaluation, dpo_llama2.py:219-231 reference log-probs)."""
    import torch.nn.functional as F
    from distributed_lion_pytorch_b200.ops import cross_entropy as ce, lm_head
    monkeypatch.setattr(lm_head, "_EMULATE", True)
    torch.manual_seed(0)
    h = torch.randn(37, 32)
    w = torch.nn.Parameter(torch.randn(101, 32))
    labels = torch.randint(0, 101, (37,))
    labels[::5] = -100
    calls = []
    real = lm_head.lm_head_log_probs
    monkeypatch.setattr(lm_head, "lm_head_log_probs", lambda *a, **k: calls.append(1) or real(*a, **k))
    ref = F.cross_entropy(h @ w.detach().t(), labels, ignore_index=-100)
    ref_lp = -F.cross_entropy(h @ w.detach().t(), labels, ignore_index=-100, reduction="none")
    with torch.no_grad():                                  # evaluation: fused
        torch.testing.assert_close(ce.linear_cross_entropy(h, w, labels), ref, atol=1e-5, rtol=1e-5)
        torch.testing.assert_close(ce.linear_log_probs(h, w, labels), ref_lp, atol=1e-5, rtol=1e-5)
    assert len(calls) == 2
    frozen = w.detach()                                    # frozen reference model with grad mode on: fused
    torch.testing.assert_close(ce.linear_log_probs(h, frozen, labels), ref_lp, atol=1e-5, rtol=1e-5)
    assert len(calls) == 3
    loss = ce.linear_cross_entropy(h, w, labels)           # training: not the no-grad kernel, and a gradient comes back
    loss.backward()
    assert len(calls) == 3 and w.grad is not None
    with torch.no_grad():                                  # an ignore_index that is a real token id: not fused
        ce.linear_cross_entropy(h, w, labels.clamp(min=0), ignore_index=3)
    assert len(calls) == 3
    monkeypatch.setenv("DLION_LMHEAD_FUSED", "0")          # switched off
    with torch.no_grad():
        torch.testing.assert_close(ce.linear_cross_entropy(h, w, labels), ref, atol=1e-5, rtol=1e-5)
    assert len(calls) == 3


def test_tensor_core_kernel_planning_is_host_only_and_sane():
    """`dlion_lmhead_plan` / `dlion_dgrad_dgelu_parts` are plain host functions of the library (callable without a GPU):
    the vocabulary split count keeps every SM busy without shredding the vocabulary, and the workspace sizes are what the
    Python wrappers allocate (csrc/lmhead_lse_sm100.cu, csrc/gemm_dgelu_sm100.cu)."""
    from distributed_lion_pytorch_b200.ops import _cabi, lm_head
    if not _cabi.available():
        pytest.skip("libdlion.so not built")
    for rows, vocab in ((128, 256), (4096, 50257), (20480, 50257), (8192, 32000), (515, 32000), (1, 8)):
        splits, ws = lm_head.plan(rows, vocab)
        m_tiles, n_tiles = (rows + 127) // 128, (vocab + 255) // 256
        assert 1 <= splits <= min(n_tiles, 64)
        assert ws == (2 * splits + 1) * rows
        units = m_tiles * splits
        waves = -(-units // 148)
        # no worse than 25 % over the ideal tile count per SM once there is more than a wave of work
        if m_tiles * n_tiles >= 4 * 148:
            assert waves * -(-n_tiles // splits) <= 1.25 * (m_tiles * n_tiles / 148) + 1
    L = _cabi.lib()
    assert L.dlion_dgrad_dgelu_parts(20480) == 160 * 4 and L.dlion_dgrad_dgelu_parts(1) == 4
    assert L.dlion_dgrad_dgelu_parts(129) == 2 * 4

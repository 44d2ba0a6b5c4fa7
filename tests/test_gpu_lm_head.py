"""tcgen05 LM-head log-sum-exp kernel (csrc/lmhead_lse_sm100.cu, ops/lm_head.py) against a plain fp32 PyTorch
reference of the same op: lse = logsumexp(h @ w.T), logp = (h @ w.T)[label] - lse.
reference: HF ForCausalLM evaluation loss (run_clm.py:626-639); TRL DPO reference log-probs (dpo_llama2.py:219-231)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(h, w, labels):
    logits = h.float() @ w.float().t()
    lse = torch.logsumexp(logits, dim=-1)
    valid = (labels >= 0) & (labels < w.shape[0])
    picked = logits.gather(1, labels.clamp(0, w.shape[0] - 1)[:, None]).squeeze(1)
    return torch.where(valid, picked - lse, torch.zeros_like(lse)), lse


def _data(rows, vocab, hidden, dtype, seed=0, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    h = (torch.randn(rows, hidden, device="cuda", generator=g) * scale).to(dtype)
    w = (torch.randn(vocab, hidden, device="cuda", generator=g) * (hidden ** -0.5)).to(dtype)
    labels = torch.randint(0, vocab, (rows,), device="cuda", generator=g)
    labels[::7] = -100
    return h, w, labels


@pytest.mark.parametrize("rows,vocab,hidden", [(128, 256, 64), (300, 1000, 72), (1000, 50257, 768), (515, 32000, 4096),
                                                (4096, 50257, 768)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_lm_head_matches_fp32_reference(rows, vocab, hidden, dtype):
    from distributed_lion_pytorch_b200.ops import lm_head
    if dtype == torch.float16 and rows > 1000:
        pytest.skip("one large fp16 case is enough")
    h, w, labels = _data(rows, vocab, hidden, dtype, seed=rows + vocab)
    assert lm_head.supported(h, w)
    logp, lse = lm_head.lm_head_log_probs(h, w, labels)
    torch.cuda.synchronize()
    ref_logp, ref_lse = _ref(h, w, labels)
    assert torch.isfinite(lse).all()
    torch.testing.assert_close(lse, ref_lse, atol=2e-3, rtol=1e-4)
    torch.testing.assert_close(logp, ref_logp, atol=3e-3, rtol=1e-4)
    assert (logp[::7] == 0).all()                       # ignored labels give exactly 0


def test_lm_head_large_logits_and_strided_views():
    """Peaked rows (|logit| ~ 60: the running maximum has to move), a row-strided activation view (ld_h > hidden) and a
    vocabulary that is not a multiple of the 256-wide tile or of the 32-column epilogue chunk."""
    from distributed_lion_pytorch_b200.ops import lm_head
    h, w, labels = _data(257, 50001, 256, torch.bfloat16, seed=3, scale=8.0)
    big = torch.zeros(257, 512, device="cuda", dtype=torch.bfloat16)
    big[:, :256] = h
    hv = big[:, :256]                                   # stride 512, no copy needed
    logp, lse = lm_head.lm_head_log_probs(hv, w, labels)
    ref_logp, ref_lse = _ref(h, w, labels)
    assert ref_lse.abs().max() > 20
    torch.testing.assert_close(lse, ref_lse, atol=5e-3, rtol=2e-4)
    torch.testing.assert_close(logp, ref_logp, atol=5e-3, rtol=2e-4)
    only_lse = lm_head.lm_head_log_probs(hv, w, None)[1]
    assert torch.equal(only_lse, lse)                   # deterministic, label-free call gives the same lse


def test_no_grad_cross_entropy_and_log_probs_take_the_fused_path(monkeypatch):
    """linear_cross_entropy / linear_log_probs without a gradient go through the tcgen05 kernel and agree with the
    cuBLAS + CE-kernel path that training uses."""
    from distributed_lion_pytorch_b200.ops import cross_entropy as ce, lm_head
    h, w, labels = _data(2050, 50257, 768, torch.bfloat16, seed=5)
    calls = []
    real = lm_head.lm_head_log_probs
    monkeypatch.setattr(lm_head, "lm_head_log_probs", lambda *a, **k: calls.append(1) or real(*a, **k))
    with torch.no_grad():
        loss_f = ce.linear_cross_entropy(h, w, labels)
        lp_f = ce.linear_log_probs(h, w, labels)
    assert len(calls) == 2
    monkeypatch.setenv("DLION_LMHEAD_FUSED", "0")
    with torch.no_grad():
        loss_s = ce.linear_cross_entropy(h, w, labels)
        lp_s = ce.linear_log_probs(h, w, labels)
    assert len(calls) == 2
    torch.testing.assert_close(loss_f, loss_s, atol=2e-3, rtol=1e-3)      # the split path rounds logits to bf16
    torch.testing.assert_close(lp_f, lp_s, atol=6e-2, rtol=1e-2)
    ref_logp, _ = _ref(h, w, labels)
    assert (lp_f - ref_logp).abs().max() <= (lp_s - ref_logp).abs().max() + 1e-4   # fp32 logits: closer to the truth
    hg = h.clone().requires_grad_(True)
    ce.linear_cross_entropy(hg, w, labels).backward()                     # with a gradient: training path, unchanged
    assert hg.grad is not None and len(calls) == 2

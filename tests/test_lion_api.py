"""API tier (single process, CPU): state_dict schema, param groups, closure, grad=None,
external LambdaLR — against the reference optimizer loaded by path."""
import copy

import pytest
import torch

from distributed_lion import Lion


def _model(seed=0, dtype=torch.float32):
    torch.manual_seed(seed)
    m = torch.nn.Sequential(torch.nn.Linear(13, 7), torch.nn.Tanh(), torch.nn.Linear(7, 3))
    return m.to(dtype)


def _same(a, b, dtype):
    """fp32: bit-exact.  bf16 on CPU: ATen's vectorised loop rounds `alpha` to bf16 and its
    scalar-tail loop does not (probed: 28 % of elements differ by one ulp between the two),
    and flat buckets move elements between the two loops — so bf16-on-CPU is compared to one
    bf16 ulp.  (On CUDA both go through one fp32 functor; the GPU tests compare exactly.)"""
    if dtype == torch.float32:
        return torch.equal(a, b)
    return torch.allclose(a.float(), b.float(), rtol=2 ** -6, atol=3e-5)


def _run(opt, model, steps=4, sched=None, seed=1):
    torch.manual_seed(seed)
    for _ in range(steps):
        x = torch.randn(5, 13).to(next(model.parameters()).dtype)
        loss = model(x).float().pow(2).mean()
        opt.zero_grad()
        loss.backward()
        opt.step()
        if sched is not None:
            sched.step()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_local_matches_reference_bitexact(ref_mod, dtype):
    m_ref, m_new = _model(0, dtype), _model(0, dtype)
    o_ref = ref_mod.Lion(m_ref.parameters(), lr=1e-3, weight_decay=0.1)
    o_new = Lion(m_new.parameters(), lr=1e-3, weight_decay=0.1)
    s_ref = torch.optim.lr_scheduler.LambdaLR(o_ref, lambda s: 1.0 / (1 + s))
    s_new = torch.optim.lr_scheduler.LambdaLR(o_new, lambda s: 1.0 / (1 + s))
    _run(o_ref, m_ref, sched=s_ref)
    _run(o_new, m_new, sched=s_new)
    for a, b in zip(m_ref.parameters(), m_new.parameters()):
        assert _same(a, b, dtype)
    sd_ref, sd_new = o_ref.state_dict(), o_new.state_dict()
    assert sd_ref["param_groups"][0].keys() == sd_new["param_groups"][0].keys()
    assert sd_ref["state"].keys() == sd_new["state"].keys()
    for k in sd_ref["state"]:
        assert list(sd_new["state"][k].keys()) == ["exp_avg"]
        assert _same(sd_ref["state"][k]["exp_avg"], sd_new["state"][k]["exp_avg"], dtype)
        assert sd_new["state"][k]["exp_avg"].dtype == dtype


def test_state_dict_roundtrip_with_reference(ref_mod):
    m_ref, m_new = _model(2), _model(2)
    o_ref = ref_mod.Lion(m_ref.parameters(), lr=1e-3, weight_decay=0.05)
    _run(o_ref, m_ref, steps=2)
    # a reference-produced optimizer.pt loads into ours ...
    m_new.load_state_dict(m_ref.state_dict())
    o_new = Lion(m_new.parameters(), lr=1e-3, weight_decay=0.05)
    o_new.load_state_dict(copy.deepcopy(o_ref.state_dict()))
    _run(o_ref, m_ref, steps=3, seed=7)
    _run(o_new, m_new, steps=3, seed=7)
    for a, b in zip(m_ref.parameters(), m_new.parameters()):
        assert torch.equal(a, b)
    # ... and ours loads into the reference
    o_ref2 = ref_mod.Lion(m_ref.parameters(), lr=1e-3, weight_decay=0.05)
    o_ref2.load_state_dict(o_new.state_dict())
    # state_dict tensors are independent copies, not bucket views
    sd = o_new.state_dict()
    t = sd["state"][0]["exp_avg"]
    assert t.untyped_storage().nbytes() == t.numel() * t.element_size()


def test_load_after_steps_folds_into_bucket():
    m1, m2 = _model(3), _model(3)
    o1, o2 = Lion(m1.parameters(), lr=1e-3), Lion(m2.parameters(), lr=1e-3)
    _run(o1, m1, steps=3)
    _run(o2, m2, steps=1, seed=99)          # o2 already has buckets
    m2.load_state_dict(m1.state_dict())
    o2.load_state_dict(o1.state_dict())
    _run(o1, m1, steps=2, seed=5)
    _run(o2, m2, steps=2, seed=5)
    for a, b in zip(m1.parameters(), m2.parameters()):
        assert torch.equal(a, b)


def test_param_groups_and_closure(ref_mod):
    m_ref, m_new = _model(4), _model(4)

    def groups(m):
        return [{"params": [p for n, p in m.named_parameters() if n.endswith("weight")], "weight_decay": 0.1},
                {"params": [p for n, p in m.named_parameters() if n.endswith("bias")], "weight_decay": 0.0,
                 "lr": 5e-4, "betas": (0.95, 0.98)}]
    o_ref = ref_mod.Lion(groups(m_ref), lr=1e-3)
    o_new = Lion(groups(m_new), lr=1e-3)
    x = torch.randn(4, 13)
    for o, m in ((o_ref, m_ref), (o_new, m_new)):
        def closure():
            o.zero_grad()
            l = m(x).pow(2).sum()
            l.backward()
            return l
        out = o.step(closure)
        assert out is not None and out.requires_grad is True or out is not None
    for a, b in zip(m_ref.parameters(), m_new.parameters()):
        assert torch.equal(a, b)
    assert len(o_new.buckets) == 2


def test_grad_none_params_are_skipped_locally(ref_mod):
    m_ref, m_new = _model(5), _model(5)
    o_ref = ref_mod.Lion(m_ref.parameters(), lr=1e-2, weight_decay=0.1)
    o_new = Lion(m_new.parameters(), lr=1e-2, weight_decay=0.1)
    for m, o in ((m_ref, o_ref), (m_new, o_new)):
        torch.manual_seed(0)
        for step in range(4):
            o.zero_grad(set_to_none=True) if o is o_ref else o.zero_grad()
            x = torch.randn(3, 13)
            h = m[0](x)
            loss = h.pow(2).mean() if step % 2 == 0 else m[2](torch.tanh(h)).pow(2).mean()
            loss.backward()
            if o is o_new and step % 2 == 0:
                for p in m[2].parameters():
                    p.grad = None       # emulate set_to_none semantics for unused params
            o.step()
    for a, b in zip(m_ref.parameters(), m_new.parameters()):
        assert torch.equal(a, b)
    assert len(o_ref.state_dict()["state"]) == len(o_new.state_dict()["state"])


def test_params_alias_flat_bucket_and_flat_grads():
    m = _model(6)
    o = Lion(m.parameters(), lr=1e-3)
    _run(o, m, steps=2)
    b = o.buckets[0]
    for p, pv, gv in zip(b.params, b.p_views, b.g_views):
        assert p.data_ptr() == pv.data_ptr()
        assert p.grad is not None and p.grad.data_ptr() == gv.data_ptr()
    assert b.numel % 8192 == 0 and all(off % 128 == 0 for off in b.offsets)


def test_constructor_validation():
    p = [torch.nn.Parameter(torch.zeros(3))]
    with pytest.raises(ValueError):
        Lion(p, lr=0.0)
    with pytest.raises(ValueError):
        Lion(p, betas=(1.1, 0.9))
    with pytest.raises(ValueError):
        Lion(p, vote="median")
    o = Lion(p)
    assert o.defaults == {"lr": 1e-4, "betas": (0.9, 0.99), "weight_decay": 0.0} or \
        {k: o.defaults[k] for k in ("lr", "betas", "weight_decay")} == {"lr": 1e-4, "betas": (0.9, 0.99), "weight_decay": 0.0}
    assert o.backend_name == "local"


def test_module_level_helpers_match_reference(ref_mod):
    import distributed_lion as ours
    t = torch.arange(13).reshape(13) > 5
    a, sa = ours.flatten_and_pad(t)
    b, sb = ref_mod.flatten_and_pad(t)
    assert torch.equal(a, b) and sa == sb and a.numel() == 16
    assert torch.equal(ours.restore_flattened_tensor(a, sa), ref_mod.restore_flattened_tensor(b, sb))
    g = torch.Generator().manual_seed(0)
    for W in (1, 2, 3, 4, 7, 8):
        votes = [torch.rand(40, generator=g) > 0.5 for _ in range(W)]
        assert torch.equal(ours.majority_vote(votes), ref_mod.majority_vote(votes))
    assert ours.exists(0) and not ours.exists(None)
    p1, p2 = torch.randn(50), None
    p2 = p1.clone()
    gr, m1 = torch.randn(50), torch.randn(50) * 0.1
    m2 = m1.clone()
    ref_mod.update_fn(p1, gr, m1, 1e-3, 0.1, 0.9, 0.99)
    ours.update_fn(p2, gr, m2, 1e-3, 0.1, 0.9, 0.99)
    assert torch.equal(p1, p2) and torch.equal(m1, m2)


def test_mixed_dtype_groups_make_separate_buckets_and_sparse_grads_are_rejected():
    torch.manual_seed(0)
    a = torch.nn.Parameter(torch.randn(40))
    b = torch.nn.Parameter(torch.randn(24).bfloat16())
    c = torch.nn.Parameter(torch.randn(8))
    opt = Lion([{"params": [a, b]}, {"params": [c], "lr": 1e-2}], lr=1e-3, weight_decay=0.1)
    for p in (a, b, c):
        p.grad = torch.randn_like(p)
    before = [p.detach().clone() for p in (a, b, c)]
    opt.step()
    assert len(opt.buckets) == 3                                     # (group 0, fp32), (group 0, bf16), (group 1, fp32)
    assert {bk.dtype for bk in opt.buckets} == {torch.float32, torch.bfloat16}
    for p, q in zip((a, b, c), before):
        assert not torch.equal(p.detach(), q)
    assert abs(float((c.detach() - before[2] * (1 - 1e-2 * 0.1)).abs().max()) - 1e-2) < 1e-6      # group lr honoured
    emb = torch.nn.Embedding(10, 4, sparse=True)
    o2 = Lion(emb.parameters(), lr=1e-3)
    emb(torch.tensor([1, 2])).sum().backward()
    with pytest.raises(RuntimeError, match="sparse"):
        o2.step()


def test_step_counter_is_settable_and_state_dict_keeps_reference_schema():
    m = _model(5)
    o = Lion(m.parameters(), lr=1e-3, max_grad_norm=1.0)
    _run(o, m, steps=3)
    assert o.steps_done == 3
    assert set(o.state_dict().keys()) == {"state", "param_groups"}
    o.steps_done = 10
    _run(o, m, steps=1)
    assert o.steps_done == 11
    with pytest.raises(ValueError):
        o.steps_done = -1


def test_add_param_group_after_first_step_is_picked_up(ref_mod):
    torch.manual_seed(0)
    a_ref, b_ref = torch.nn.Linear(6, 4), torch.nn.Linear(4, 2)
    a_new, b_new = copy.deepcopy(a_ref), copy.deepcopy(b_ref)

    def run(opt, a, b, late):
        torch.manual_seed(3)
        for s in range(4):
            if s == 2:
                opt.add_param_group({"params": list(b.parameters()), "lr": 5e-3, "weight_decay": 0.2})
            x = torch.randn(3, 6)
            opt.zero_grad()
            b(a(x)).pow(2).mean().backward()
            opt.step()

    o_ref = ref_mod.Lion(a_ref.parameters(), lr=1e-3)
    o_new = Lion(a_new.parameters(), lr=1e-3)
    run(o_ref, a_ref, b_ref, True)
    run(o_new, a_new, b_new, True)
    for x, y in zip(list(a_ref.parameters()) + list(b_ref.parameters()),
                    list(a_new.parameters()) + list(b_new.parameters())):
        assert torch.equal(x, y)
    assert len(o_new.buckets) == 2


class _Intermittent(torch.nn.Module):
    """A parameter (``b``) that is part of the graph only on even calls — conditionally-used branches,
    alternating heads and routed experts look like this to the optimizer."""

    def __init__(self):
        super().__init__()
        torch.manual_seed(7)
        self.a = torch.nn.Linear(6, 4)
        self.b = torch.nn.Linear(4, 4)
        self.calls = 0

    def forward(self, x):
        h = self.a(x)
        if self.calls % 2 == 0:
            h = self.b(h)
        self.calls += 1
        return h


def test_parameter_without_gradient_is_left_untouched_like_the_reference(ref_mod):
    """reference: ``if p.grad is not None`` (distributed_lion.py:180) — a parameter that received no gradient in
    a step gets no weight decay, no momentum decay and no sign(beta1*m) move.  With flat buckets ``p.grad`` aliases
    the bucket from the second step on and is never None again; the per-step "written" marks must keep the
    reference's behaviour."""
    m_ref, m_new = _Intermittent(), _Intermittent()
    o_ref = ref_mod.Lion(m_ref.parameters(), lr=1e-2, weight_decay=0.1)
    o_new = Lion(m_new.parameters(), lr=1e-2, weight_decay=0.1)
    torch.manual_seed(3)
    xs = [torch.randn(5, 6) for _ in range(6)]
    for x in xs:
        for model, opt in ((m_ref, o_ref), (m_new, o_new)):
            opt.zero_grad()
            model(x).pow(2).mean().backward()
            opt.step()
    for (n, a), b in zip(m_ref.named_parameters(), m_new.parameters()):
        assert torch.equal(a, b), n
    for a, b in zip(m_ref.parameters(), m_new.parameters()):
        assert torch.equal(o_ref.state[a]["exp_avg"], o_new.state[b]["exp_avg"])


def test_clip_grad_norm_eager_path_matches_torch():
    """On CPU (no kernels) ``Lion.clip_grad_norm_`` scales the gradients in place like torch's utility."""
    m1, m2 = _model(4), _model(4)
    o1 = Lion(m1.parameters(), lr=1e-3)
    o2 = Lion(m2.parameters(), lr=1e-3)
    torch.manual_seed(0)
    x = torch.randn(5, 13)
    for step in range(3):
        for m, o in ((m1, o1), (m2, o2)):
            o.zero_grad()
            (m(x).pow(2).mean() * 100).backward()
        n1 = o1.clip_grad_norm_(0.5)
        n2 = torch.nn.utils.clip_grad_norm_(m2.parameters(), 0.5)
        assert torch.allclose(n1, n2)
        o1.step()
        o2.step()
    for a, b in zip(m1.parameters(), m2.parameters()):
        assert torch.equal(a, b)


def test_gradients_written_by_hand_into_the_aliased_views_are_used():
    """From the second step on ``p.grad`` aliases the flat bucket.  A caller who fills it without a backward pass
    (``p.grad.copy_(...)``) leaves no "written" mark; the parameter must still be updated (it is only skipped when its
    gradient also still holds the zeros of ``zero_grad()``)."""
    torch.manual_seed(0)
    p = torch.nn.Parameter(torch.randn(40))
    q = torch.nn.Parameter(torch.randn(24))
    opt = Lion([p, q], lr=1e-2, weight_decay=0.0)
    (p.sum() + q.sum()).backward()                    # step 1: ordinary backward, buckets get built
    opt.step()
    opt.zero_grad()
    before_p, before_q = p.detach().clone(), q.detach().clone()
    p.grad.copy_(torch.ones(40))                      # step 2: by hand, q gets nothing
    opt.step()
    assert not torch.equal(p.detach(), before_p)      # p moved
    assert torch.equal(q.detach(), before_q)          # q (no gradient this step) is left untouched, as upstream

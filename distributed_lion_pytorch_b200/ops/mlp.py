"""GPT-2 MLP (Conv1D → tanh-GELU → Conv1D) as ONE autograd node, so that the backward can run the down-projection's
data-gradient GEMM, the GELU backward and the up-projection's bias gradient as a single tcgen05 kernel
(``csrc/gemm_dgelu_sm100.cu``): the [tokens, 4·hidden] product is multiplied by gelu'(pre) on its way out of TMEM and
reaches HBM once, already as d(pre), with its column sums reduced on the side.

forward   act, pre = GELU_AUX_BIAS GEMM (cuBLASLt epilogue, ``ops/lt_gemm.py``);  y = act @ W_proj + b_proj
backward  dW_proj += act^T dy (cuBLAS, accumulated in the epilogue); db_proj from the producer hand-off or one pass;
          d(pre), db_fc = dgrad_dgelu(dy, W_proj, pre)            <- the hand-written kernel
          dx = d(pre) @ W_fc^T;  dW_fc += x^T d(pre)

reference: HF ``GPT2MLP`` forward/backward under autograd (modeling_gpt2 via run_clm.py:425-433 and
accelerator.backward, async_trainer.py:33): 2 GEMMs + ~8 elementwise kernels forward, 4 GEMMs + ~10 elementwise /
reduction kernels backward.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch

from . import _cabi, lt_gemm, mark_grad_written, require_library

_ENABLED = os.environ.get("DLION_FUSED_MLP", "1") == "1"
_EMULATE = os.environ.get("DLION_EMULATE_KERNELS") == "1"   # CPU stand-ins, to test the autograd plumbing
_STATUS: Dict[int, torch.Tensor] = {}


def _status(dev: torch.device) -> torch.Tensor:
    st = _STATUS.get(dev.index)
    if st is None:
        st = _STATUS[dev.index] = torch.zeros(1, dtype=torch.int32, device=dev)
    return st


def dgrad_dgelu(dy: torch.Tensor, w: torch.Tensor, pre: torch.Tensor, bias_grad_out: Optional[torch.Tensor] = None,
                accumulate: bool = False, want_bias_grad: bool = True):
    """``dpre = (dy @ w.T) * gelu_tanh'(pre)`` with ``w`` [n, k]; returns ``(dpre, db)`` where ``db`` is the column sum
    of ``dpre`` in ``pre.dtype`` (written / accumulated into ``bias_grad_out`` when given; ``None`` if not wanted)."""
    if not dy.is_cuda:                       # torch stand-in with the kernel's contract (CPU tests of the plumbing)
        x = pre.float()
        kk, cc = 0.7978845608028654, 0.044715
        sg = 0.5 * (1.0 + torch.tanh(kk * (x + cc * x * x * x)))
        dpre = ((dy.float() @ w.float().t()) * (sg + sg * (1.0 - sg) * x * (2 * kk + 6 * kk * cc * x * x))).to(pre.dtype)
        if not want_bias_grad:
            return dpre, None
        sums = dpre.float().sum(0)
        if bias_grad_out is None:
            return dpre, sums.to(pre.dtype)
        bias_grad_out.copy_((sums + bias_grad_out.float()) if accumulate else sums)
        return dpre, bias_grad_out
    assert dy.dtype == w.dtype == pre.dtype and dy.dtype in (torch.bfloat16, torch.float16)
    m, k = dy.shape
    n = w.shape[0]
    assert w.shape[1] == k and pre.shape == (m, n)
    dev = dy.device
    L = _cabi.lib()
    with torch.cuda.device(dev):
        dpre = torch.empty_like(pre)
        m_tiles = int(L.dlion_dgrad_dgelu_parts(m))
        part = torch.empty(m_tiles, n, dtype=torch.float32, device=dev)
        st = _cabi.stream_ptr(dev)
        dt = _cabi.dtype_code(pre.dtype)
        _cabi.check(L.dlion_dgrad_dgelu(dy.data_ptr(), dy.stride(0), w.data_ptr(), w.stride(0), pre.data_ptr(), pre.stride(0),
                                        dpre.data_ptr(), dpre.stride(0), part.data_ptr(), _status(dev).data_ptr(), m, n, k, dt,
                                        st), "dgrad_dgelu")
        db = None
        if want_bias_grad:
            db = bias_grad_out if bias_grad_out is not None else torch.empty(n, dtype=pre.dtype, device=dev)
            _cabi.check(L.dlion_dgrad_dgelu_bias_finalize(part.data_ptr(), db.data_ptr(), m_tiles, n, int(accumulate), dt, st),
                        "dgrad_dgelu_bias_finalize")
    return dpre, db


def _ok16(t: torch.Tensor) -> bool:
    return t.dim() == 2 and t.stride(1) == 1 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0


class _GeluMLP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w_fc, b_fc, w_proj, b_proj):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        inner = w_fc.shape[1]
        pre = torch.empty(x2.shape[0], inner, dtype=x.dtype, device=x.device)
        act = torch.empty_like(pre)
        if not (x.is_cuda and lt_gemm.gemm(x2, w_fc, act, epilogue=lt_gemm.GELU_AUX_BIAS, bias=b_fc, aux=pre)):
            from .activation import gelu_tanh
            torch.addmm(b_fc, x2, w_fc, out=pre)
            act = gelu_tanh(pre)
        y = x.new_empty(*x.shape[:-1], w_proj.shape[1])
        torch.addmm(b_proj, act, w_proj, out=y.view(-1, w_proj.shape[1]))
        ctx.save_for_backward(x2, pre, act, w_fc, b_fc, w_proj, b_proj)
        ctx.x_shape = x.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        from .linear import _can_accumulate, bias_grad
        x2, pre, act, w_fc, b_fc, w_proj, b_proj = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw_fc = db_fc = dw_proj = db_proj = None
        # ---- down-projection: weight and bias gradients
        if ctx.needs_input_grad[3]:
            if _can_accumulate(w_proj, dy2):
                mark_grad_written(w_proj)
                w_proj.grad.addmm_(act.t(), dy2)
            else:
                dw_proj = act.t() @ dy2
        if ctx.needs_input_grad[4]:
            db_proj = bias_grad(dy2, b_proj, dy)
        # ---- d(pre) and the up-projection's bias gradient out of one tensor-core kernel
        want_db = ctx.needs_input_grad[2]
        acc = want_db and _can_accumulate(b_fc, pre) and b_fc.grad.shape == b_fc.shape and b_fc.grad.dtype == pre.dtype
        if acc:
            mark_grad_written(b_fc)
        dpre, db = dgrad_dgelu(dy2, w_proj, pre, bias_grad_out=b_fc.grad if acc else None, accumulate=acc,
                               want_bias_grad=want_db)
        db_fc = None if acc else db
        # ---- up-projection
        if ctx.needs_input_grad[0]:
            dx = (dpre @ w_fc.t()).view(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            if _can_accumulate(w_fc, dpre):
                mark_grad_written(w_fc)
                w_fc.grad.addmm_(x2.t(), dpre)
            else:
                dw_fc = x2.t() @ dpre
        return dx, dw_fc, db_fc, dw_proj, db_proj


def gelu_mlp(x: torch.Tensor, w_fc: torch.Tensor, b_fc: torch.Tensor, w_proj: torch.Tensor,
             b_proj: torch.Tensor) -> Optional[torch.Tensor]:
    """``gelu_tanh(x @ w_fc + b_fc) @ w_proj + b_proj`` (HF Conv1D weights, [in, out]) as one autograd node, or ``None``
    when the fused backward does not apply (CPU, fp32, autograd off, odd shapes, library missing, switched off)."""
    if not (_ENABLED and torch.is_grad_enabled()):
        return None
    if x.is_cuda:
        if not (require_library() and lt_gemm.available()) or x.dtype not in (torch.bfloat16, torch.float16):
            return None
    elif not _EMULATE:
        return None
    if b_fc is None or b_proj is None:
        return None
    if not (x.dtype == w_fc.dtype == b_fc.dtype == w_proj.dtype == b_proj.dtype):
        return None
    inner, out = w_proj.shape
    if w_fc.shape[1] != inner or inner % 8 or out % 8 or w_fc.shape[0] % 8 or not (_ok16(w_fc) and _ok16(w_proj)) \
            or not w_fc.is_contiguous() or (x.is_cuda and b_fc.data_ptr() % 16):
        return None
    if not (x.requires_grad or w_fc.requires_grad or w_proj.requires_grad):
        return None
    return _GeluMLP.apply(x, w_fc, b_fc, w_proj, b_proj)

"""cuBLASLt GEMMs with training epilogues (``csrc/lt_gemm.cu``): GELU+aux+bias forward, dGELU+bias-grad
dgrad, weight-gradient GEMM that also emits the bias gradient — the fusions PyTorch does not expose.

reference: HF GPT-2 ``MLP`` (Conv1D → NewGELU → Conv1D) and ``Conv1D`` backward under autograd, reached
from run_clm.py:588-639: two GEMMs plus ~8 elementwise / reduction kernels per direction upstream.

Every function returns ``False`` when the library has no kernel for the request (or cuBLASLt is not
loadable); callers keep their unfused path for that case, so nothing here is required for correctness.
"""
from __future__ import annotations

import ctypes as C
import logging
import os
from typing import Dict, Optional

import torch

from . import _cabi, require_library
from ._cabi import LtArgs

NONE, BIAS, GELU_AUX_BIAS, DGELU_BGRAD, BGRAD_B, DGELU = 0, 1, 2, 3, 4, 5
_WORKSPACE_BYTES = 32 << 20
_ENABLED = os.environ.get("DLION_LT_EPILOGUES", "1") == "1"
_workspaces: Dict[torch.device, torch.Tensor] = {}
_unsupported = set()          # problem signatures the library refused once: do not ask again


def available(device: Optional[torch.device] = None) -> bool:
    if not _ENABLED or not torch.cuda.is_available() or not require_library():
        return False
    return bool(_cabi.lib().dlion_lt_available())


def _workspace(device: torch.device) -> torch.Tensor:
    ws = _workspaces.get(device)
    if ws is None:
        ws = _workspaces[device] = torch.empty(_WORKSPACE_BYTES, dtype=torch.uint8, device=device)
    return ws


def _rowmajor(t: torch.Tensor) -> bool:
    return t.dim() == 2 and t.stride(1) == 1 and t.stride(0) >= t.shape[1]


def gemm(a: torch.Tensor, b: torch.Tensor, d: torch.Tensor, *, trans_a: bool = False, trans_b: bool = False,
         alpha: float = 1.0, beta: float = 0.0, epilogue: int = NONE, bias: Optional[torch.Tensor] = None,
         bias_grad: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None, algo_rank: int = 0) -> bool:
    """``d[m,n] = alpha * op(a) @ op(b) + beta * d`` with the epilogue described in ``include/dlion.h``.
    ``a`` is stored [m,k] ([k,m] with ``trans_a``), ``b`` [k,n] ([n,k] with ``trans_b``); all row-major."""
    if not (_rowmajor(a) and _rowmajor(b) and _rowmajor(d)) or not (a.dtype == b.dtype == d.dtype):
        return False
    m, n = d.shape
    k = a.shape[0] if trans_a else a.shape[1]
    if (a.shape[1] if trans_a else a.shape[0]) != m or (b.shape[0] if trans_b else b.shape[1]) != n \
            or (b.shape[1] if trans_b else b.shape[0]) != k:
        raise ValueError(f"lt_gemm shape mismatch: a{tuple(a.shape)} b{tuple(b.shape)} d{tuple(d.shape)} "
                         f"trans_a={trans_a} trans_b={trans_b}")
    sig = (m, n, k, trans_a, trans_b, epilogue, a.dtype, beta != 0.0, algo_rank)
    if sig in _unsupported:
        return False
    if aux is not None and not (_rowmajor(aux) and aux.shape == d.shape and aux.dtype == d.dtype):
        return False
    args = LtArgs(a.data_ptr(), b.data_ptr(), d.data_ptr(), m, n, k, a.stride(0), b.stride(0), d.stride(0),
                  int(trans_a), int(trans_b), alpha, beta, epilogue, _cabi.ptr(bias), _cabi.ptr(bias_grad), _cabi.ptr(aux),
                  aux.stride(0) if aux is not None else 0, _workspace(d.device).data_ptr(), _WORKSPACE_BYTES,
                  _cabi.dtype_code(d.dtype), algo_rank)
    with torch.cuda.device(d.device):
        rc = _cabi.lib().dlion_lt_gemm(C.byref(args), _cabi.stream_ptr(d.device))
    if rc == -2:                # DLION_E_UNSUPPORTED: remember, say so once, and let the caller fall back
        _unsupported.add(sig)
        logging.getLogger("dlion.lt_gemm").warning(
            "cuBLASLt declined %s (%s); using the unfused path for this shape",
            sig, _cabi.lib().dlion_last_error().decode(errors="replace"))
        return False
    if rc != 0:                 # not _cabi.check: these are the library's kernels, not counted as ours
        raise RuntimeError(f"libdlion lt_gemm failed (code {rc}): "
                           f"{_cabi.lib().dlion_last_error().decode(errors='replace')}")
    return True


# ---------------------------------------------------------------------- linear + GELU in one GEMM
class _LinearGelu(torch.autograd.Function):
    """``gelu_tanh(x @ w + b)`` with an HF Conv1D weight ([in, out]): the GEMM's GELU_AUX_BIAS epilogue writes
    the activation and its pre-image in one pass (90 µs instead of 126 µs for GEMM + GELU kernel at the GPT-2
    shape).  The backward keeps this package's own kernels: cuBLASLt's DGELU / BGRAD epilogues were measured
    and lose on this library version (DGELU_BGRAD 314 µs vs 175 µs for GEMM + gelu_bwd + column-sum; BGRAD on
    the weight-gradient GEMM 85 µs vs 79 µs) — see DESIGN §8."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        pre = torch.empty(x2.shape[0], weight.shape[1], dtype=x.dtype, device=x.device)
        act = x.new_empty(*x.shape[:-1], weight.shape[1])
        if not gemm(x2, weight, act.view(-1, weight.shape[1]), epilogue=GELU_AUX_BIAS, bias=bias, aux=pre):
            from .activation import gelu_tanh
            torch.addmm(bias, x2, weight, out=pre)
            act = gelu_tanh(pre).view(*x.shape[:-1], weight.shape[1])
        ctx.save_for_backward(x2, pre, weight, bias)
        ctx.x_shape = x.shape
        return act

    @staticmethod
    def backward(ctx, dact):
        from .linear import _can_accumulate, bias_grad
        x2, pre, weight, bias = ctx.saved_tensors
        d2 = dact.reshape(-1, dact.shape[-1])
        if not d2.is_contiguous():
            d2 = d2.contiguous()
        dpre = torch.empty_like(pre)
        dx = dw = db = None
        want_db = ctx.needs_input_grad[2]
        from . import linear as _lin
        rows, cols = pre.shape
        fuse_db = want_db and _lin._FUSE_COLSUM and cols % 8 == 0 \
            and int(_cabi.lib().dlion_colsum_tiles(cols)) <= _lin._MAX_TILES and bias.dtype == pre.dtype
        with torch.cuda.device(pre.device):
            if fuse_db:
                # d(pre) and the c_fc bias gradient from ONE pass over dY (the column sums are reduced while the
                # activation backward has the values in registers)
                L = _cabi.lib()
                acc = _can_accumulate(bias, dpre) and bias.grad.shape == (cols,)
                out = bias.grad if acc else torch.empty(cols, dtype=bias.dtype, device=pre.device)
                partial = torch.empty(int(L.dlion_colsum_parts(rows, cols)), cols, dtype=torch.float32, device=pre.device)
                if acc:
                    from . import mark_grad_written
                    mark_grad_written(bias)
                _cabi.check(L.dlion_gelu_tanh_bwd_colsum(pre.data_ptr(), d2.data_ptr(), dpre.data_ptr(), rows, cols,
                                                         partial.data_ptr(), _lin._tickets(pre.device).data_ptr(),
                                                         out.data_ptr(), int(acc), _cabi.dtype_code(pre.dtype),
                                                         _cabi.stream_ptr(pre.device)), "gelu_bwd_colsum")
                db = None if acc else out
                want_db = False
            else:
                _cabi.check(_cabi.lib().dlion_gelu_tanh_bwd(pre.data_ptr(), d2.data_ptr(), dpre.data_ptr(), pre.numel(),
                                                            _cabi.dtype_code(pre.dtype), _cabi.stream_ptr(pre.device)),
                            "gelu_bwd")
        if ctx.needs_input_grad[0]:
            dx = (dpre @ weight.t()).view(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            if _can_accumulate(weight, dpre):
                from . import mark_grad_written
                mark_grad_written(weight)
                weight.grad.addmm_(x2.t(), dpre)
            else:
                dw = x2.t() @ dpre
        if want_db:
            db = bias_grad(dpre, bias)
        return dx, dw, db


def linear_gelu(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor) -> Optional[torch.Tensor]:
    """Fused ``gelu_tanh(x @ weight + bias)`` (``weight`` [in, out]) or ``None`` when it does not apply (CPU, no
    cuBLASLt, fp32, shapes that are not multiples of 8, autograd off) — the caller then runs linear + GELU."""
    if bias is None or not (x.is_cuda and torch.is_grad_enabled() and available()):
        return None
    if not (x.dtype == weight.dtype == bias.dtype) or x.dtype not in (torch.bfloat16, torch.float16):
        return None
    if weight.shape[1] % 8 or weight.shape[0] % 8 or not weight.is_contiguous() or weight.data_ptr() % 16 \
            or bias.data_ptr() % 16:
        return None
    return _LinearGelu.apply(x, weight, bias)

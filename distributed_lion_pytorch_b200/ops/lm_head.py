"""N5, no-grad side — LM-head projection fused with log-sum-exp on the tcgen05 tensor cores.

``lm_head_log_probs(h, weight, labels)`` returns ``(logp, lse)`` — per-token ``log softmax(h @ weight.T)[label]`` and
``logsumexp(h @ weight.T)`` as fp32 ``[N]`` — from ONE kernel (``csrc/lmhead_lse_sm100.cu``: TMA tile loads,
``tcgen05.mma`` into TMEM, online log-sum-exp epilogue out of TMEM) plus a tiny merge; no ``[N, vocab]`` logits block is
written, the head is not padded or copied.  It serves every place that needs log-probabilities but no gradient:
evaluation loss (``Trainer.evaluate`` → ``linear_cross_entropy`` under ``no_grad``), the frozen reference model of DPO
and the forward half of ``linear_log_probs`` (whose backward recomputes blocks anyway).

reference: HF ``ForCausalLM`` evaluation loss (run_clm.py:626-639); TRL ``DPOTrainer`` reference log-probs
(dpo_llama2.py:219-231) — upstream both materialise fp32 ``[batch, seq, vocab]`` logits.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Tuple

import torch

from . import _cabi, require_library

_STATUS: Dict[int, torch.Tensor] = {}
_EMULATE = os.environ.get("DLION_EMULATE_KERNELS") == "1"   # CPU stand-in, to test the dispatch around the kernel


def enabled() -> bool:
    return os.environ.get("DLION_LMHEAD_FUSED", "1") != "0"


def supported(h: torch.Tensor, weight: torch.Tensor, dtype: Optional[torch.dtype] = None) -> bool:
    """16-bit CUDA operands, hidden a multiple of 8, library present and the path not switched off."""
    dt = dtype or h.dtype
    if _EMULATE and not h.is_cuda:
        return enabled() and h.dim() == 2 and weight.dim() == 2 and h.shape[1] == weight.shape[1] and h.shape[0] > 0
    return (h.is_cuda and weight.is_cuda and enabled() and dt in (torch.bfloat16, torch.float16)
            and h.dim() == 2 and weight.dim() == 2 and h.shape[1] == weight.shape[1] and h.shape[1] % 8 == 0
            and h.shape[0] > 0 and require_library())


def plan(rows: int, vocab: int) -> Tuple[int, int]:
    """(vocabulary splits, workspace floats) the kernel will use for this problem size."""
    s, ws = C.c_int(0), C.c_int64(0)
    _cabi.check(_cabi.lib().dlion_lmhead_plan(rows, vocab, C.byref(s), C.byref(ws)), "lmhead_plan")
    return s.value, ws.value


def _rowmajor16(t: torch.Tensor) -> torch.Tensor:
    if t.stride(1) == 1 and t.stride(0) % 8 == 0 and t.stride(0) >= t.shape[1] and t.data_ptr() % 16 == 0:
        return t
    return t.contiguous()


def lm_head_log_probs(h: torch.Tensor, weight: torch.Tensor, labels: Optional[torch.Tensor],
                      want_lse: bool = True) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """h [N, H], weight [V, H] (same 16-bit dtype, CUDA), labels [N] int64 or None → (logp | None, lse | None)."""
    if not h.is_cuda:                        # torch stand-in with the kernel's contract: fp32 logits, never stored
        logits = h.detach().float() @ weight.detach().float().t()
        lse = torch.logsumexp(logits, dim=-1)
        if labels is None:
            return None, lse
        valid = (labels >= 0) & (labels < weight.shape[0])
        picked = logits.gather(1, labels.clamp(0, weight.shape[0] - 1)[:, None]).squeeze(1)
        return torch.where(valid, picked - lse, torch.zeros_like(lse)), (lse if want_lse else None)
    assert h.dtype == weight.dtype and h.dtype in (torch.bfloat16, torch.float16), (h.dtype, weight.dtype)
    h, weight = _rowmajor16(h.detach()), _rowmajor16(weight.detach())
    N, H = h.shape
    V = weight.shape[0]
    dev = h.device
    _, ws_floats = plan(N, V)
    with torch.cuda.device(dev):
        ws = torch.empty(ws_floats, dtype=torch.float32, device=dev)
        logp = torch.empty(N, dtype=torch.float32, device=dev) if labels is not None else None
        lse = torch.empty(N, dtype=torch.float32, device=dev) if (want_lse or labels is None) else None
        st = _STATUS.get(dev.index)
        if st is None:
            st = _STATUS[dev.index] = torch.zeros(1, dtype=torch.int32, device=dev)
        lab = labels.contiguous() if labels is not None else None
        _cabi.check(_cabi.lib().dlion_lmhead_logprobs(
            h.data_ptr(), h.stride(0), weight.data_ptr(), weight.stride(0), lab.data_ptr() if lab is not None else None,
            logp.data_ptr() if logp is not None else None, lse.data_ptr() if lse is not None else None, ws.data_ptr(),
            ws_floats, st.data_ptr(), N, V, H, _cabi.dtype_code(h.dtype), _cabi.stream_ptr(dev)), "lmhead_logprobs")
    return logp, lse

"""ctypes binding of ``_lib/libdlion.so`` (declarations: ``include/dlion.h``).

The library is plain C ABI + raw device pointers, so a call costs one ctypes dispatch
(~2 µs) and is legal inside CUDA-graph capture.  ``lib()`` raises if the shared object is
missing — on a GPU box a silent PyTorch fallback would hide a broken build.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

import torch

_LIB: Optional[C.CDLL] = None
_LIB_PATH = Path(__file__).resolve().parent.parent / "_lib" / "libdlion.so"

DT_F32, DT_BF16, DT_F16 = 0, 1, 2
_DTYPES = {torch.float32: DT_F32, torch.bfloat16: DT_BF16, torch.float16: DT_F16}
VOTE_CODES = {("majority", "neg"): 0, ("majority", "zero"): 1, ("average", "neg"): 2, ("average", "zero"): 2}
MAX_RANKS = 16
FUSED_MAX_CTAS = 1024


class FusedArgs(C.Structure):
    _fields_ = [
        ("p", C.c_void_p), ("g", C.c_void_p), ("m", C.c_void_p), ("n", C.c_int64),
        ("elem_base", C.c_int64), ("dtype", C.c_int), ("world", C.c_int), ("rank", C.c_int),
        ("live_mask", C.c_uint32), ("wait_mask", C.c_uint32),
        ("lr", C.c_float), ("decay", C.c_float), ("beta1", C.c_float), ("omb1", C.c_float),
        ("beta2", C.c_float), ("omb2", C.c_float),
        ("vote", C.c_int), ("stoch_r", C.c_float), ("seed", C.c_uint64), ("step", C.c_uint32),
        ("epoch", C.c_uint64),
        ("arena", C.c_void_p * MAX_RANKS), ("flags", C.c_void_p * MAX_RANKS), ("mc_arena", C.c_void_p),
        ("arena_words", C.c_int64), ("grid", C.c_int), ("emulate_ranks", C.c_int),
        ("emu_stride", C.c_int64), ("status", C.c_void_p), ("spin_limit", C.c_int64), ("sr_key", C.c_uint64),
        ("clip_sumsq", C.c_void_p), ("clip_max_norm", C.c_float), ("zero_g", C.c_int),
        ("wait_timeout_us", C.c_uint64),
    ]


class LtArgs(C.Structure):
    """``dlion_lt_args`` (include/dlion.h): one cuBLASLt GEMM with a training epilogue."""
    _fields_ = [("a", C.c_void_p), ("b", C.c_void_p), ("d", C.c_void_p),
                ("m", C.c_int64), ("n", C.c_int64), ("k", C.c_int64),
                ("lda", C.c_int64), ("ldb", C.c_int64), ("ldd", C.c_int64),
                ("trans_a", C.c_int), ("trans_b", C.c_int), ("alpha", C.c_float), ("beta", C.c_float),
                ("epilogue", C.c_int), ("bias", C.c_void_p), ("bias_grad", C.c_void_p), ("aux", C.c_void_p),
                ("ld_aux", C.c_int64), ("workspace", C.c_void_p), ("workspace_bytes", C.c_int64), ("dtype", C.c_int),
                ("algo_rank", C.c_int)]


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DTYPES[dt]
    except KeyError:
        raise TypeError(f"libdlion supports float32/bfloat16/float16, got {dt}") from None


def lib_path() -> Path:
    return Path(os.environ.get("DLION_LIB", _LIB_PATH))


def available() -> bool:
    return lib_path().exists()


def _declare(L: C.CDLL) -> None:
    vp, i64, i32, u32, u64, f32 = C.c_void_p, C.c_int64, C.c_int, C.c_uint32, C.c_uint64, C.c_float
    L.dlion_last_error.restype = C.c_char_p
    L.dlion_version.restype = i32
    L.dlion_sumsq_scratch_bytes.restype = i64
    L.dlion_norm_bwd_partial_rows.restype = i32
    sigs = {
        "dlion_device_sm_count": [C.POINTER(i32)],
        "dlion_lion_local": [vp, vp, vp, i64, i32, f32, f32, f32, f32, f32, f32, vp],
        "dlion_lion_local_sr": [vp, vp, vp, i64, i32, f32, f32, f32, f32, f32, f32, u64, i64, vp],
        "dlion_lion_local_clip": [vp, vp, vp, i64, i32, f32, f32, f32, f32, f32, f32, u64, i64, vp, f32, i32, vp],
        "dlion_sign_pack_clip": [vp, vp, vp, i64, i32, f32, f32, f32, u64, u32, u32, vp, f32, vp],
        "dlion_vote_update_clip": [vp, vp, vp, vp, i64, i32, u32, i64, i32, f32, f32, f32, f32, i32, u64, vp, f32, i32, vp],
        "dlion_sign_pack": [vp, vp, vp, i64, i32, f32, f32, f32, u64, u32, u32, vp],
        "dlion_vote_update": [vp, vp, vp, vp, i64, i32, u32, i64, i32, f32, f32, f32, f32, i32, vp],
        "dlion_vote_update_sr": [vp, vp, vp, vp, i64, i32, u32, i64, i32, f32, f32, f32, f32, i32, u64, vp],
        "dlion_fused_step": [C.POINTER(FusedArgs), vp],
        "dlion_fused_tile": [i32],
        "dlion_fused_max_grid": [i32, C.POINTER(i32)],
        "dlion_multi_gather": [C.POINTER(vp), C.POINTER(vp), C.POINTER(i64), i32, i32, i32, vp],
        "dlion_sumsq": [vp, i64, i32, vp, vp, vp],
        "dlion_clip_scale": [vp, i64, i32, vp, f32, vp],
        "dlion_norm_fwd": [vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, i32, f32, u64, i32, vp],
        "dlion_norm_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, f32, u64, i32, vp],
        "dlion_norm_bwd_parts": [i64, i32],
        "dlion_norm_bwd_finalize": [vp, vp, vp, vp, i32, i32, i32, vp],
        "dlion_colsum_parts": [i64, i32],
        "dlion_colsum": [vp, i64, i32, vp, i32, vp],
        "dlion_colsum_tiles": [i32],
        "dlion_colsum_fused": [vp, i64, i32, vp, vp, vp, i32, i32, vp],
        "dlion_norm_bwd_finalize_acc": [vp, vp, vp, vp, i32, i32, i32, i32, vp],
        "dlion_gelu_tanh_bwd_colsum": [vp, vp, vp, i64, i32, vp, vp, vp, i32, i32, vp],
        "dlion_qkv_pack_colsum": [vp, vp, vp, C.POINTER(i64), vp, i32, i32, i32, i32, vp, vp, vp, i32, i32, vp],
        "dlion_norm_bwd_dxsum_supported": [i32, i32],
        "dlion_norm_bwd_dxsum_parts": [i64, i32],
        "dlion_norm_bwd_dxsum": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, f32, u64, i32, vp],
        "dlion_norm_bwd_finalize_dxsum": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp],
        "dlion_embed_fwd": [vp, vp, vp, vp, i64, i32, i32, i64, i32, vp],
        "dlion_embed_bwd_sorted": [vp, vp, vp, vp, i64, i32, i64, i32, i32, vp],
        "dlion_gelu_tanh_fwd": [vp, vp, i64, i32, vp],
        "dlion_gelu_tanh_bwd": [vp, vp, vp, i64, i32, vp],
        "dlion_swiglu_fwd": [vp, vp, vp, i64, i32, vp],
        "dlion_swiglu_bwd": [vp, vp, vp, vp, vp, i64, i32, vp],
        "dlion_rope": [vp, vp, vp, vp, i64, i32, i32, i32, i32, i64, i64, i32, i32, vp],
        "dlion_ce_fwd_bwd": [vp, vp, vp, i64, i32, i64, i64, f32, i32, i32, vp],
        "dlion_ce_bwd_rowscale": [vp, vp, vp, i64, i32, i64, i64, i32, vp],
        "dlion_lmhead_plan": [i32, i32, C.POINTER(i32), C.POINTER(i64)],
        "dlion_lmhead_logprobs": [vp, i64, vp, i64, vp, vp, vp, vp, i64, vp, i32, i32, i32, i32, vp],
        "dlion_dgrad_dgelu_parts": [i64],
        "dlion_dgrad_dgelu_set_profile": [vp],
        "dlion_dgrad_dgelu": [vp, i64, vp, i64, vp, i64, vp, i64, vp, vp, i32, i32, i32, i32, vp],
        "dlion_dgrad_dgelu_bias_finalize": [vp, vp, i32, i32, i32, i32, vp],
        "dlion_lt_available": [],
        "dlion_lt_gemm": [C.POINTER(LtArgs), vp],
        "dlion_q4_quantize": [vp, vp, vp, i64, i32, C.POINTER(f32), i32, vp],
        "dlion_q4_dequantize": [vp, vp, vp, i64, i32, C.POINTER(f32), i32, vp],
    }
    for name, argtypes in sigs.items():
        fn = getattr(L, name)          # a missing symbol is a broken build: fail loudly
        fn.argtypes = argtypes
        fn.restype = i32


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        path = lib_path()
        if not path.exists():
            raise RuntimeError(
                f"{path} is missing: build it with `python -m distributed_lion_pytorch_b200.build` "
                "(or __graft_entry__.build()).  CUDA tensors never fall back to eager PyTorch silently.")
        L = C.CDLL(str(path))
        _declare(L)
        _LIB = L
    return _LIB


LAUNCHES = 0     # C-ABI calls that enqueued our kernels (bench.py reports it as gpu_launches)


def check(rc: int, what: str = "") -> None:
    global LAUNCHES
    LAUNCHES += 1
    if rc != 0:
        msg = lib().dlion_last_error().decode(errors="replace")
        raise RuntimeError(f"libdlion {what} failed (code {rc}): {msg}")


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()

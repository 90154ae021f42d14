"""ctypes binding of libdptb200.so (the C ABI declared in include/dptb200.h).

The kernels are plain CUDA with ``extern "C"`` launchers, so the binding is pointer + stream
passing; no torch extension is needed on this path.  On a machine with a GPU the library MUST be
present — ops raise instead of silently falling back to PyTorch.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_NATIVE_DIR = Path(__file__).resolve().parent.parent / "_native"
_LIB_PATH = _NATIVE_DIR / "libdptb200.so"
_lib = None

ERRORS = {-1: "unsupported shape/alignment", -2: "driver entry point unavailable",
          -3: "tensor-map encode rejected", -4: "invalid argument", -5: "timeout"}

vp, i32, i64, f32, f64 = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_double

_SIGS = {
    "dpt_conv_fprop": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp, i32, i32, vp],
    "dpt_conv_dgrad": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp, i32, i32, i32, vp],
    "dpt_conv_dgrad_ex": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp, i32, i32, i32,
                          i32, i32, vp],
    "dpt_bn_fwd_train_small_f32": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, i32, vp],
    "dpt_bn_bwd_reduce_ex": [vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, vp],
    "dpt_conv_wgrad": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "dpt_gemm_bf16": [i32, vp, vp, vp, i32, i32, i32, i64, i64, i64, i32, vp, vp, vp, i32, vp],
    "dpt_bn_stats": [vp, vp, vp, i64, i32, vp],
    "dpt_bn_fwd_train": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, i32, vp],
    "dpt_bn_fwd_eval": [vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, i32, vp],
    "dpt_bn_bwd_reduce": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, vp],
    "dpt_bn_bwd_fused": [vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, vp],
    "dpt_bn_bwd_apply": [vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, vp],
    "dpt_maxpool3x3s2_fwd": [vp, vp, vp, i32, i32, i32, i32, vp],
    "dpt_maxpool3x3s2_bwd": [vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "dpt_head_fwd": [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "dpt_head_bwd": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "dpt_cross_entropy": [vp, vp, vp, vp, vp, vp, i32, i32, vp],
    "dpt_im2col_nchw_f32": [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "dpt_pad_rows_f32_to_bf16": [vp, vp, i32, i32, i32, vp],
    "dpt_augment_crop_flip_norm": [vp, vp, vp, i32, i32, i32, i32, vp, vp, vp],
    "dpt_nchw_f32_to_nhwc_bf16": [vp, vp, i32, i32, i32, i32, vp],
    "dpt_amp_unscale": [vp, i64, vp, vp, vp],
    "dpt_amp_update_scale": [vp, vp, vp, vp, f32, f32, i32, vp],
    "dpt_sgd_flat": [vp, vp, vp, vp, i64, f32, f32, f32, f32, i32, vp, vp, vp],
    "dpt_adam_flat": [vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, i32, vp, vp, vp, vp],
    "dpt_cast_f32_to_bf16": [vp, vp, i64, vp],
    "dpt_igemm_set_persist": [i64, i32],
    "dpt_igemm_set_bn256": [i32],
    "dpt_gemm2sm_bf16": [vp, vp, vp, i32, i32, i32, i64, i64, i64, i32, i32, vp],
    "dpt_pdl_skip_next": [],
    "dpt_set_lowp_dtype": [i32],
    "dpt_get_lowp_dtype": [],
    "dpt_symm_alloc": [C.c_size_t, C.POINTER(vp), vp],
    "dpt_symm_open": [vp, C.POINTER(vp)],
    "dpt_symm_close": [vp],
    "dpt_symm_free": [vp],
    "dpt_comm_create": [i32, i32, C.POINTER(vp), f64, C.POINTER(vp)],
    "dpt_comm_destroy": [vp],
    "dpt_comm_epochs": [vp, vp, i32],
    "dpt_allreduce_fused": [vp, C.POINTER(vp), i64, f32, vp, C.POINTER(vp), C.POINTER(vp),
                            C.POINTER(vp), vp, f32, f32, f32, i32, i32, vp],
    "dpt_allreduce_fused_mcpush": [vp, C.POINTER(vp), i64, f32, vp, C.POINTER(vp), C.POINTER(vp),
                                   C.POINTER(vp), vp, vp, vp, f32, f32, f32, i32, i32, vp],
    "dpt_allreduce_oneshot": [vp, C.POINTER(vp), vp, i64, f32, vp, C.POINTER(vp), i32, vp],
    "dpt_allreduce_oneshot_push": [vp, vp, C.POINTER(vp), i64, i64, f32, vp, vp, vp, vp, vp, f32, f32,
                                   f32, i32, i32, vp],
    "dpt_comm_set_trace": [vp, vp, i32],
    "dpt_comm_skip_closing_barrier_next": [],
    "dpt_trace_stamp": [vp, i32, i32, vp],
    "dpt_allreduce_nvls": [vp, vp, i64, f32, vp, C.POINTER(vp), i32, vp],
    "dpt_allreduce_nvls_sgd": [vp, vp, i64, f32, vp, C.POINTER(vp), vp, vp, vp, vp, f32, f32, f32, i32,
                               i32, vp],
    "dpt_broadcast": [vp, C.POINTER(vp), i64, i32, i32, vp],
    "dpt_barrier": [vp, vp],
}


def available() -> bool:
    return _LIB_PATH.exists()


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not _LIB_PATH.exists():
            raise RuntimeError(
                f"{_LIB_PATH} is missing: run `python build_native.py` (or __graft_entry__.build()) "
                "— the CUDA ops never fall back to PyTorch kernels")
        _lib = C.CDLL(str(_LIB_PATH), mode=os.RTLD_GLOBAL)
        for name, sig in _SIGS.items():
            fn = getattr(_lib, name)
            fn.argtypes = sig
            fn.restype = i32
        _lib.dpt_comm_flag_bytes.restype = C.c_size_t
        _lib.dpt_comm_flag_bytes.argtypes = []
        _lib.dpt_comm_error.restype = C.c_uint
        _lib.dpt_comm_error.argtypes = [vp]
    return _lib


def ptr(t) -> int | None:
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def check(code: int, what: str) -> None:
    if code == 0:
        return
    if code < 0:
        raise RuntimeError(f"{what}: {ERRORS.get(code, code)}")
    raise RuntimeError(f"{what}: CUDA error {code}")


def call(name: str, *args) -> None:
    check(getattr(lib(), name)(*args), name)


def set_lowp_dtype(dtype) -> None:
    """16-bit element type used by every kernel launched afterwards (bf16 or fp16)."""
    lib().dpt_set_lowp_dtype({torch.float16: 1, torch.float32: 2}.get(dtype, 0))


def pdl_skip_next() -> None:
    """Call after a cross-stream event wait: the next kernel launch is fully ordered."""
    lib().dpt_pdl_skip_next()


LAUNCHES = 0  # kernels launched through this binding (bench.py reports it)


def count(n: int = 1) -> None:
    global LAUNCHES
    LAUNCHES += n

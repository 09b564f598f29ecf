"""ctypes binding of ``libb200kernels.so`` (C ABI declared in include/b200_kernels.h).

The library is loaded lazily. On a machine with a GPU a missing library is a hard
error (no silent fallback to eager PyTorch); on a CPU-only machine ``available()`` is
False and the pure-PyTorch reference ops are used instead.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_LIB = None
_LIB_ERR = None
LIB_PATH = Path(__file__).resolve().parent / "_lib" / "libb200kernels.so"

vp, i32, i64, f32, u64 = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong

_SIGNATURES = {
    "b200_norm_bwd_grid": ([i32], i32),
    "b200_rmsnorm_fwd": ([vp, vp, vp, vp, vp, vp, i32, i32, f32, vp], i32),
    "b200_rmsnorm_bwd": ([vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp], i32),
    "b200_layernorm_fwd": ([vp, vp, vp, vp, vp, vp, i32, i32, f32, vp], i32),
    "b200_layernorm_bwd": ([vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp], i32),
    "b200_rope": ([vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp], i32),
    "b200_glu_fwd": ([vp, vp, i64, i32, i32, vp], i32),
    "b200_glu_bwd": ([vp, vp, vp, i64, i32, i32, vp], i32),
    "b200_bias_gelu_fwd": ([vp, vp, vp, i64, i32, i32, vp], i32),
    "b200_bias_gelu_bwd": ([vp, vp, vp, vp, i64, i32, i32, vp], i32),
    "b200_dropout_add_fwd": ([vp, vp, vp, vp, i64, i32, f32, u64, u64, vp], i32),
    "b200_dropout_bwd": ([vp, vp, i64, i32, f32, u64, u64, vp], i32),
    "b200_accum_bf16_to_f32": ([vp, vp, i64, f32, i32, vp], i32),
    "b200_cast_f32_to_bf16": ([vp, vp, i64, vp], i32),
    "b200_softmax_fwd": ([vp, vp, vp, i32, i32, i32, i32, i32, f32, i32, vp], i32),
    "b200_softmax_bwd": ([vp, vp, vp, i64, i32, f32, vp], i32),
    "b200_ipc_export_range": ([vp, vp, vp], i32),
    "b200_embed_fwd": ([vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp], i32),
    "b200_embed_bwd": ([vp, vp, i32, vp, i32, i32, i32, i64, i32, vp], i32),
    "b200_gemm_fp8": ([vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, vp, vp, vp], i32),
    "b200_fp8_quantize": ([vp, vp, vp, i32, i32, i64, vp, vp, i32, vp], i32),
    "b200_xent_stats": ([vp, vp, vp, i32, i32, i64, vp], i32),
    "b200_xent_push_grid": ([i32], i32),
    "b200_xent_stats_push": ([vp, vp, vp, vp, i32, i32, i32, i32, i64, vp], i32),
    "b200_xent_combine_wait": ([vp, vp, C.c_uint, vp, vp, i32, i32, vp], i32),
    "b200_xent_combine": ([vp, vp, vp, i32, i32, vp], i32),
    "b200_xent_bwd": ([vp, vp, vp, vp, i32, i32, i64, vp], i32),
    "b200_adam_step": ([vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, i32, i32, i32, vp, vp, vp], i32),
    "b200_sgd_step": ([vp, vp, vp, vp, i64, f32, f32, f32, i32, vp, vp, vp], i32),
    "b200_sumsq_scratch_elems": ([], i32),
    "b200_sumsq": ([vp, i64, vp, vp, i32, vp], i32),
    "b200_scale": ([vp, i64, vp, f32, vp], i32),
    "b200_gemm_bf16": ([vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, i32, vp], i32),
    "b200_gemm_bf16_variant": ([vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, i32, i32, vp], i32),
    "b200_attn_fwd": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, i64, i64,
                       i64, i64, f32, i32, vp], i32),
    "b200_attn_fwd_v2": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, i64, i64,
                          i64, i64, f32, i32, vp], i32),
    "b200_attn_bwd_scratch_elems": ([i32, i32, i32, i32, i32], i64),
    "b200_attn_bwd": ([vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32,
                       i64, i64, i64, i64, i64, i64, i64, i64, i64, i64, i64, i64, i64, i64,
                       f32, i32, vp], i32),
    "b200_attn_bwd_v2": ([vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32,
                          i64, i64, i64, i64, i64, i64, i64, i64, i64, i64, i64, i64, i64, i64,
                          f32, i32, vp], i32),
    "b200_symm_alloc": ([i64, vp], i32),
    "b200_symm_free": ([vp], i32),
    "b200_symm_export": ([vp, vp], i32),
    "b200_symm_import": ([vp, vp], i32),
    "b200_symm_unmap": ([vp], i32),
    "b200_symm_signal": ([vp, i32, C.c_uint, vp], i32),
    "b200_symm_wait": ([vp, i32, i32, C.c_uint, vp], i32),
    "b200_copy2": ([vp, vp, vp, i64, vp], i32),
    "b200_comm_flag_words": ([i32, i32], i32),
    "b200_comm_shard_ready_word": ([i32], i32),
    "b200_comm_consumed_word": ([i32], i32),
    "b200_stage_shard": ([vp, vp, vp, i64, vp, i32, i32, C.c_uint, i32, vp, C.c_uint, vp, i32, C.c_uint, vp], i32),
    "b200_comm_debug_read": ([vp], i32),
    "b200_dp_reduce": ([vp, i32, i32, i64, f32, vp, C.c_uint, i32, vp], i32),
    "b200_adam_step_push": ([vp, vp, vp, vp, vp, i32, i64, f32, f32, f32, f32, f32, i32, i32, i32, vp, vp, vp], i32),
    "b200_dp_wait": ([vp, i32, i32, C.c_uint, C.c_uint, vp], i32),
    "b200_dp_debug_read": ([vp], i32),
    "b200_copy_async": ([vp, vp, i64, vp], i32),
    "b200_gemm_allgather": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i64, i64, C.c_uint, vp, C.c_uint, vp], i32),
    "b200_gemm_reduce_scatter": ([vp, vp, i32, vp, vp, vp, i32, i32, i32, i32, i64, i64, vp], i32),
    "b200_gemm_allreduce": ([vp, vp, vp, i32, vp, vp, i32, i32, i32, i32, i64, i64, vp], i32),
}

_OPTIONAL = set()


def _load():
    global _LIB, _LIB_ERR
    if _LIB is not None or _LIB_ERR is not None:
        return _LIB
    if os.environ.get("B200_DISABLE_KERNELS") == "1":
        _LIB_ERR = "disabled by B200_DISABLE_KERNELS=1"
        return None
    if not LIB_PATH.exists():
        try:
            from .build import build_kernels
            build_kernels()
        except Exception as e:  # pragma: no cover - build environment problem
            _LIB_ERR = f"build failed: {e}"
            return None
    try:
        lib = C.CDLL(str(LIB_PATH))
    except OSError as e:
        _LIB_ERR = str(e)
        return None
    for name, (argtypes, restype) in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if name in _OPTIONAL:
                continue
            raise
        fn.argtypes = argtypes
        fn.restype = restype
    _LIB = lib
    return _LIB


def available() -> bool:
    """True when the CUDA kernels can be used (GPU present and library loadable)."""
    if not torch.cuda.is_available():
        return False
    lib = _load()
    if lib is None:
        if os.environ.get("B200_DISABLE_KERNELS") == "1":
            return False
        raise RuntimeError(
            f"GPU present but libb200kernels.so could not be loaded ({_LIB_ERR}); "
            "run `python -m epfl_megatron_b200.ops.build`")
    return True


def lib():
    l = _load()
    if l is None:
        raise RuntimeError(f"libb200kernels.so unavailable: {_LIB_ERR}")
    return l


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    return None if t is None else t.data_ptr()


class KernelError(RuntimeError):
    pass


# Launch counter: bench.py reports how many of OUR kernels ran in the timed region.
_launches = 0


def launch_count() -> int:
    return _launches


_SYNC_CALLS = os.environ.get("B200_SYNC_CALLS", "0") == "1"


def _sync_check(name):
    """B200_SYNC_CALLS=1 (debugging): wait for the device after every entry point so that an
    asynchronous fault is reported against the kernel that caused it."""
    import torch
    try:
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        raise KernelError(f"{name}: device fault surfaced after this call: {e}") from None


def call(name: str, *args, launches: int = 1):
    """Invoke an entry point and turn its return code into an exception."""
    global _launches
    rc = getattr(lib(), name)(*args)
    if _SYNC_CALLS and launches:
        _sync_check(name)
    if rc != 0:
        if rc > 0:
            raise KernelError(f"{name}: CUDA error {rc}")
        raise KernelError(f"{name}: argument error {rc} "
                          "(-1 bad argument, -2 unsupported shape/alignment, -3 too large); "
                          f"args={args!r}")
    _launches += launches
    return rc


def try_call(name: str, *args, launches: int = 1) -> bool:
    """Like ``call`` but returns False on 'unsupported shape' so callers can fall back."""
    global _launches
    rc = getattr(lib(), name)(*args)
    if rc == 0:
        _launches += launches
        return True
    if rc == -2:
        return False
    if rc > 0:
        raise KernelError(f"{name}: CUDA error {rc}")
    raise KernelError(f"{name}: argument error {rc}")

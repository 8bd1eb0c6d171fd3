"""ctypes binding of libb200petals.so (the sm_100a kernel library, csrc/).

The library is built in-tree by ``csrc/build.py`` (also run by
``__graft_entry__.build``) and loaded once per process.  On a machine with a
GPU a missing library is a hard error: the CUDA path must never silently fall
back to torch ops.
"""
from __future__ import annotations

import ctypes
import os
import threading

_LIB = None
_LOCK = threading.Lock()
LIB_PATH = os.environ.get("B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200petals.so")


class NativeError(RuntimeError):
    pass


class LLChainArgs(ctypes.Structure):
    """Mirror of ``b200_ll_chain``."""

    _fields_ = [("seq", ctypes.c_void_p), ("x_ll", ctypes.c_void_p), ("res_ll", ctypes.c_void_p),
                ("out_ll", ctypes.c_void_p), ("ldx_ll", ctypes.c_int32), ("ldr_ll", ctypes.c_int32),
                ("ldo_ll", ctypes.c_int32), ("x_op", ctypes.c_int32), ("res_op", ctypes.c_int32),
                ("out_op", ctypes.c_int32), ("half", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class GemmArgs(ctypes.Structure):
    """Mirror of ``b200_gemm_args`` (include/b200_petals.h)."""

    _fields_ = [
        ("w_packed", ctypes.c_void_p),
        ("x", ctypes.c_void_p),
        ("out", ctypes.c_void_p),
        ("residual", ctypes.c_void_p),
        ("bias", ctypes.c_void_p),
        ("norm_w", ctypes.c_void_p),
        ("norm_b", ctypes.c_void_p),
        ("scratch", ctypes.c_void_p),
        ("counters", ctypes.c_void_p),
        ("T", ctypes.c_int32),
        ("N", ctypes.c_int32),
        ("K", ctypes.c_int32),
        ("ldx", ctypes.c_int32),
        ("ldo", ctypes.c_int32),
        ("ldr", ctypes.c_int32),
        ("n_tiles", ctypes.c_int32),
        ("dual", ctypes.c_int32),
        ("epilogue", ctypes.c_int32),
        ("norm_mode", ctypes.c_int32),
        ("norm_eps", ctypes.c_float),
        ("dtype", ctypes.c_int32),
        ("out_fp32", ctypes.c_int32),
        ("grid", ctypes.c_int32),
        ("pdl", ctypes.c_int32),
        ("debug", ctypes.c_int32),
        ("fp8", ctypes.c_int32),
        ("w_scale", ctypes.c_void_p),
        ("prefetch", ctypes.c_void_p * 2),
        ("prefetch_bytes", ctypes.c_int64 * 2),
        ("ll_buf", ctypes.c_void_p),
        ("rstd", ctypes.c_void_p),
        ("chain", LLChainArgs),
    ]


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is None:
            if not os.path.exists(LIB_PATH):
                raise NativeError(
                    f"{LIB_PATH} is missing: run `python csrc/build.py` (or __graft_entry__.build())"
                )
            L = ctypes.CDLL(LIB_PATH)
            L.b200_last_error.restype = ctypes.c_char_p
            L.b200_version.restype = ctypes.c_int
            L.b200_sm_count.restype = ctypes.c_int
            L.b200_gemm.argtypes = [ctypes.POINTER(GemmArgs), ctypes.c_void_p]
            L.b200_gemm.restype = ctypes.c_int
            L.b200_gemm_prefill.argtypes = [ctypes.POINTER(GemmArgs), ctypes.c_void_p]
            L.b200_gemm_prefill.restype = ctypes.c_int
            L.b200_gemm_scratch_bytes.argtypes = [ctypes.c_int32]
            L.b200_gemm_scratch_bytes.restype = ctypes.c_int64
            L.b200_gemm_ll_bytes.argtypes = []
            L.b200_gemm_ll_bytes.restype = ctypes.c_int64
            L.b200_gemm_token_tile.argtypes = [ctypes.c_int32, ctypes.c_int32]
            L.b200_gemm_token_tile.restype = ctypes.c_int32
            _register_optional(L)
            _LIB = L
    return _LIB


def _register_optional(L: ctypes.CDLL) -> None:
    """Prototypes of entry points added after the GEMM (present in full builds)."""
    from . import _prototypes  # noqa: WPS433 (late import keeps this module light)

    _prototypes.register(L)


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b200_last_error()
        raise NativeError(f"{what or 'native call'} failed ({rc}): {msg.decode() if msg else ''}")


def current_stream_ptr() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream

"""ctypes prototypes for the non-GEMM entry points of libb200petals.so."""
from __future__ import annotations

import ctypes as C


def _proto(L, name, argtypes, restype=C.c_int):
    fn = getattr(L, name, None)
    if fn is None:  # older build without this entry point
        return
    fn.argtypes = argtypes
    fn.restype = restype


def register(L) -> None:
    vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
    # filled in as kernels are added (see include/b200_petals.h)
    for name, args, res in PROTOTYPES:
        _proto(L, name, args, res)


PROTOTYPES: list = []

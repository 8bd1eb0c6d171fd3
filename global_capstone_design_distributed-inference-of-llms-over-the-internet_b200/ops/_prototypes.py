"""ctypes prototypes for the non-GEMM entry points of libb200petals.so."""
from __future__ import annotations

import ctypes as C

vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float


from .native import LLChainArgs as _LLChainArgs  # noqa: E402


class AttnDecodeArgs(C.Structure):
    """Mirror of ``b200_attn_decode_args``."""

    _fields_ = [
        ("qkv", vp), ("out", vp), ("kv_ptrs", vp), ("pos", vp), ("inv_freq", vp), ("scratch", vp),
        ("counters", vp),
        ("T", i32), ("n_heads", i32), ("n_kv_heads", i32), ("head_dim", i32), ("s_max", i32),
        ("splits", i32), ("ldqkv", i32), ("ldo", i32),
        ("scale", f32), ("dtype", i32), ("pdl", i32), ("ll_scratch", vp), ("chain", _LLChainArgs),
    ]


class SampleRow(C.Structure):
    """Mirror of ``b200::SampleRow`` (csrc/sampling.cu)."""

    _fields_ = [("temperature", f32), ("top_p", f32), ("penalty", f32), ("top_k", i32), ("n_recent", i32),
                ("recent", i32 * 50)]


class HopArgs(C.Structure):
    """Mirror of ``b200_hop_args``."""

    _fields_ = [("x", vp), ("extra", vp), ("ring", vp), ("counter", vp),
                ("T", i32), ("hidden", i32), ("ldx", i32), ("n_extra", i32),
                ("depth", i32), ("t_max", i32), ("pairs_per_token", i32), ("timeout_ms", i32), ("pdl", i32)]


PROTOTYPES = [
    ("b200_hop_ring_bytes", [i32, i32, i32], i64),
    ("b200_hop_send", [C.POINTER(HopArgs), vp], C.c_int),
    ("b200_hop_recv", [C.POINTER(HopArgs), vp], C.c_int),
    ("b200_hop_init", [], C.c_int),
    ("b200_ll_publish", [vp, i32, vp, i32, i32, i32, vp, i32, i32, vp], C.c_int),
    ("b200_trace_set", [vp, i32], C.c_int),
    ("b200_argmax_partial", [vp, i32, i32, i32, vp, vp, vp, vp, i32, i32, vp, vp, i32, vp], C.c_int),
    ("b200_sample_row_bytes", [], i32),
    ("b200_sample", [vp, i32, i32, i32, vp, vp, vp, vp, i32, vp], C.c_int),
    ("b200_sample_state_bytes", [], i32),
    ("b200_sample_stateful", [vp, i32, i32, i32, vp, vp, vp, vp, vp, i32, vp], C.c_int),
    ("b200_gemm_static_scratch", [vp, vp], C.c_int),
    ("b200_attn_decode", [C.POINTER(AttnDecodeArgs), vp], C.c_int),
    ("b200_attn_decode_scratch_bytes", [i32, i32, i32, i32], i64),
    ("b200_rope_append", [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp], C.c_int),
    ("b200_attn_prefill", [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, f32, i32, i32, vp], C.c_int),
    ("b200_attn_prefill_tc", [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, f32, i32, i32, vp], C.c_int),
    ("b200_embed", [vp, vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp], C.c_int),
    ("b200_norm", [vp, vp, vp, vp, i32, i32, i32, i32, f32, i32, i32, i32, vp], C.c_int),
    ("b200_rms_rstd", [vp, vp, i32, i32, i32, f32, i32, i32, vp], C.c_int),
    ("b200_argmax", [vp, i32, i32, i32, vp, vp, vp, vp, i32, i32, vp], C.c_int),
]


def register(L) -> None:
    for name, args, res in PROTOTYPES:
        fn = getattr(L, name, None)
        if fn is None:  # older build without this entry point
            continue
        fn.argtypes = args
        fn.restype = res

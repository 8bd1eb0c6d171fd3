"""Python front-ends of the session-KV attention kernels (csrc/attention.cu)."""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

from . import native
from ._prototypes import AttnDecodeArgs
from .gemm import dtype_code


def rope_inv_freq(head_dim: int, theta: float, scaling: Optional[dict] = None, device="cpu") -> torch.Tensor:
    """HF ``LlamaRotaryEmbedding`` inverse frequencies, including the llama3 rescaling."""
    import math

    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    if scaling and scaling.get("rope_type", scaling.get("type")) == "llama3":
        factor = float(scaling["factor"])
        lo, hi = float(scaling["low_freq_factor"]), float(scaling["high_freq_factor"])
        old = float(scaling["original_max_position_embeddings"])
        wavelen = 2 * math.pi / inv
        scaled = torch.where(wavelen > old / lo, inv / factor, inv)
        smooth = (old / wavelen - lo) / (hi - lo)
        mid = (1 - smooth) * inv / factor + smooth * inv
        is_mid = (wavelen >= old / hi) & (wavelen <= old / lo)
        inv = torch.where(is_mid, mid, scaled)
    return inv.to(device)


class DecodeAttnWorkspace:
    def __init__(self, device, max_tokens: int, n_heads: int, n_kv_heads: int, head_dim: int, splits: int):
        L = native.lib()
        nbytes = L.b200_attn_decode_scratch_bytes(max_tokens, n_heads, head_dim, splits)
        self.scratch = torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=device)
        self.counters = torch.zeros(max_tokens * n_kv_heads, dtype=torch.int32, device=device)
        # (value, tag) words of the ticket-free split merge; B200_ATTN_LL_MERGE=0 falls back to the ticket path
        self.ll = (torch.zeros(max(nbytes // 4, 1), dtype=torch.int64, device=device)
                   if os.environ.get("B200_ATTN_LL_MERGE", "1") != "0" else None)
        self.splits = splits
        self.max_tokens = max_tokens


def attn_decode(qkv, out, kv_ptrs, pos, inv_freq, ws: DecodeAttnWorkspace, n_heads, n_kv_heads, head_dim, s_max,
                scale=None, pdl=False, chain=None):
    a = fill_decode_args(qkv, out, kv_ptrs, pos, inv_freq, ws, n_heads, n_kv_heads, head_dim, s_max, scale, pdl)
    if chain is not None:
        a.chain = chain
    native.check(native.lib().b200_attn_decode(ctypes.byref(a), native.current_stream_ptr()), "b200_attn_decode")
    return out


def fill_decode_args(qkv, out, kv_ptrs, pos, inv_freq, ws, n_heads, n_kv_heads, head_dim, s_max, scale=None,
                     pdl=False) -> AttnDecodeArgs:
    T = qkv.shape[0]
    assert T <= ws.max_tokens
    a = AttnDecodeArgs()
    a.qkv, a.out = qkv.data_ptr(), out.data_ptr()
    a.kv_ptrs, a.pos = kv_ptrs.data_ptr(), pos.data_ptr()
    a.inv_freq = inv_freq.data_ptr() if inv_freq is not None else None
    a.scratch, a.counters = ws.scratch.data_ptr(), ws.counters.data_ptr()
    a.ll_scratch = ws.ll.data_ptr() if getattr(ws, "ll", None) is not None else None
    a.T, a.n_heads, a.n_kv_heads, a.head_dim, a.s_max = T, n_heads, n_kv_heads, head_dim, s_max
    a.splits, a.ldqkv, a.ldo = ws.splits, qkv.stride(0), out.stride(0)
    a.scale = scale if scale is not None else head_dim ** -0.5
    a.dtype, a.pdl = dtype_code(qkv.dtype), 1 if pdl else 0
    return a


def rope_append(qkv, kcache, vcache, inv_freq, n_heads, n_kv_heads, head_dim, s_max, pos0, pdl=False):
    rc = native.lib().b200_rope_append(
        qkv.data_ptr(), kcache.data_ptr(), vcache.data_ptr(),
        inv_freq.data_ptr() if inv_freq is not None else None, qkv.shape[0], n_heads, n_kv_heads, head_dim,
        s_max, qkv.stride(0), pos0, dtype_code(qkv.dtype), 1 if pdl else 0, native.current_stream_ptr())
    native.check(rc, "b200_rope_append")


TC_PREFILL_MIN_TOKENS = 32


def attn_prefill(qkv, out, kcache, vcache, n_heads, n_kv_heads, head_dim, s_max, pos0, scale=None, pdl=False,
                 tensor_cores=None):
    """Causal attention of a chunk against the session cache.  Chunks of >= 32 tokens run on the
    tcgen05 kernel (csrc/attn_prefill_tc.cu); smaller ones on the CUDA-core kernel.  ``B200_PREFILL_TC=0``
    forces the latter."""
    import os

    if tensor_cores is None:
        tensor_cores = qkv.shape[0] >= TC_PREFILL_MIN_TOKENS and os.environ.get("B200_PREFILL_TC", "1") != "0"
    fn = native.lib().b200_attn_prefill_tc if tensor_cores else native.lib().b200_attn_prefill
    rc = fn(
        qkv.data_ptr(), out.data_ptr(), kcache.data_ptr(), vcache.data_ptr(), qkv.shape[0], n_heads, n_kv_heads,
        head_dim, s_max, qkv.stride(0), out.stride(0), pos0,
        scale if scale is not None else head_dim ** -0.5, dtype_code(qkv.dtype), 1 if pdl else 0,
        native.current_stream_ptr())
    native.check(rc, "b200_attn_prefill")
    return out

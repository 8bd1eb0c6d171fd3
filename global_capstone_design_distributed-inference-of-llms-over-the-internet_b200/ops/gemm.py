"""Python front-end of the swap-AB stream-K tcgen05 GEMM (csrc/gemm.cu).

``linear(x, w)`` computes ``epilogue(norm(x) @ W^T)`` for a packed weight with
optional fused RMSNorm/LayerNorm prologue (decode-size inputs), bias,
GELU-tanh / SwiGLU activation and residual add.  Replaces the reference's
separate torch ops (petals/llama/block.py:88-90,151,208-238).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import torch

from . import native
from .packing import PackedWeight

EPI_NONE, EPI_GELU_TANH, EPI_SWIGLU = 0, 1, 2
NORM_NONE, NORM_RMS, NORM_LAYER = 0, 1, 2
MAX_FUSED_NORM_TOKENS = 32


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float16:
        return 0
    if dt == torch.bfloat16:
        return 1
    raise ValueError(f"unsupported activation dtype {dt}")


@dataclass
class GemmWorkspace:
    """Stream-K partial-tile scratch + per-tile arrival counters (kept zero)."""

    scratch: torch.Tensor
    counters: torch.Tensor
    max_tokens: int
    ll: Optional[torch.Tensor] = None     # ticket-free fix-up words (b200_gemm_args.ll_buf), zeroed once

    @classmethod
    def create(cls, device, max_tokens: int = 16, max_units: int = 1 << 16) -> "GemmWorkspace":
        with torch.cuda.device(device):
            nbytes = native.lib().b200_gemm_scratch_bytes(int(max_tokens))
        scratch = torch.empty(nbytes // 4, dtype=torch.float32, device=device)
        counters = torch.zeros(max_units, dtype=torch.int32, device=device)
        ll = None
        if _ll_fixup_enabled():
            ll = torch.zeros(native.lib().b200_gemm_ll_bytes() // 8, dtype=torch.int64, device=device)
        return cls(scratch, counters, max_tokens, ll)


_default_ws: dict = {}


def default_workspace(device, tokens: int) -> GemmWorkspace:
    key = (torch.device(device).index or 0)
    ws = _default_ws.get(key)
    if ws is None or ws.max_tokens < tokens:
        ws = GemmWorkspace.create(device, max(tokens, 16))
        _default_ws[key] = ws
    return ws


def linear(
    x: torch.Tensor,
    w: PackedWeight,
    *,
    out: Optional[torch.Tensor] = None,
    residual: Optional[torch.Tensor] = None,
    epilogue: int = EPI_NONE,
    norm_mode: int = NORM_NONE,
    norm_w: Optional[torch.Tensor] = None,
    norm_b: Optional[torch.Tensor] = None,
    norm_eps: float = 1e-5,
    out_fp32: bool = False,
    workspace: Optional[GemmWorkspace] = None,
    pdl: bool = False,
    grid: int = 0,
    debug: int = 0,
    prefetch=None,
    rstd: Optional[torch.Tensor] = None,
    chain=None,
) -> torch.Tensor:
    """x: [T, K] (16-bit, last dim contiguous) -> [T, N].

    ``chain``: a :class:`petals_b200.ops.native.LLChainArgs` making this launch a link of a flag-chained decode step
    (see ``b200_ll_chain`` in the header)."""
    a, out = fill_args(x, w, out=out, residual=residual, epilogue=epilogue, norm_mode=norm_mode, norm_w=norm_w,
                       norm_b=norm_b, norm_eps=norm_eps, out_fp32=out_fp32, workspace=workspace, pdl=pdl, grid=grid,
                       debug=debug, prefetch=prefetch, rstd=rstd, chain=chain)
    if uses_prefill_kernel(x.shape[0], w, norm_mode):
        # token-tile-filling inputs: TMA-fed kernel with the epilogue overlapped (csrc/gemm_prefill.cu)
        rc = native.lib().b200_gemm_prefill(ctypes.byref(a), native.current_stream_ptr())
        native.check(rc, "b200_gemm_prefill")
        return out
    rc = native.lib().b200_gemm(ctypes.byref(a), native.current_stream_ptr())
    native.check(rc, "b200_gemm")
    return out


PREFILL_MIN_TOKENS = 64


def uses_prefill_kernel(T: int, w: PackedWeight, norm_mode: int = NORM_NONE) -> bool:
    """The routing rule of :func:`linear` (callers that cache argument structs launch the kernel themselves)."""
    return T >= PREFILL_MIN_TOKENS and norm_mode == NORM_NONE and not w.fp8 and _prefill_kernel_enabled()


def _ll_fixup_enabled() -> bool:
    import os

    return os.environ.get("B200_GEMM_LL_FIXUP", "1") != "0"


def _prefill_kernel_enabled() -> bool:
    import os

    return os.environ.get("B200_PREFILL_GEMM", "1") != "0"


def fill_args(x, w, *, out=None, residual=None, epilogue=EPI_NONE, norm_mode=NORM_NONE, norm_w=None, norm_b=None,
              norm_eps=1e-5, out_fp32=False, workspace=None, pdl=False, grid=0, debug=0, prefetch=None, rstd=None,
              chain=None):
    """Builds the ``b200_gemm_args`` for one call (also used to describe ops of the persistent kernel)."""
    assert x.dim() == 2 and x.stride(1) == 1 and x.is_cuda
    T, K = x.shape
    assert K == w.k_padded, f"activation width {K} != packed K {w.k_padded}"
    if w.dual:
        epilogue = EPI_SWIGLU
    if out is None:
        out = torch.empty((T, w.n), dtype=torch.float32 if out_fp32 else x.dtype, device=x.device)
    ws = workspace or default_workspace(x.device, T)
    assert w.n_tiles * ((T + 255) // 256 + 1) <= ws.counters.numel() or T <= 256
    a = native.GemmArgs()
    a.w_packed = w.data.data_ptr()
    a.x = x.data_ptr()
    a.out = out.data_ptr()
    a.residual = residual.data_ptr() if residual is not None else None
    a.bias = w.bias.data_ptr() if w.bias is not None else None
    a.norm_w = norm_w.data_ptr() if norm_w is not None else None
    a.norm_b = norm_b.data_ptr() if norm_b is not None else None
    a.scratch = ws.scratch.data_ptr()
    a.counters = ws.counters.data_ptr()
    a.T, a.N, a.K = T, w.n, K
    a.ldx = x.stride(0)
    a.ldo = out.stride(0)
    a.ldr = residual.stride(0) if residual is not None else 0
    a.n_tiles = w.n_tiles
    a.dual = 1 if w.dual else 0
    a.epilogue = epilogue
    a.norm_mode = norm_mode
    a.norm_eps = norm_eps
    a.dtype = dtype_code(x.dtype)
    a.out_fp32 = 1 if out_fp32 else 0
    a.grid = grid
    a.pdl = 1 if pdl else 0
    a.debug = debug
    a.fp8 = (2 if getattr(w, "mx", False) else 1) if w.fp8 else 0
    a.w_scale = w.scale.data_ptr() if w.scale is not None else None
    a.ll_buf = ws.ll.data_ptr() if ws.ll is not None else None
    a.rstd = rstd.data_ptr() if rstd is not None else None
    if chain is not None:
        a.chain = chain
    # prefetch: up to two (PackedWeight | tensor, first_byte, n_bytes) ranges to warm the L2 with
    for i, (t, off, nbytes) in enumerate(prefetch or ()):
        data = t.data if hasattr(t, "n_tiles") else t
        total = data.numel() * data.element_size()
        off = min(off, total) & ~15
        a.prefetch[i] = data.data_ptr() + off
        a.prefetch_bytes[i] = max(0, min(nbytes, total - off))
    return a, out


def chain_capable(w: PackedWeight, T: int = 1) -> bool:
    """Would b200_gemm, as a link of a flag chain, exchange the stream-K partials of this weight through the
    per-launch-tagged wide words (the only exchange two overlapping launches may share)?  Mirrors the grid choice
    in csrc/gemm.cu: 2..9 CTAs per output tile, each with at least one k-block."""
    if w.fp8 or T > 16:
        return False
    sms = native.lib().b200_sm_count()
    kbpt = w.k_padded // 64
    parts = min((2 * sms) // w.n_tiles, 9, kbpt)
    return parts >= 2 and parts * w.n_tiles <= 512

"""Python front-ends of csrc/elementwise.cu (embedding, norm, argmax)."""
from __future__ import annotations

from typing import Optional

import torch

from . import native
from .gemm import dtype_code


def embed(ids: torch.Tensor, wte: torch.Tensor, out: torch.Tensor, wpe: Optional[torch.Tensor] = None,
          pos: Optional[torch.Tensor] = None, pos0: int = 0, pdl: bool = False):
    assert ids.dtype == torch.int64 and ids.is_cuda
    T, H = out.shape
    rc = native.lib().b200_embed(ids.data_ptr(), pos.data_ptr() if pos is not None else None, pos0, wte.data_ptr(),
                                 wpe.data_ptr() if wpe is not None else None, out.data_ptr(), T, H, wte.shape[0],
                                 wpe.shape[0] if wpe is not None else 1 << 30, dtype_code(out.dtype),
                                 1 if pdl else 0, native.current_stream_ptr())
    native.check(rc, "b200_embed")
    return out


def norm(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], eps: float, layer_norm: bool,
         out: Optional[torch.Tensor] = None, pdl: bool = False):
    if out is None:
        out = torch.empty_like(x)
    T, H = x.shape
    rc = native.lib().b200_norm(x.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, out.data_ptr(),
                                T, H, x.stride(0), out.stride(0), eps, 1 if layer_norm else 0, dtype_code(x.dtype),
                                1 if pdl else 0, native.current_stream_ptr())
    native.check(rc, "b200_norm")
    return out


def rms_rstd(x: torch.Tensor, eps: float, out: torch.Tensor, pdl: bool = False) -> torch.Tensor:
    """out[t] = 1/sqrt(mean(x[t]^2) + eps) (fp32), bit-identical to the statistics the GEMM would compute itself."""
    T, K = x.shape
    assert out.dtype == torch.float32 and out.numel() >= T
    rc = native.lib().b200_rms_rstd(x.data_ptr(), out.data_ptr(), T, K, x.stride(0), eps, dtype_code(x.dtype),
                                    1 if pdl else 0, native.current_stream_ptr())
    native.check(rc, "b200_rms_rstd")
    return out


class ArgmaxWorkspace:
    def __init__(self, device, max_tokens: int, chunks: int = 64):
        self.chunks = chunks
        self.part_val = torch.empty(max_tokens * chunks, dtype=torch.float32, device=device)
        self.part_idx = torch.empty(max_tokens * chunks, dtype=torch.int32, device=device)
        self.counters = torch.zeros(max_tokens, dtype=torch.int32, device=device)


def argmax(logits: torch.Tensor, out_ids: Optional[torch.Tensor], ws: ArgmaxWorkspace, pdl: bool = False,
           idx_offset: int = 0, init_pair: Optional[torch.Tensor] = None, out_pair: Optional[torch.Tensor] = None):
    """First index of the row maximum.  For a vocabulary SHARD: ``idx_offset`` is added to the indices,
    ``init_pair`` ([T, 2] int32: float bits, index) is another shard's winner to beat (lower index wins ties)
    and ``out_pair`` receives this shard's winner in the same form."""
    assert logits.dtype == torch.float32 and (out_ids is None or out_ids.dtype == torch.int64)
    T, V = logits.shape
    for t in (init_pair, out_pair):
        assert t is None or (t.dtype == torch.int32 and t.is_contiguous() and t.numel() >= 2 * T)
    rc = native.lib().b200_argmax_partial(logits.data_ptr(), logits.stride(0), T, V, ws.part_val.data_ptr(),
                                          ws.part_idx.data_ptr(), ws.counters.data_ptr(),
                                          out_ids.data_ptr() if out_ids is not None else None, ws.chunks, idx_offset,
                                          init_pair.data_ptr() if init_pair is not None else None,
                                          out_pair.data_ptr() if out_pair is not None else None,
                                          1 if pdl else 0, native.current_stream_ptr())
    native.check(rc, "b200_argmax")
    return out_ids


def ll_publish(x: torch.Tensor, out_ll: torch.Tensor, seq: torch.Tensor, op: int, pdl: bool = False) -> None:
    """Head of a flag chain: x [T, H] (16-bit) -> LL words in out_ll [T, ld] (int64) tagged (seq << 12 | op)."""
    T, H = x.shape
    rc = native.lib().b200_ll_publish(x.data_ptr(), x.stride(0), out_ll.data_ptr(), out_ll.stride(0), T, H,
                                      seq.data_ptr(), op, 1 if pdl else 0, native.current_stream_ptr())
    native.check(rc, "b200_ll_publish")

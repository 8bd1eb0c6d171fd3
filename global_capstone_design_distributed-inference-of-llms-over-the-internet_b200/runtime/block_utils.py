"""Block sizing helpers (upstream petals/server/block_utils.py:12-65, server.py:160-209, 275-326).

* ``resolve_block_dtype``     "auto" -> bf16 on CUDA, fp32 on CPU.
* ``get_block_size``          bytes of one decoder block for a weight format (bf16/fp16 = 2 B,
                              fp8 = 1 B + per-row scales; upstream: NF4 = 4.25 bit, INT8 = 1 B).
* ``default_limits``          ``inference_max_length`` 8192 (GQA) / 2048 and ``attn_cache_tokens``
                              16384 / 4096, as upstream picks them.
* ``choose_num_blocks``       how many blocks fit a GPU once the KV budget and a safety reserve
                              are set aside (B200: 180 GB per GPU).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ..models.config import ModelConfig


def resolve_block_dtype(cfg: ModelConfig, dtype, device) -> torch.dtype:
    if dtype not in (None, "auto"):
        return dtype
    return torch.bfloat16 if torch.device(device).type == "cuda" else torch.float32


def get_block_size(cfg: ModelConfig, dtype: torch.dtype = torch.bfloat16, weight_format: str = "native") -> int:
    params = cfg.layer_param_count()
    if weight_format == "fp8":
        rows = cfg.qkv_size + cfg.hidden + (2 if cfg.is_llama else 1) * cfg.intermediate + cfg.hidden
        return params + 4 * rows
    if weight_format == "mxfp8":
        return params + params // 32          # one UE8M0 byte per 32 elements of K (OCP MX)
    return params * torch.empty((), dtype=dtype).element_size()


def default_limits(cfg: ModelConfig) -> Tuple[int, int]:
    """(inference_max_length, attn_cache_tokens)."""
    gqa = cfg.n_kv_heads < cfg.n_heads
    return (8192, 16384) if gqa else (2048, 4096)


def cache_bytes_per_block(cfg: ModelConfig, attn_cache_tokens: int, dtype: torch.dtype = torch.bfloat16) -> int:
    return 2 * cfg.kv_size * attn_cache_tokens * torch.empty((), dtype=dtype).element_size()


def choose_num_blocks(cfg: ModelConfig, device, dtype: torch.dtype = torch.bfloat16, weight_format: str = "native",
                      attn_cache_tokens: Optional[int] = None, total_memory: Optional[int] = None,
                      reserve_bytes: int = 4 << 30) -> int:
    if total_memory is None:
        d = torch.device(device)
        if d.type != "cuda":
            raise ValueError("choose_num_blocks needs total_memory for non-CUDA devices")
        total_memory = torch.cuda.get_device_properties(d).total_memory
    tokens = attn_cache_tokens if attn_cache_tokens is not None else default_limits(cfg)[1]
    per_block = get_block_size(cfg, dtype, weight_format) + cache_bytes_per_block(cfg, tokens, dtype)
    n = int((total_memory - reserve_bytes) // per_block)
    if n < 1:
        raise ValueError(f"not enough memory for a single block ({total_memory} B total, {per_block} B per block)")
    return min(n, cfg.n_layers)

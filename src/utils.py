"""Cache / position helpers (reference: src/utils.py:5-64).

The reference needs these to shuttle HF ``Cache`` objects and legacy ``(k, v)`` tuples
around; here the cache is a :class:`SessionKV`, so the helpers reduce to views of it.  They
keep the reference's names so scripts written against it still read naturally."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from petals_b200.runtime.kv_cache import SessionKV


def extract_kv_tuple(past: Optional[SessionKV], layer: int = 0) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """(k, v) of one layer trimmed to the cached length: ``[n_kv, length, head_dim]`` each."""
    if past is None:
        return None
    return past.k(layer)[:, : past.length], past.v(layer)[:, : past.length]


def past_length(past: Optional[SessionKV]) -> int:
    return 0 if past is None else past.length


def default_position_ids(past: Optional[SessionKV], seq_len: int, device=None) -> torch.Tensor:
    p0 = past_length(past)
    return torch.arange(p0, p0 + seq_len, device=device, dtype=torch.long).unsqueeze(0)


def normalize_cache(past):
    """Identity for SessionKV (the reference converts HF Cache -> legacy tuples here)."""
    return past

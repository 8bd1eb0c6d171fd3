"""Token sampling with the reference's semantics (src/rpc_handler.py:327-403).

* ``temperature <= 0``  -> greedy arg-max (no repetition penalty), ``:331-335``.
* otherwise: count-based repetition penalty over the last 50 generated tokens
  (``penalty ** count``; an extra ``penalty ** 3`` when the last three tokens are
  identical) ``:352-372``; ``softmax(logits / T)``; top-k (keep k largest probs)
  ``:376-379``; top-p over the sorted probs keeping entries whose cumulative sum is
  ``<= top_p`` and always the first ``:383-392``; renormalise; multinomial.

The reference mutates the logits with one host-synchronising indexing op per
penalised token; here the whole thing is a handful of vectorised device ops and one
``.item()`` at the end.  Handler defaults 0.8 / 0.9 / 0 and repetition penalty 1.5
(``:71-73, 164``) are kept in ``SamplingParams``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence

import torch


@dataclass
class SamplingParams:
    temperature: float = 0.8
    top_p: float = 0.9
    top_k: int = 0
    repetition_penalty: float = 1.5

    @classmethod
    def from_metadata(cls, md: dict) -> "SamplingParams":
        return cls(float(md.get("temperature", 0.8)), float(md.get("top_p", 0.9)), int(md.get("top_k", 0)),
                   float(md.get("repetition_penalty", 1.5)))


def apply_repetition_penalty(logits: torch.Tensor, generated: Sequence[int], penalty: float) -> torch.Tensor:
    """logits: [V] (modified copy returned)."""
    if penalty == 1.0 or not generated:
        return logits
    V = logits.shape[-1]
    recent = [t for t in generated[-50:] if 0 <= t < V]
    logits = logits.clone()
    if recent:
        ids = torch.tensor(recent, device=logits.device, dtype=torch.long)
        uniq, counts = torch.unique(ids, return_counts=True)
        pen = torch.pow(torch.tensor(penalty, device=logits.device, dtype=torch.float32), counts.float())
        vals = logits[uniq].float()
        logits[uniq] = torch.where(vals > 0, vals / pen, vals * pen).to(logits.dtype)
    if len(generated) >= 3 and len(set(generated[-3:])) == 1 and 0 <= generated[-1] < V:
        r = generated[-1]
        strong = penalty ** 3
        v = logits[r].float()
        logits[r] = (v / strong if v > 0 else v * strong).to(logits.dtype)
    return logits


def filtered_probs(logits: torch.Tensor, params: SamplingParams,
                   generated: Optional[Sequence[int]] = None) -> torch.Tensor:
    """The exact categorical distribution the reference samples from. logits: [V]."""
    temp = max(params.temperature, 1e-5)
    logits = apply_repetition_penalty(logits.float(), list(generated or []), params.repetition_penalty)
    probs = torch.softmax(logits / temp, dim=-1)
    if 0 < params.top_k < probs.numel():
        tv, ti = torch.topk(probs, params.top_k)
        probs = torch.zeros_like(probs).scatter(-1, ti, tv)
    if 0.0 < params.top_p < 1.0:
        sp, si = torch.sort(probs, descending=True)
        keep = torch.cumsum(sp, -1) <= params.top_p
        keep[0] = True
        f = sp * keep
        f = f / f.sum()
        probs = torch.zeros_like(probs).scatter(-1, si, f)
    return probs / probs.sum()


def sample_token(logits: torch.Tensor, params: SamplingParams, generated: Optional[Sequence[int]] = None,
                 generator: Optional[torch.Generator] = None) -> int:
    """logits: [V] or [1, V]."""
    logits = logits.reshape(-1)
    if params.temperature <= 0.0:
        return int(torch.argmax(logits).item())
    probs = filtered_probs(logits, params, generated)
    return int(torch.multinomial(probs, 1, generator=generator).item())

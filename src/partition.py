"""GPT-2 layer-span partitioning (the reference's ``partition.py`` is a Llama-only re-export
shim, src/partition.py:1-8, even though its README and launcher default to ``gpt2`` — SURVEY B2).

GPT-2 differences handled natively: LayerNorm with bias, fused ``c_attn`` with bias, GELU-tanh
MLP, learned positions added at the embedding, ``Conv1D`` weights stored ``[in, out]``
(transposed once at load), LM head tied to ``wte``.

``load_stage_model`` here also accepts Llama models and dispatches by the model's family, so
``from src.partition import ...`` works as the generic entry point.
"""
from __future__ import annotations

from typing import Optional

import torch

from petals_b200.runtime.stages import Stage0, StageLast, StageSegment, StageWeights, load_stage_weights  # noqa: F401


def load_stage_model(model_name: str, device, role: str, *, start: int = 0, end: Optional[int] = None,
                     dtype: torch.dtype = torch.float16, use_cpu_offload: bool = False, seed: int = 0) -> StageWeights:
    return load_stage_weights(model_name, device, role, start=start, end=end, dtype=dtype,
                              use_cpu_offload=use_cpu_offload, seed=seed, families=("gpt2", "llama"))

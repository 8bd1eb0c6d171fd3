"""Llama-family layer-span partitioning: ``load_stage_model`` + ``Stage0/StageSegment/StageLast``.

Same four public names and call shapes as the reference module (src/llama_partition.py:76-550),
re-built on the framework's span engines: only the requested blocks are read from the
checkpoint, prefill is causal, the KV cache is a pre-allocated per-session buffer, and on a
B200 the blocks run on the sm_100a kernels.  Any number of stages is supported
(``--splits a,b,c,...``), not just the reference's fixed four.
"""
from __future__ import annotations

from typing import Optional

import torch

from petals_b200.runtime.stages import Stage0 as _Stage0
from petals_b200.runtime.stages import StageLast as _StageLast
from petals_b200.runtime.stages import StageSegment as _StageSegment
from petals_b200.runtime.stages import StageWeights, load_stage_weights

FAMILY = ("llama",)


def load_stage_model(model_name: str, device, role: str, *, start: int = 0, end: Optional[int] = None,
                     dtype: torch.dtype = torch.float16, use_cpu_offload: bool = False, seed: int = 0) -> StageWeights:
    """Weights of one stage. role: 'stage0' (embed + [0,end)), 'segment' ([start,end)), 'last' ([start,L) + head)."""
    return load_stage_weights(model_name, device, role, start=start, end=end, dtype=dtype,
                              use_cpu_offload=use_cpu_offload, seed=seed, families=FAMILY)


def _check(full: StageWeights):
    if full.cfg.family not in FAMILY:
        raise ValueError(f"llama_partition handles Llama/Mistral-style models, got {full.cfg.family!r}; "
                         "use src.partition for GPT-2")


class Stage0(_Stage0):
    def __init__(self, full: StageWeights, end: int, **kw):
        _check(full)
        super().__init__(full, end, **kw)


class StageSegment(_StageSegment):
    def __init__(self, full: StageWeights, start: int, end: int, gpu_device=None, keep_layers_on_gpu: int = 0, **kw):
        _check(full)
        super().__init__(full, start, end, gpu_device=gpu_device, keep_layers_on_gpu=keep_layers_on_gpu, **kw)


class StageLast(_StageLast):
    def __init__(self, full: StageWeights, start: int, gpu_device=None, keep_layers_on_gpu: int = 0, **kw):
        _check(full)
        super().__init__(full, start, gpu_device=gpu_device, keep_layers_on_gpu=keep_layers_on_gpu, **kw)

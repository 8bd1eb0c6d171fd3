"""CPU-offload execution: weights live in pinned host memory and stream through the GPU.

The reference moves layer *i* to the GPU and layer *i-1* back to the CPU inside every
forward with synchronous ``.to()`` calls, for every server stage unless
``--keep_layers_on_gpu >= span`` (src/llama_partition.py:161-220,233-246,290-293).
Here offload is an explicit opt-in (``--use_cpu_offload``): the first
``keep_layers_on_gpu`` blocks stay resident, the rest are copied host->device on a
side stream one block ahead of the compute (two device slots, so the copy of block i+1
overlaps block i), and nothing is ever copied back (weights are read-only).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ..models.config import ModelConfig
from ..models.torch_ref import TorchSpan
from .kv_cache import SessionKV


class OffloadedSpan(TorchSpan):
    def __init__(self, cfg: ModelConfig, weights: Dict[str, object], device, dtype, keep_layers_on_gpu: int = 0):
        dev = torch.device(device)
        host_layers: List[Dict[str, torch.Tensor]] = weights["layers"]  # type: ignore[assignment]
        glob = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in weights.items() if k != "layers"}
        glob["layers"] = []
        super().__init__(cfg, glob, dev, dtype)
        self.keep = max(0, min(keep_layers_on_gpu, len(host_layers)))
        self.resident = [{k: v.to(dev) for k, v in L.items()} for L in host_layers[: self.keep]]
        self.host = [{k: (v.pin_memory() if dev.type == "cuda" else v) for k, v in L.items()}
                     for L in host_layers[self.keep:]]
        self.copy_stream = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self.slots: List[Optional[Dict[str, torch.Tensor]]] = [None, None]
        self.bytes_streamed = 0

    def _fetch(self, j: int, slot: int):
        """Start copying streamed block j into a device slot; returns an event."""
        L = self.host[j]
        if self.copy_stream is None:
            self.slots[slot] = L
            return None
        with torch.cuda.stream(self.copy_stream):
            self.slots[slot] = {k: v.to(self.device, non_blocking=True) for k, v in L.items()}
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self.bytes_streamed += sum(v.numel() * v.element_size() for v in L.values())
        return ev

    @torch.inference_mode()
    def forward(self, h: torch.Tensor, kv: SessionKV, pos0: Optional[int] = None) -> torch.Tensor:
        pos0 = kv.length if pos0 is None else pos0
        T = h.shape[0]
        fn = self._llama_layer if self.cfg.is_llama else self._gpt2_layer
        li = 0
        for L in self.resident:
            h = fn(h, L, kv, li, pos0)
            li += 1
        ev = self._fetch(0, 0) if self.host else None
        for j in range(len(self.host)):
            nxt = self._fetch(j + 1, (j + 1) & 1) if j + 1 < len(self.host) else None
            if ev is not None:
                torch.cuda.current_stream(self.device).wait_event(ev)
            h = fn(h, self.slots[j & 1], kv, li, pos0)
            if self.copy_stream is not None:
                # the slot may be overwritten two blocks later: make the copy stream wait for this compute
                self.copy_stream.wait_stream(torch.cuda.current_stream(self.device))
            li += 1
            ev = nxt
        kv.length = pos0 + T
        kv.touch()
        return h

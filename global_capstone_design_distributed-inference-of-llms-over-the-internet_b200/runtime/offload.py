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


class OffloadedNativeSpan:
    """The same streaming scheme with the sm_100a kernels doing the compute (``--use_cpu_offload`` on a B200).

    Blocks are packed into the GEMM's tiled layout ONCE (on the GPU, one block at a time) and the packed image of
    every non-resident block is parked in pinned host memory; at run time two device slots receive block i+1 on a
    side stream while the kernels of block i run from the other slot - one ``cudaMemcpyAsync`` per packed tensor,
    no re-packing, nothing copied back.  Built by :func:`make_offloaded_native_span`; behaves like a NativeSpan
    (``prefill`` / ``decode`` / ``forward``), without CUDA graphs because the weight pointers alternate between slots.
    """


def make_offloaded_native_span(cfg: ModelConfig, weights: Dict[str, object], device, dtype, keep_layers_on_gpu: int = 0,
                               **span_kw):
    from ..ops import packing
    from .native_span import NativeSpan

    dev = torch.device(device)
    host_layers: List[Dict[str, torch.Tensor]] = weights["layers"]  # type: ignore[assignment]
    keep = max(0, min(keep_layers_on_gpu, len(host_layers)))
    w = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in weights.items() if k != "layers"}
    w["layers"] = [{k: v.to(dev) for k, v in L.items()} for L in host_layers[:keep]]

    class _Span(NativeSpan, OffloadedNativeSpan):
        def _iter_layers(self):
            li = 0
            for P in self.layers:
                yield li, P
                li += 1
            cur = torch.cuda.current_stream(self.device)
            ev = self._fetch(0, 0) if self.host_blocks else None
            for j in range(len(self.host_blocks)):
                nxt = self._fetch(j + 1, (j + 1) & 1) if j + 1 < len(self.host_blocks) else None
                cur.wait_event(ev)
                yield li, self.slots[j & 1]
                self.copy_stream.wait_stream(cur)      # the slot is overwritten two blocks later: not before this compute
                li += 1
                ev = nxt

        def _fetch(self, j: int, slot: int):
            with torch.cuda.stream(self.copy_stream):
                for name, dst in self.slot_tensors[slot].items():
                    src = self.host_blocks[j][name]
                    dst.copy_(src, non_blocking=True)
                    self.bytes_streamed += src.numel() * src.element_size()
                ev = torch.cuda.Event()
                ev.record(self.copy_stream)
            return ev

    n_stream = len(host_layers) - keep
    span = _Span(cfg, w, dev, dtype, use_graphs=False, streamed_layers=n_stream, **span_kw)
    span.copy_stream = torch.cuda.Stream(device=dev)
    span.bytes_streamed = 0
    span.host_blocks = []
    names = ("qkv", "o", "mlp_in", "down")
    first = None
    with torch.cuda.device(dev):
        for L in host_layers[keep:]:
            Ld = {k: v.to(dev) for k, v in L.items()}
            P = {"qkv": packing.pack_weight(Ld["qkv_w"], Ld.get("qkv_b")), "o": packing.pack_weight(Ld["o_w"], Ld.get("o_b")),
                 "down": packing.pack_weight(Ld["down_w"], Ld.get("down_b"))}
            P["mlp_in"] = packing.pack_dual(Ld["gate_w"], Ld["up_w"]) if cfg.is_llama else packing.pack_weight(Ld["fc_w"], Ld.get("fc_b"))
            blk = {n: P[n].data.cpu().pin_memory() for n in names}
            for k in ("ln1_w", "ln1_b", "ln2_w", "ln2_b"):
                if Ld.get(k) is not None:
                    blk[k] = Ld[k].cpu().pin_memory()
            for n in names:                     # biases are tiny: they travel with the block too
                if P[n].bias is not None:
                    blk[n + "_bias"] = P[n].bias.cpu().pin_memory()
            span.host_blocks.append(blk)
            if first is None:
                first = P
            del Ld
        # two device slots shaped like a packed block
        span.slot_tensors, span.slots = [], []
        for _ in range(2 if n_stream else 0):
            t = {k: torch.empty_like(v, device=dev) for k, v in span.host_blocks[0].items()}
            span.slot_tensors.append(t)
            from dataclasses import replace
            slot = {"ln1_w": t["ln1_w"], "ln1_b": t.get("ln1_b"), "ln2_w": t["ln2_w"], "ln2_b": t.get("ln2_b")}
            for n in names:
                slot[n] = replace(first[n], data=t[n], bias=t.get(n + "_bias"))
            span.slots.append(slot)
    return span

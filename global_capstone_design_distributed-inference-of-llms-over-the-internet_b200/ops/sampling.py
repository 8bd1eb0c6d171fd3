"""Device sampling kernel front-end (csrc/sampling.cu)."""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch

from . import native
from ._prototypes import SampleRow


class DeviceSampler:
    """Samples one token per row of fp32 logits entirely on the GPU."""

    def __init__(self, device, max_rows: int, vocab: int):
        L = native.lib()
        assert L.b200_sample_row_bytes() == ctypes.sizeof(SampleRow)
        self.device, self.max_rows, self.vocab = torch.device(device), max_rows, vocab
        self.rows_host = torch.zeros(max_rows * ctypes.sizeof(SampleRow), dtype=torch.uint8).pin_memory()
        self.rows_dev = torch.zeros_like(self.rows_host, device=self.device)
        self.probs = torch.empty(max_rows, vocab, dtype=torch.float32, device=self.device)
        self.ids = torch.zeros(max_rows, dtype=torch.int64, device=self.device)

    def sample(self, logits: torch.Tensor, params, generated: Sequence[Sequence[int]],
               uniforms: Optional[torch.Tensor] = None, generator: Optional[torch.Generator] = None) -> torch.Tensor:
        """logits [T, V] fp32 (device); ``params``: one SamplingParams or a list; returns ids [T] int64 (device)."""
        T, V = logits.shape
        assert T <= self.max_rows and V == self.vocab and logits.dtype == torch.float32
        plist = params if isinstance(params, (list, tuple)) else [params] * T
        arr = (SampleRow * T).from_buffer(self.rows_host.numpy())
        for t in range(T):
            r, pr = arr[t], plist[t]
            r.temperature, r.top_p, r.penalty, r.top_k = pr.temperature, pr.top_p, pr.repetition_penalty, pr.top_k
            rec = list(generated[t])[-50:] if generated and generated[t] else []
            r.n_recent = len(rec)
            for j, tok in enumerate(rec):
                r.recent[j] = int(tok)
        n = T * ctypes.sizeof(SampleRow)
        self.rows_dev[:n].copy_(self.rows_host[:n], non_blocking=True)
        if uniforms is None:
            uniforms = torch.rand(T, device=self.device, generator=generator)
        rc = native.lib().b200_sample(logits.data_ptr(), logits.stride(0), T, V, self.rows_dev.data_ptr(),
                                      uniforms.data_ptr(), self.probs.data_ptr(), self.ids.data_ptr(), 0,
                                      native.current_stream_ptr())
        native.check(rc, "b200_sample")
        return self.ids[:T]

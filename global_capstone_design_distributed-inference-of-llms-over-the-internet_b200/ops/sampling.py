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


def philox_uniform(seed: int, offset: int) -> float:
    """Host restatement of the kernel's generator: first word of Philox4x32-10(key = seed, counter = offset) as a float
    in [0, 1) with 24 random bits."""
    M = 0xFFFFFFFF
    c = [offset & M, (offset >> 32) & M, 0, 0]
    k0, k1 = seed & M, (seed >> 32) & M
    for _ in range(10):
        p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
        c = [((p1 >> 32) ^ c[1] ^ k0) & M, p1 & M, ((p0 >> 32) ^ c[3] ^ k1) & M, p0 & M]
        k0, k1 = (k0 + 0x9E3779B9) & M, (k1 + 0xBB67AE85) & M
    return (c[0] >> 8) / 16777216.0


class SamplerPool:
    """Device-resident sampling state of many sessions (b200_sample_stateful): per session one SampleRow (parameters +
    window of recent tokens) and one Philox stream (seed, draws so far), both updated by the kernel itself."""

    def __init__(self, device, n_rows: int, vocab: int, max_batch: int = 32):
        L = native.lib()
        assert L.b200_sample_row_bytes() == ctypes.sizeof(SampleRow) and L.b200_sample_state_bytes() == 16
        self.device, self.n_rows, self.vocab = torch.device(device), n_rows, vocab
        self.row_bytes = ctypes.sizeof(SampleRow)
        self.rows = torch.zeros(n_rows * self.row_bytes, dtype=torch.uint8, device=self.device)
        self.states = torch.zeros(n_rows, 2, dtype=torch.int64, device=self.device)
        self.probs = torch.empty(max_batch, vocab, dtype=torch.float32, device=self.device)
        self._free = list(range(n_rows - 1, -1, -1))

    def open(self, params, seed: int, recent: Sequence[int] = ()) -> int:
        """A row for a new session; returns its index."""
        idx = self._free.pop()
        host = torch.zeros(self.row_bytes, dtype=torch.uint8)
        r = SampleRow.from_buffer(host.numpy())
        r.temperature, r.top_p, r.penalty, r.top_k = params.temperature, params.top_p, params.repetition_penalty, params.top_k
        rec = list(recent)[-50:]
        r.n_recent = len(rec)
        for j, tok in enumerate(rec):
            r.recent[j] = int(tok)
        self.rows[idx * self.row_bytes:(idx + 1) * self.row_bytes].copy_(host)
        self.states[idx] = torch.tensor([seed & 0x7FFFFFFFFFFFFFFF, 0], dtype=torch.int64)
        return idx

    def close(self, idx: int) -> None:
        self._free.append(idx)

    def sample(self, logits: torch.Tensor, row_index: torch.Tensor, out_ids: torch.Tensor, pdl: bool = False) -> torch.Tensor:
        """logits [T, V] fp32, row_index [T] int32 (device) -> out_ids [T] int64; asynchronous, graph-capturable."""
        T, V = logits.shape
        assert V == self.vocab and T <= self.probs.shape[0] and row_index.dtype == torch.int32
        rc = native.lib().b200_sample_stateful(logits.data_ptr(), logits.stride(0), T, V, self.rows.data_ptr(),
                                               self.states.data_ptr(), row_index.data_ptr(), self.probs.data_ptr(),
                                               out_ids.data_ptr(), 1 if pdl else 0, native.current_stream_ptr())
        native.check(rc, "b200_sample_stateful")
        return out_ids

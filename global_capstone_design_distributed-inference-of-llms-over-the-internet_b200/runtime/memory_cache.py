"""Byte-budget allocator for attention caches with blocking admission.

Upstream Petals' ``MemoryCache`` (petals/server/memory_cache.py:26-225) hands out integer
handles against a global byte budget shared by all connection-handler processes; a request
that does not fit waits (up to a timeout) for other sessions to free memory and fails with
``AllocationFailed`` otherwise, and tensors are materialised lazily in the runtime process.

Here there is one process per GPU, so handles map directly to device tensors, but the contract
is the same: ``with cache.allocate_cache(descr, ..., timeout=t) as handles:`` blocks while the
budget is exhausted, accounts ``enqueued_size`` for waiters, frees on exit.

:class:`~petals_b200.runtime.kv_cache.KVCacheManager` allocates every session cache through one of
these (budget = ``attn_cache_tokens`` x bytes per token), so a request that arrives while the budget
is taken WAITS up to ``alloc_timeout`` for another session to close or expire, as upstream's
``_allocate_cache`` does (petals/server/handler.py:532-547).
"""
from __future__ import annotations

import contextlib
import itertools
import threading
import time
from dataclasses import dataclass
from typing import Dict, Iterator, Optional, Sequence, Tuple

import torch

Handle = int


class AllocationFailed(RuntimeError):
    """The cache budget cannot hold the request (petals/server/memory_cache.py:26)."""


@dataclass(frozen=True)
class TensorDescriptor:
    shape: Tuple[int, ...]
    dtype: torch.dtype
    device: torch.device

    @property
    def nbytes(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n * torch.empty((), dtype=self.dtype).element_size()


class MemoryCache:
    def __init__(self, max_size_bytes: Optional[int], alloc_timeout: float = 10.0):
        self.max_size_bytes = max_size_bytes if max_size_bytes is not None else 2 ** 64 - 1
        self.alloc_timeout = alloc_timeout
        self._used = 0
        self._enqueued = 0
        self._cv = threading.Condition()
        self._tensors: Dict[Handle, torch.Tensor] = {}
        self._sizes: Dict[Handle, int] = {}
        self._ids = itertools.count(1)

    @property
    def current_size_bytes(self) -> int:
        return self._used

    @property
    def enqueued_size_bytes(self) -> int:
        return self._enqueued

    @property
    def bytes_left(self) -> int:
        return self.max_size_bytes - self._used

    @contextlib.contextmanager
    def allocate_cache(self, *descriptors: TensorDescriptor, timeout: Optional[float] = None) -> Iterator[Sequence[Handle]]:
        need = sum(d.nbytes for d in descriptors)
        if need > self.max_size_bytes:
            raise AllocationFailed(f"requested {need} B but the cache holds at most {self.max_size_bytes} B")
        timeout = self.alloc_timeout if timeout is None else timeout
        deadline = time.monotonic() + timeout
        with self._cv:
            self._enqueued += need
            try:
                while self._used + need > self.max_size_bytes:
                    left = deadline - time.monotonic()
                    if left <= 0 or not self._cv.wait(left):
                        if self._used + need > self.max_size_bytes:
                            raise AllocationFailed(f"could not allocate {need} B in {timeout} s "
                                                   f"({self.bytes_left} B free, {self._enqueued - need} B queued ahead)")
                self._used += need
            finally:
                self._enqueued -= need
            handles = []
            for d in descriptors:
                h = next(self._ids)
                self._sizes[h] = d.nbytes
                handles.append(h)
        try:
            for h, d in zip(handles, descriptors):     # materialise outside the lock
                self._tensors[h] = torch.zeros(d.shape, dtype=d.dtype, device=d.device)
            yield handles
        finally:
            with self._cv:
                for h in handles:
                    self._tensors.pop(h, None)
                    self._used -= self._sizes.pop(h, 0)
                self._cv.notify_all()

    def use_cache(self, *handles: Handle) -> Sequence[torch.Tensor]:
        return [self._tensors[h] for h in handles]

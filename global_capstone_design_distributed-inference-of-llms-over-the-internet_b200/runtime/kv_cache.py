"""Per-session KV cache: pre-allocated to ``max_length`` and appended in place.

The reference keeps ``session_id -> tuple((k, v), ...)`` and grows every tensor
with ``torch.cat`` on every token, never evicts and never enforces ``max_length``
(src/rpc_handler.py:70,185,266; petals/llama/block.py:123-128).
Here a session owns ``[n_layers, 2, n_kv, max_len, head_dim]`` from its first
request; the manager enforces a token budget (upstream ``attn_cache_tokens``,
petals/server/server.py:204-209), evicts idle sessions after a TTL, supports
explicit close, and rollback to an earlier position (``start_from_position``,
petals/server/block_functions.py:163-168).
"""
from __future__ import annotations

import contextlib
import threading
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, Optional

import torch

from .memory_cache import AllocationFailed, MemoryCache, TensorDescriptor  # noqa: F401  (AllocationFailed re-exported)


@dataclass
class SessionKV:
    session_id: str
    data: torch.Tensor            # [n_layers, 2, n_kv, max_len, head_dim]
    max_len: int
    length: int = 0               # tokens cached so far
    last_used: float = field(default_factory=time.monotonic)
    limit: Optional[int] = None   # request-level max_length (<= max_len), enforced by the handler
    release: Optional[Callable[[], None]] = None   # returns the bytes to the manager's MemoryCache

    @property
    def n_layers(self) -> int:
        return self.data.shape[0]

    def k(self, layer: int) -> torch.Tensor:
        return self.data[layer, 0]

    def v(self, layer: int) -> torch.Tensor:
        return self.data[layer, 1]

    def rollback(self, position: int) -> None:
        if not 0 <= position <= self.length:
            raise ValueError(f"cannot roll session back to {position} (length {self.length})")
        self.length = position

    def touch(self) -> None:
        self.last_used = time.monotonic()


class KVCacheManager:
    def __init__(self, n_layers: int, n_kv_heads: int, head_dim: int, dtype: torch.dtype, device,
                 max_tokens: Optional[int] = None, default_max_length: int = 2048, session_ttl: float = 600.0,
                 alloc_timeout: float = 0.0):
        self.n_layers, self.n_kv, self.head_dim = n_layers, n_kv_heads, head_dim
        self.dtype, self.device = dtype, torch.device(device)
        self.max_tokens = max_tokens              # None = unlimited
        self.default_max_length = default_max_length
        self.session_ttl = session_ttl
        self.alloc_timeout = alloc_timeout        # how long a request may wait for cache bytes (0 = fail at once)
        self.sessions: Dict[str, SessionKV] = {}
        self._lock = threading.Lock()
        # every session buffer is carved from this byte budget; waiters are woken when a session closes
        self.cache = MemoryCache(None if max_tokens is None else max_tokens * self.bytes_per_token(), alloc_timeout)

    # ---- accounting
    @property
    def tokens_used(self) -> int:
        return sum(s.max_len for s in self.sessions.values())

    @property
    def tokens_left(self) -> Optional[int]:
        return None if self.max_tokens is None else self.max_tokens - self.tokens_used

    def bytes_per_token(self) -> int:
        return 2 * self.n_layers * self.n_kv * self.head_dim * torch.empty((), dtype=self.dtype).element_size()

    # ---- lifecycle
    def get(self, session_id: str) -> Optional[SessionKV]:
        s = self.sessions.get(session_id)
        if s is not None:
            s.touch()
        return s

    def _drop(self, session_id: str) -> bool:
        s = self.sessions.pop(session_id, None)
        if s is None:
            return False
        if s.release is not None:
            s.release()
        return True

    def allocate(self, session_id: str, max_length: Optional[int] = None, timeout: Optional[float] = None) -> SessionKV:
        """A zeroed cache of ``max_length`` tokens for ``session_id`` (an existing, large enough one is reset and
        reused).  When the token budget is taken the call first evicts expired / long-idle sessions and then WAITS
        up to ``timeout`` (default: the manager's ``alloc_timeout``) for other sessions to be closed;
        ``AllocationFailed`` after that."""
        max_length = int(max_length or self.default_max_length)
        with self._lock:
            self.evict_expired()
            old = self.sessions.get(session_id)
            if old is not None and old.max_len >= max_length:
                old.length = 0
                old.touch()
                return old
            self._drop(session_id)
            if self.max_tokens is not None and self.tokens_used + max_length > self.max_tokens:
                self._evict_lru_until(max_length)
        desc = TensorDescriptor((self.n_layers, 2, self.n_kv, max_length, self.head_dim), self.dtype, self.device)
        stack = contextlib.ExitStack()
        try:
            handles = stack.enter_context(self.cache.allocate_cache(desc, timeout=timeout))   # may block (no lock held)
        except AllocationFailed as e:
            raise AllocationFailed(f"KV budget exhausted: need {max_length} tokens, {self.tokens_left} left of "
                                   f"{self.max_tokens} ({e})") from None
        data = self.cache.use_cache(*handles)[0]
        s = SessionKV(session_id, data, max_length, release=stack.close)
        with self._lock:
            self._drop(session_id)          # a concurrent allocate for the same id loses
            self.sessions[session_id] = s
        return s

    def reorder(self, session_ids, hypo_ids) -> None:
        """Beam-search bookkeeping: after the call, session ``session_ids[i]`` continues the hypothesis
        that ``session_ids[hypo_ids[i]]`` held (upstream reorders cache rows by ``hypo_ids`` inside one
        batched cache, petals/server/backend.py:154-158; sessions here are separate buffers, so the
        surviving prefixes are copied, cached length included)."""
        sess = [self.sessions[s] for s in session_ids]
        src_idx = [int(h) for h in hypo_ids]
        if len(src_idx) != len(sess) or any(not 0 <= h < len(sess) for h in src_idx):
            raise ValueError("hypo_ids must index into session_ids")
        need = {h for i, h in enumerate(src_idx) if h != i}
        snap = {h: (sess[h].data[:, :, :, : sess[h].length].clone(), sess[h].length) for h in need}
        for i, h in enumerate(src_idx):
            if h == i:
                continue
            data, length = snap[h]
            if length > sess[i].max_len:
                raise ValueError("destination session is too small for the source hypothesis")
            sess[i].data[:, :, :, :length].copy_(data)
            sess[i].length = length
            sess[i].touch()

    def close(self, session_id: str) -> bool:
        with self._lock:
            return self._drop(session_id)

    def evict_expired(self, now: Optional[float] = None) -> int:
        now = time.monotonic() if now is None else now
        dead = [k for k, s in self.sessions.items() if now - s.last_used > self.session_ttl]
        for k in dead:
            self._drop(k)
        return len(dead)

    def _evict_lru_until(self, need: int) -> None:
        # only idle sessions (unused for > 10% of the TTL) are fair game
        now = time.monotonic()
        for k, s in sorted(self.sessions.items(), key=lambda kv: kv[1].last_used):
            if self.tokens_used + need <= self.max_tokens:
                break
            if now - s.last_used > 0.1 * self.session_ttl:
                self._drop(k)

"""Priority task pool: one compute thread drains (priority, submit-time)-ordered tasks.

Upstream Petals gives each backend three pools (inference / forward / backward) whose tasks
are ordered by a prioritizer and executed by one Runtime thread
(petals/server/task_pool.py:17-177, task_prioritizer.py:6-20: inference 1.0 beats
forward/backward 2.0); different-length requests are never batched together.  Here a single
pool per GPU plays both roles: RPC handlers submit closures with a priority, the worker runs
them one at a time on the device stream, and a task larger than ``max_task_tokens`` is
rejected up front (task_pool.py:103-105).
"""
from __future__ import annotations

import concurrent.futures as cf
import heapq
import itertools
import threading
import time
from abc import ABC, abstractmethod
from typing import Any, Callable, Optional

PRIORITY_INFERENCE = 1.0
PRIORITY_FORWARD = 2.0


class TaskPrioritizerBase(ABC):
    @abstractmethod
    def prioritize(self, kind: str, *, points: float = 0.0, **kwargs) -> float:
        """Smaller runs first."""


class DummyTaskPrioritizer(TaskPrioritizerBase):
    def prioritize(self, kind: str, *, points: float = 0.0, **kwargs) -> float:
        return PRIORITY_INFERENCE if kind == "inference" else PRIORITY_FORWARD


class PrioritizedTaskPool:
    def __init__(self, name: str = "pool", max_task_tokens: Optional[int] = None,
                 prioritizer: Optional[TaskPrioritizerBase] = None):
        self.name = name
        self.max_task_tokens = max_task_tokens
        self.prioritizer = prioritizer or DummyTaskPrioritizer()
        self._heap: list = []
        self._seq = itertools.count()
        self._cv = threading.Condition()
        self._stop = False
        self.completed = 0
        self._thread = threading.Thread(target=self._run, name=f"{name}-runtime", daemon=True)
        self._thread.start()

    def submit(self, fn: Callable[..., Any], *args, kind: str = "inference", tokens: int = 0,
               priority: Optional[float] = None, **kwargs) -> "cf.Future":
        if self.max_task_tokens is not None and tokens > self.max_task_tokens:
            raise ValueError(f"task of {tokens} tokens exceeds max_task_tokens={self.max_task_tokens}")
        fut: cf.Future = cf.Future()
        pr = self.prioritizer.prioritize(kind) if priority is None else priority
        with self._cv:
            if self._stop:
                raise RuntimeError("pool is shut down")
            heapq.heappush(self._heap, (pr, time.monotonic(), next(self._seq), fn, args, kwargs, fut))
            self._cv.notify()
        return fut

    def _run(self) -> None:
        while True:
            with self._cv:
                while not self._heap and not self._stop:
                    self._cv.wait()
                if self._stop and not self._heap:
                    return
                _, _, _, fn, args, kwargs, fut = heapq.heappop(self._heap)
            if not fut.set_running_or_notify_cancel():
                continue
            try:
                fut.set_result(fn(*args, **kwargs))
            except BaseException as e:  # noqa: BLE001 - delivered through the future
                fut.set_exception(e)
            self.completed += 1

    @property
    def queued(self) -> int:
        with self._cv:
            return len(self._heap)

    def shutdown(self) -> None:
        with self._cv:
            self._stop = True
            for item in self._heap:
                item[-1].cancel()
            self._heap.clear()
            self._cv.notify_all()

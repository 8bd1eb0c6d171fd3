"""Priority task pool: one compute thread drains (priority, submit-time)-ordered tasks.

Upstream Petals gives each backend three pools (inference / forward / backward) whose tasks
are ordered by a prioritizer and executed by one Runtime thread
(petals/server/task_pool.py:17-177, task_prioritizer.py:6-20: inference 1.0 beats
forward/backward 2.0); different-length requests are never batched together.  Here a single
pool per GPU plays both roles: RPC handlers submit closures with a priority, the worker runs
them one at a time on the device stream, and a task larger than ``max_task_tokens`` is
rejected up front (task_pool.py:103-105).

Batching: upstream's pools merge queued tasks of the SAME shape into one batch
(task_pool.py:117-131).  Here single-token decode steps of different sessions are the one
case worth merging - the decode kernels take up to 32 sessions per launch sequence at little
more than the cost of one - so ``submit_batch_item`` queues an item for a batch function and
the worker, when it reaches such an item, takes every other queued item of the same function
along (up to its ``max_batch``) and runs them in one call.  Clients that step in lockstep send their
requests within a few hundred microseconds of each other, so the worker may *linger*: when it reaches a
batch item and fewer items are queued than the caller expects (``expect()``: the number of sessions
that have been stepping lately), it waits up to ``batch_linger`` seconds for the rest before it runs
the step - one short wait instead of a step for the early bird and a second step for everybody else.
The wait ends early when a quarter of the linger passes without a new arrival (the missing sessions
are not on their way).  Measured on Llama-2-7B, 16 client processes against one B200 stage: greedy
batching settles into two alternating groups of 8 (1694 tok/s), waiting for the expected 16 gives one
batch per round (2478 tok/s).  ``linger_max_missing`` can restrict the wait to "few items missing".
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


def _BATCH():  # marker in the `fn` slot of a heap entry: args = (batch_fn, item, max_batch)
    raise RuntimeError("batch marker is not callable")


class TaskPrioritizerBase(ABC):
    @abstractmethod
    def prioritize(self, kind: str, *, points: float = 0.0, **kwargs) -> float:
        """Smaller runs first."""


class DummyTaskPrioritizer(TaskPrioritizerBase):
    def prioritize(self, kind: str, *, points: float = 0.0, **kwargs) -> float:
        return PRIORITY_INFERENCE if kind == "inference" else PRIORITY_FORWARD


class PrioritizedTaskPool:
    def __init__(self, name: str = "pool", max_task_tokens: Optional[int] = None,
                 prioritizer: Optional[TaskPrioritizerBase] = None, batch_linger: float = 0.0,
                 expect: Optional[Callable[[], int]] = None, linger_max_missing: int = 1 << 30,
                 thread_init: Optional[Callable[[], None]] = None):
        self.name = name
        self.batch_linger = float(batch_linger)
        self.expect = expect
        self.linger_max_missing = int(linger_max_missing)
        self.thread_init = thread_init          # runs first on the worker thread (e.g. select the stage's GPU)
        self.max_task_tokens = max_task_tokens
        self.prioritizer = prioritizer or DummyTaskPrioritizer()
        self._heap: list = []
        self._seq = itertools.count()
        self._cv = threading.Condition()
        self._stop = False
        self.completed = 0
        self.batches = 0           # batch calls made / items they carried (rpc_info reports the mean batch)
        self.batched_items = 0
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

    def submit_batch_item(self, batch_fn: Callable[[list], list], item: Any, *, kind: str = "inference",
                          tokens: int = 1, max_batch: int = 32, priority: Optional[float] = None) -> "cf.Future":
        """Queue ``item`` for ``batch_fn``; the worker calls ``batch_fn([item, ...])`` with every queued item of
        the same function (submit order, at most ``max_batch``) and resolves each future with the matching entry
        of the returned list (an entry that is an exception instance fails only that item)."""
        if self.max_task_tokens is not None and tokens > self.max_task_tokens:
            raise ValueError(f"task of {tokens} tokens exceeds max_task_tokens={self.max_task_tokens}")
        fut: cf.Future = cf.Future()
        pr = self.prioritizer.prioritize(kind) if priority is None else priority
        with self._cv:
            if self._stop:
                raise RuntimeError("pool is shut down")
            heapq.heappush(self._heap, (pr, time.monotonic(), next(self._seq), _BATCH, (batch_fn, item, max_batch), {}, fut))
            self._cv.notify()
        return fut

    def _take_batch(self, first) -> list:
        """Called with the lock held: `first` plus the queued batch items of the same function."""
        batch_fn, _, max_batch = first[4]
        group = [first]
        if max_batch > 1:
            rest = []
            for ent in sorted(self._heap):
                if len(group) < max_batch and ent[3] is _BATCH and ent[4][0] == batch_fn:
                    group.append(ent)
                else:
                    rest.append(ent)
            if len(group) > 1:
                self._heap[:] = rest
                heapq.heapify(self._heap)
        return group

    def _linger(self, first) -> None:
        """Lock held: wait (releasing the lock) until as many items of ``first``'s function are queued as the caller
        expects - at most ``batch_linger`` seconds in total, and only while items keep coming (a quiet gap of a
        quarter of the linger ends the wait: the missing sessions are in the other phase, not on their way)."""
        batch_fn, _, max_batch = first[4]
        deadline = time.monotonic() + self.batch_linger
        gap = self.batch_linger / 4
        seen = -1
        while not self._stop:
            want = min(int(self.expect()), max_batch)
            have = 1 + sum(1 for e in self._heap if e[3] is _BATCH and e[4][0] == batch_fn)
            left = deadline - time.monotonic()
            if have >= want or left <= 0 or want - have > self.linger_max_missing:
                return
            if have == seen:
                return                              # a whole gap without a new arrival
            seen = have
            self._cv.wait(min(left, gap))

    def _run(self) -> None:
        if self.thread_init is not None:
            self.thread_init()
        while True:
            with self._cv:
                while not self._heap and not self._stop:
                    self._cv.wait()
                if self._stop and not self._heap:
                    return
                ent = heapq.heappop(self._heap)
                if ent[3] is _BATCH and self.batch_linger > 0.0 and self.expect is not None:
                    self._linger(ent)
                group = self._take_batch(ent) if ent[3] is _BATCH else None
            if group is not None:
                self._run_batch(group)
                continue
            _, _, _, fn, args, kwargs, fut = ent
            if not fut.set_running_or_notify_cancel():
                continue
            try:
                fut.set_result(fn(*args, **kwargs))
            except BaseException as e:  # noqa: BLE001 - delivered through the future
                fut.set_exception(e)
            self.completed += 1

    def _run_batch(self, group: list) -> None:
        live = [e for e in group if e[6].set_running_or_notify_cancel()]
        if not live:
            return
        batch_fn = live[0][4][0]
        try:
            results = batch_fn([e[4][1] for e in live])
            if len(results) != len(live):
                raise RuntimeError(f"batch function returned {len(results)} results for {len(live)} items")
        except BaseException as e:  # noqa: BLE001
            results = [e] * len(live)
        for ent, r in zip(live, results):
            if isinstance(r, BaseException):
                ent[6].set_exception(r)
            else:
                ent[6].set_result(r)
        self.completed += len(live)
        self.batches += 1
        self.batched_items += len(live)

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

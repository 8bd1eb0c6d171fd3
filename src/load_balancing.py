"""Throughput-weighted block placement (Petals paper, Appendix D) for replica spans.

Behavioural contract (checked against the reference's functions on seeded swarms in
tests/test_load_balancing.py; reference: src/load_balancing.py:17-366):

* a server's span = its contiguous run of announced blocks; span throughput = the minimum
  its blocks announce; per-block swarm throughput = the sum over the spans covering it
  (spans visited in ``str(peer_id)`` order so every server computes the same floats);
* rule 1 (join): serve the ``num_blocks``-wide window with the lexicographically smallest
  ``(window min, window mean, start)``, never starting below ``min_block`` (the blocks the
  client runs itself);
* rule 2 (re-balance): pretend to leave, pick the best window again, then let every other
  server greedily re-place itself (random order, at most 10 sweeps); move only if the swarm
  bottleneck would improve by more than ``1 - balance_quality``.

Deliberate deviations from the reference: the bottleneck is taken over the
blocks servers can actually cover, ``[min_block, total_blocks)`` — the reference takes it over
all blocks, which makes it 0 whenever the client keeps a prefix and disables rule 2 entirely —
and OFFLINE servers are ignored when building spans.  With ``min_block == 0`` the decisions are
identical to the reference.
"""
from __future__ import annotations

import logging
from dataclasses import dataclass
from enum import Enum
from typing import Dict, Hashable, List, Optional, Sequence

import numpy as np

logger = logging.getLogger(__name__)

EPS = 1e-3
MAX_REBALANCE_SWEEPS = 10


class ServerState(Enum):
    JOINING = "joining"
    ONLINE = "online"
    OFFLINE = "offline"


@dataclass
class ServerInfo:
    peer_id: Hashable
    state: ServerState
    throughput: float
    start_block: int
    end_block: int
    server_address: Optional[str] = None

    @property
    def num_blocks(self) -> int:
        return self.end_block - self.start_block


@dataclass
class RemoteModuleInfo:
    uid: str                              # "block_<i>"
    server_info: Optional[ServerInfo] = None


@dataclass
class RemoteSpanInfo:
    peer_id: Hashable
    start: int
    end: int
    length: int
    throughput: float

    def __post_init__(self):
        self.length = self.end - self.start

    def move_to(self, new_start: int) -> None:
        self.start, self.end = new_start, new_start + self.length


def _block_index(uid: str) -> Optional[int]:
    try:
        return int(uid.rsplit("_", 1)[-1])
    except ValueError:
        return None


def compute_spans(module_infos: Sequence[RemoteModuleInfo],
                  min_state: ServerState = ServerState.JOINING) -> Dict[Hashable, RemoteSpanInfo]:
    """One span per peer: its last contiguous run of announced blocks."""
    allowed = {ServerState.JOINING, ServerState.ONLINE} if min_state == ServerState.JOINING else {ServerState.ONLINE}
    per_peer: Dict[Hashable, Dict[int, float]] = {}
    for info in module_infos:
        srv = info.server_info
        if srv is None or srv.state not in allowed:
            continue
        idx = _block_index(info.uid)
        if idx is None:
            logger.warning("cannot parse block index from uid %r", info.uid)
            continue
        per_peer.setdefault(srv.peer_id, {})[idx] = float(srv.throughput)
    spans: Dict[Hashable, RemoteSpanInfo] = {}
    for peer, blocks in per_peer.items():
        order = sorted(blocks)
        run_start = order[0]
        run_min = blocks[order[0]]
        prev = order[0]
        for b in order[1:]:
            if b == prev + 1:
                run_min = min(run_min, blocks[b])
            else:
                run_start, run_min = b, blocks[b]
            prev = b
        spans[peer] = RemoteSpanInfo(peer, run_start, prev + 1, prev + 1 - run_start, run_min)
    return spans


def compute_throughputs(spans: Dict[Hashable, RemoteSpanInfo], total_blocks: int) -> np.ndarray:
    out = np.zeros(total_blocks)
    for span in sorted(spans.values(), key=lambda s: str(s.peer_id)):
        out[span.start: span.end] += span.throughput
    return out


def _choose_best_start(throughputs: np.ndarray, num_blocks: int, min_block: int = 0) -> int:
    n = len(throughputs)
    if n < num_blocks:
        return max(0, int(min_block))
    last = n - num_blocks
    lo = int(max(0, min(min_block, last)))
    win = np.lib.stride_tricks.sliding_window_view(throughputs, num_blocks)[lo: last + 1]
    mins, means = win.min(axis=1), win.mean(axis=1)
    # lexicographic (min, mean, start): lexsort's LAST key is the primary one
    best = np.lexsort((np.arange(len(mins)), means, mins))[0]
    return lo + int(best)


def _infer_total_blocks(module_infos: Sequence[RemoteModuleInfo], fallback: int) -> int:
    top = max((i for i in (_block_index(m.uid) for m in module_infos) if i is not None), default=0)
    return top + 1 if top > 0 else fallback


def choose_best_blocks(num_blocks: int, module_infos: Sequence[RemoteModuleInfo],
                       total_blocks: Optional[int] = None, min_block: int = 0) -> List[int]:
    """Rule 1: where should a joining server with room for ``num_blocks`` blocks go?"""
    if total_blocks is None:
        total_blocks = _infer_total_blocks(module_infos, num_blocks)
    thr = compute_throughputs(compute_spans(module_infos), total_blocks)
    start = _choose_best_start(thr, num_blocks, min_block)
    return list(range(start, start + num_blocks))


def swarm_bottleneck(throughputs: np.ndarray, min_block: int = 0) -> float:
    seg = throughputs[min(int(min_block), max(len(throughputs) - 1, 0)):]
    return float(seg.min()) if len(seg) else 0.0


def should_choose_other_blocks(local_peer_id: Hashable, module_infos: Sequence[RemoteModuleInfo],
                               balance_quality: float = 0.75, total_blocks: Optional[int] = None,
                               min_block: int = 0, rng=None) -> bool:
    """Rule 2: would the swarm bottleneck improve enough if this server moved?"""
    if balance_quality > 1.0:
        return True                      # forced re-placement (debug switch, as in the reference)
    if total_blocks is None:
        total_blocks = _infer_total_blocks(module_infos, 32)
    spans = compute_spans(module_infos)
    thr = compute_throughputs(spans, total_blocks)
    initial = swarm_bottleneck(thr, min_block)

    me = next((pid for pid in spans if str(pid) == str(local_peer_id)), None)
    if me is None:
        logger.warning("local peer %s is not announced in the swarm (%d spans)", str(local_peer_id)[:16], len(spans))
        return False
    mine = spans[me]
    lo, hi = max(0, min(mine.start, total_blocks - 1)), min(mine.end, total_blocks)
    if hi > lo:
        thr[lo:hi] -= mine.throughput * (1 + EPS)
    if initial > EPS and swarm_bottleneck(thr, min_block) <= 0:
        return False                     # leaving would disconnect the chain
    target = _choose_best_start(thr, mine.length, min_block)
    if target == mine.start:
        return False
    thr[mine.start: mine.end] += mine.throughput * EPS
    mine.move_to(target)
    thr[mine.start: mine.end] += mine.throughput

    shuffle = (rng.shuffle if rng is not None else np.random.shuffle)
    for _ in range(MAX_REBALANCE_SWEEPS):
        order = list(spans.keys())
        shuffle(order)
        moved = False
        for pid in order:
            sp = spans[pid]
            thr[sp.start: sp.end] -= sp.throughput * (1 + EPS)
            cand = _choose_best_start(thr, sp.length, min_block)
            thr[sp.start: sp.end] += sp.throughput * EPS
            if cand != sp.start:
                sp.move_to(cand)
                moved = True
            thr[sp.start: sp.end] += sp.throughput
        if not moved:
            break
    new = swarm_bottleneck(thr, min_block)
    if new < initial or new < EPS:
        return False
    quality = initial / new
    logger.info("Swarm balance quality: %.1f%% (initial=%.2f, new=%.2f)", quality * 100, initial, new)
    return quality < balance_quality - EPS

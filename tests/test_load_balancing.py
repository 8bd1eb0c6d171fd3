"""Placement decisions: golden values, parity with the reference's functions on seeded
swarms (min_block=0), and the B13 fix on shifted swarms."""
import random

import numpy as np
import pytest

from src import load_balancing as lb
from tests.refstub import load_reference_module


def swarm(spans, mod=lb):
    infos = []
    for pid, (s, e, t) in spans.items():
        si = mod.ServerInfo(pid, mod.ServerState.ONLINE, t, s, e)
        infos += [mod.RemoteModuleInfo(f"block_{b}", si) for b in range(s, e)]
    return infos


def test_golden_join_placements():
    assert lb.choose_best_blocks(8, swarm({"a": (8, 16, 10)}), 32, 8)[0] == 16
    assert lb.choose_best_blocks(8, swarm({"a": (8, 16, 10), "b": (16, 24, 10)}), 32, 8)[0] == 24
    full = {"a": (8, 16, 10), "b": (16, 24, 10), "c": (24, 32, 10)}
    assert lb.choose_best_blocks(8, swarm(full), 32, 8)[0] == 8            # lowest start on ties
    full["b"] = (16, 24, 2)
    assert lb.choose_best_blocks(8, swarm(full), 32, 8)[0] == 16           # weakest window
    sp = lb.compute_spans(swarm({"a": (8, 16, 10), "a2": (8, 16, 10), "a3": (8, 16, 10), "b": (16, 24, 10),
                                 "c": (24, 32, 10)}))
    assert lb.compute_throughputs(sp, 32).tolist() == [0] * 8 + [30] * 8 + [10] * 16


def test_spans_contiguity_and_min_throughput():
    si = lb.ServerInfo("p", lb.ServerState.ONLINE, 5.0, 0, 4)
    si2 = lb.ServerInfo("p", lb.ServerState.ONLINE, 3.0, 0, 4)
    infos = [lb.RemoteModuleInfo("block_0", si), lb.RemoteModuleInfo("block_1", si2), lb.RemoteModuleInfo("block_2", si),
             lb.RemoteModuleInfo("block_9", si)]
    sp = lb.compute_spans(infos)["p"]
    assert (sp.start, sp.end, sp.length) == (9, 10, 1)      # the last contiguous run wins, like the reference
    off = lb.ServerInfo("q", lb.ServerState.OFFLINE, 5.0, 0, 2)
    assert "q" not in lb.compute_spans([lb.RemoteModuleInfo("block_0", off)])


def _random_swarm(rng, total, n_servers, mod):
    spans = {}
    for i in range(n_servers):
        length = rng.choice([2, 4, 8])
        start = rng.randrange(0, total - length + 1)
        spans[f"peer{i:02d}"] = (start, start + length, rng.choice([1.0, 2.5, 10.0, 30.0]))
    return spans


@pytest.mark.parametrize("seed", range(12))
def test_parity_with_reference_functions(seed):
    ref = load_reference_module("load_balancing")
    if ref is None:
        pytest.skip("reference sources not available")
    rng = random.Random(seed)
    total = 24
    spans = _random_swarm(rng, total, rng.randrange(2, 7), lb)
    ours, theirs = swarm(spans, lb), swarm(spans, ref)
    a = lb.compute_throughputs(lb.compute_spans(ours), total)
    b = ref.compute_throughputs(ref.compute_spans(theirs), total)
    assert np.allclose(a, b)
    for nb in (2, 4, 8):
        assert lb.choose_best_blocks(nb, ours, total, 0) == ref.choose_best_blocks(nb, theirs, total, 0)
        assert lb.choose_best_blocks(nb, ours, total, 4) == ref.choose_best_blocks(nb, theirs, total, 4)
    for pid in spans:
        np.random.seed(seed)
        x = lb.should_choose_other_blocks(pid, swarm(spans, lb), 0.75, total, 0)
        np.random.seed(seed)
        y = ref.should_choose_other_blocks(pid, swarm(spans, ref), 0.75, total, 0)
        assert x == y, f"rebalance decision differs for {pid}"


def test_b13_rebalance_fires_with_client_prefix():
    bad = {"a": (8, 16, 10), "a2": (8, 16, 10), "a3": (8, 16, 10), "b": (16, 24, 10), "c": (24, 32, 1)}
    np.random.seed(0)
    assert lb.should_choose_other_blocks("a", swarm(bad), 0.75, 32, 8) is True
    ref = load_reference_module("load_balancing")
    if ref is not None:
        np.random.seed(0)
        assert ref.should_choose_other_blocks("a", swarm(bad, ref), 0.75, 32, 8) is False   # documents B13
    balanced = {"a": (8, 16, 10), "b": (16, 24, 10), "c": (24, 32, 10)}
    assert lb.should_choose_other_blocks("a", swarm(balanced), 0.75, 32, 8) is False
    assert lb.should_choose_other_blocks("zzz", swarm(balanced), 0.75, 32, 8) is False      # unknown peer
    assert lb.should_choose_other_blocks("a", swarm(balanced), 1.5, 32, 8) is True          # forced

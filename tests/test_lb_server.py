"""Load-balanced servers: join-time placement over a live DHT, module routing by the client,
re-placement when the swarm is imbalanced."""
import threading

import torch

from tests.test_cpu_pipeline import _args


def test_lb_servers_cover_the_model_and_client_routes_by_module():
    from src.main import parse_splits, run_rank0, run_stage_server_with_load_balancing

    model, splits = "tiny-llama", "1"            # client keeps block 0; servers cover [1, 4)
    cuts = parse_splits(splits)
    dev = torch.device("cpu")
    handles = []
    try:
        a = _args(model, splits, 1, "", seed=4, use_load_balancing=True, num_blocks=2, total_blocks=4,
                  network_bandwidth_mbps=1e6)
        h1 = run_stage_server_with_load_balancing(a, dev, cuts, block=False)
        handles.append(h1)
        peers = h1.dht.get_visible_maddrs()[0]
        assert h1.blocks == [1, 2]                                       # empty swarm: lowest start >= min_block
        b = _args(model, splits, 1, peers, seed=4, use_load_balancing=True, num_blocks=2, total_blocks=4,
                  network_bandwidth_mbps=1e6)
        h2 = run_stage_server_with_load_balancing(b, dev, cuts, block=False)
        handles.append(h2)
        assert h2.blocks == [2, 3]                                       # weakest window includes uncovered block 3
        # [1,2) + ... : block 1 only by h1, 2 by both, 3 only by h2 -> route h1[1,3) then needs a server starting at 3
        c = _args(model, splits, 1, peers, seed=4, use_load_balancing=True, num_blocks=1, total_blocks=4,
                  network_bandwidth_mbps=1e6)
        h3 = run_stage_server_with_load_balancing(c, dev, cuts, block=False)
        handles.append(h3)
        assert h3.blocks in ([1], [3])                                   # tie between the single-covered blocks
        if h3.blocks == [1]:
            d = _args(model, splits, 1, peers, seed=4, use_load_balancing=True, num_blocks=1, total_blocks=4,
                      network_bandwidth_mbps=1e6)
            h4 = run_stage_server_with_load_balancing(d, dev, cuts, block=False)
            handles.append(h4)
            assert h4.blocks == [3]
        cargs = _args(model, splits, 0, peers, seed=4, use_load_balancing=True, total_blocks=4, max_new_tokens=6)
        out, stats = run_rank0(cargs, dev, cuts)
        # oracle: fixed 2-stage pipeline on the same weights
        from src.main import run_stage_server_fixed

        f = run_stage_server_fixed(_args(model, splits, 1, "", seed=4), dev, cuts, block=False)
        handles.append(f)
        ref, _ = run_rank0(_args(model, splits, 0, f.dht.get_visible_maddrs()[0], seed=4, max_new_tokens=6), dev, cuts)
        assert out == ref
    finally:
        for h in handles:
            h.stop()


def test_client_enters_a_span_mid_way_when_windows_overlap():
    """Two servers with overlapping windows [1,3) and [2,4): no server STARTS at block 3, so after [1,3) the client
    asks the second server for blocks [3,4) of its span only (a view with caches of its own).  Same tokens as a
    fixed two-stage pipeline on the same weights, in unary and in streaming mode."""
    from src.main import parse_splits, run_rank0, run_stage_server_fixed, run_stage_server_with_load_balancing

    model, splits = "tiny-llama", "1"
    cuts = parse_splits(splits)
    dev = torch.device("cpu")
    handles = []
    try:
        kw = dict(seed=4, use_load_balancing=True, num_blocks=2, total_blocks=4, network_bandwidth_mbps=1e6)
        h1 = run_stage_server_with_load_balancing(_args(model, splits, 1, "", **kw), dev, cuts, block=False)
        handles.append(h1)
        peers = h1.dht.get_visible_maddrs()[0]
        h2 = run_stage_server_with_load_balancing(_args(model, splits, 1, peers, **kw), dev, cuts, block=False)
        handles.append(h2)
        assert (h1.blocks, h2.blocks) == ([1, 2], [2, 3])
        outs = []
        for stream in (False, True):
            cargs = _args(model, splits, 0, peers, seed=4, use_load_balancing=True, total_blocks=4, max_new_tokens=6,
                          stream=stream)
            outs.append(run_rank0(cargs, dev, cuts)[0])
        assert h2.handler._views, "the second server served a view over part of its span"
        assert list(h2.handler._views) == [(3, 4)]
        f = run_stage_server_fixed(_args(model, splits, 1, "", seed=4), dev, cuts, block=False)
        handles.append(f)
        ref, _ = run_rank0(_args(model, splits, 0, f.dht.get_visible_maddrs()[0], seed=4, max_new_tokens=6), dev, cuts)
        assert outs[0] == ref and outs[1] == ref
        # the sessions' caches were freed in the view's manager too
        assert all(not m.sessions for _, m, _ in h2.handler._views.values())
    finally:
        for h in handles:
            h.stop()

"""Kill-a-stage rerouting with exact continuation (the reference's manual test_fault_tolerance.py
+ kill_stage.py, scripted).  Two replicas serve every stage; mid-generation the replica in use dies
(server stopped => connection refused) or drops one request; the client must exclude it, replay the
journal into the other replica — WITHOUT the in-flight step — and produce exactly the
tokens of an undisturbed run."""
import threading

import pytest
import torch

from tests.test_cpu_pipeline import _args


def _start(model, splits, cuts, stage, peers, seed=11):
    from src.main import run_stage_server_fixed

    return run_stage_server_fixed(_args(model, splits, stage, peers, seed=seed), torch.device("cpu"), cuts, block=False)


@pytest.mark.parametrize("routing_push", [False, True])
def test_stage_kill_and_replay_is_token_exact(routing_push):
    from src.main import parse_splits, run_rank0

    model, splits = "tiny-llama-gqa", "1,3"
    cuts = parse_splits(splits)
    dev = torch.device("cpu")
    handles = []
    try:
        first = _start(model, splits, cuts, 1, "")
        peers = first.dht.get_visible_maddrs()[0]
        handles = [first, _start(model, splits, cuts, 2, peers), _start(model, splits, cuts, 1, peers),
                   _start(model, splits, cuts, 2, peers)]
        n_new = 10
        ref, _ = run_rank0(_args(model, splits, 0, peers, seed=11, max_new_tokens=n_new), dev, cuts)

        # whichever stage-2 replica the next run picks "crashes" while serving its 3rd request
        finals = [h for h in handles if h.handler.final_stage]
        fired = []
        for h in finals:
            h.handler.die_after = h.handler.requests + 3
            h.handler.die_guard = lambda: (not fired) and (fired.append(1) or True)   # only one replica dies
        out, stats = run_rank0(_args(model, splits, 0, peers, seed=11, max_new_tokens=n_new, push=routing_push,
                                     request_timeout=5.0), dev, cuts)
        killed = [h for h in finals if h.handler.die_after is None]
        assert killed, "the fault was never injected"
        assert stats["recoveries"] >= 1
        assert out == ref, "continuation after recovery must be token-exact"
    finally:
        for h in handles:
            try:
                h.stop()
            except Exception:
                pass


def test_dropped_request_is_retried_without_duplicate_token():
    from src.main import parse_splits, run_rank0

    model, splits = "tiny-gpt2", "2"
    cuts = parse_splits(splits)
    dev = torch.device("cpu")
    h1 = _start(model, splits, cuts, 1, "")
    peers = h1.dht.get_visible_maddrs()[0]
    h2 = _start(model, splits, cuts, 1, peers)
    try:
        ref, _ = run_rank0(_args(model, splits, 0, peers, seed=11, max_new_tokens=8), dev, cuts)
        h1.handler.fail_next = 1
        h2.handler.fail_next = 1     # whichever replica is chosen fails once with a server-side error
        out, stats = run_rank0(_args(model, splits, 0, peers, seed=11, max_new_tokens=8), dev, cuts)
        assert out == ref and stats["recoveries"] >= 1
    finally:
        h1.stop()
        h2.stop()


def test_kv_length_mismatch_is_an_error_not_a_warning():
    """The reference only warns when cur_len disagrees with the cached length (B4)."""
    from petals_b200.net.rpc import RpcClient, RpcError
    from src.main import parse_splits

    model, splits = "tiny-gpt2", "2"
    cuts = parse_splits(splits)
    h = _start(model, splits, cuts, 1, "")
    try:
        c = RpcClient()
        x = torch.randn(1, 4, 128)
        md = {"session_id": "s", "seq_len": 4, "cur_len": 4, "is_prefill": True, "max_length": 32, "temperature": 0.0}
        c.call(h.rpc.address, "StageConnectionHandler.rpc_forward", "k", [x], md)
        bad = dict(md, is_prefill=False, seq_len=1, cur_len=9)
        with pytest.raises(RpcError, match="KV length mismatch"):
            c.call(h.rpc.address, "StageConnectionHandler.rpc_forward", "k", [x[:, :1]], bad)
        ok = dict(md, is_prefill=False, seq_len=1, cur_len=5)
        c.call(h.rpc.address, "StageConnectionHandler.rpc_forward", "k", [x[:, :1]], ok)
        over = dict(md, is_prefill=True, seq_len=4, cur_len=4, max_length=3)
        with pytest.raises(RpcError, match="max_length"):
            c.call(h.rpc.address, "StageConnectionHandler.rpc_forward", "k", [x], over)
        _, info = c.call(h.rpc.address, "StageConnectionHandler.rpc_info", "k", [], {})
        assert info["sessions"] >= 1 and info["final_stage"] is True
        _, r = c.call(h.rpc.address, "StageConnectionHandler.rpc_close_session", "k", [], {"session_id": "s"})
        assert r["closed"] is True
        with pytest.raises(RpcError, match="Missing past_key_values"):
            c.call(h.rpc.address, "StageConnectionHandler.rpc_forward", "k", [x[:, :1]], ok)
        c.close()
    finally:
        h.stop()


def test_push_chain_transient_error_downstream_is_recovered():
    """Three server stages, servers push to each other; the LAST one fails one request with a server-side error
    while every server stays alive.  The stages before it have already consumed the step, so the client must
    rebuild the whole suffix (not blame the first hop) and still produce the undisturbed tokens."""
    from src.main import parse_splits, run_rank0

    model, splits = "tiny-llama-gqa", "1,2,3"
    cuts = parse_splits(splits)
    dev = torch.device("cpu")
    handles = []
    try:
        first = _start(model, splits, cuts, 1, "")
        peers = first.dht.get_visible_maddrs()[0]
        handles = [first, _start(model, splits, cuts, 2, peers), _start(model, splits, cuts, 3, peers)]
        ref, _ = run_rank0(_args(model, splits, 0, peers, seed=11, max_new_tokens=8), dev, cuts)
        last = handles[2]
        assert last.handler.final_stage
        base = last.handler.requests

        def arm():                       # fail the 4th request of the next run (a decode step)
            import time
            while last.handler.requests < base + 3:
                time.sleep(0.005)
            last.handler.fail_next = 1

        t = threading.Thread(target=arm, daemon=True)
        t.start()
        out, stats = run_rank0(_args(model, splits, 0, peers, seed=11, max_new_tokens=8, push=True,
                                     request_timeout=5.0), dev, cuts)
        assert stats["recoveries"] >= 1
        assert out == ref
    finally:
        for h in handles:
            try:
                h.stop()
            except Exception:
                pass


def test_registry_survives_the_bootstrap_node():
    """Servers learn of each other through the bootstrap node; when it dies mid-generation the survivors keep
    announcing to each other and the client (which has talked to the swarm) still finds a replacement."""
    from src.main import parse_splits, run_rank0

    model, splits = "tiny-llama-gqa", "2"
    cuts = parse_splits(splits)
    dev = torch.device("cpu")
    handles = []
    try:
        boot = _start(model, splits, cuts, 1, "")
        peers = boot.dht.get_visible_maddrs()[0]
        a = _start(model, splits, cuts, 1, peers)
        # b is bootstrapped from a (not from the bootstrap node): discovery must be transitive
        b = _start(model, splits, cuts, 1, a.dht.get_visible_maddrs()[0])
        handles = [boot, a, b]
        assert len(b.dht.known_peers()) >= 3 and len(boot.dht.known_peers()) >= 3
        ref, _ = run_rank0(_args(model, splits, 0, peers, seed=11, max_new_tokens=8), dev, cuts)
        # the bootstrap node (DHT + stage server) dies while serving its 3rd request, whichever replica is in use
        fired = []
        for h in handles:
            h.handler.die_after = h.handler.requests + 3
            h.handler.die_guard = lambda: (not fired) and (fired.append(1) or True)
        boot.handler.on_die = lambda: boot.stop(announce=False)
        # the client joins through `a`, the registry's bootstrap node goes away under it
        boot_only = boot.dht
        out, stats = run_rank0(_args(model, splits, 0, a.dht.get_visible_maddrs()[0], seed=11, max_new_tokens=8,
                                     request_timeout=5.0), dev, cuts)
        assert out == ref
        boot.stop(announce=False)
        del boot_only
        # with the bootstrap node gone, a fresh client that only knows `b` still sees `a`'s record
        from petals_b200.net.dht import DHT
        from src.dht_utils import get_stage_entries

        c = DHT(start=True, initial_peers=[b.dht.get_visible_maddrs()[0]], client_mode=True)
        ids = {str(e.get("peer_id")) for e in get_stage_entries(c, 1)}
        assert str(a.peer_id) in ids and str(b.peer_id) in ids
        c.shutdown()
    finally:
        for h in handles:
            try:
                h.stop()
            except Exception:
                pass


@pytest.mark.parametrize("push", [False, True])
def test_concurrent_sessions_survive_a_replica_kill(push):
    """Six client sessions step concurrently against batching servers (two replicas per stage); one replica "crashes"
    mid-generation.  Every session - the ones that were on the dead replica and the ones batched next to them on the
    survivors - must end with exactly the tokens it produces alone."""
    import threading

    from src.main import parse_splits, run_rank0

    model, splits = "tiny-llama-gqa", "1,3"
    cuts = parse_splits(splits)
    dev = torch.device("cpu")
    prompts = ["alpha beta", "one two three four", "x", "the quick brown fox", "a b c d e f g", "hello there"]
    handles = []
    try:
        first = _start(model, splits, cuts, 1, "")
        peers = first.dht.get_visible_maddrs()[0]
        handles = [first, _start(model, splits, cuts, 2, peers), _start(model, splits, cuts, 1, peers),
                   _start(model, splits, cuts, 2, peers)]
        n_new = 12

        def client(i):
            return run_rank0(_args(model, splits, 0, peers, seed=11, max_new_tokens=n_new, prompt=prompts[i],
                                   request_timeout=10.0, no_repeat_stop=True, push=push and armed), dev, cuts)

        armed = False
        ref = [client(i)[0] for i in range(len(prompts))]
        armed = True
        fired = []
        for h in handles:
            h.handler.die_after = h.handler.requests + 20          # somewhere in the middle of the concurrent run
            h.handler.die_guard = lambda: (not fired) and (fired.append(1) or True)   # only one server dies
        out, rec = [None] * len(prompts), [0] * len(prompts)

        def run(i):
            out[i], st = client(i)
            rec[i] = st["recoveries"]

        ts = [threading.Thread(target=run, args=(i,)) for i in range(len(prompts))]
        for t in ts:
            t.start()
        for t in ts:
            t.join(120)
        assert fired, "the fault was never injected"
        assert sum(rec) >= 1, "somebody must have had to recover"
        assert out == ref, "every concurrent session continues token-exact"
        batched = any(h.handler.pool.batches and h.handler.pool.batched_items > h.handler.pool.batches for h in handles)
        assert batched, "the servers did batch concurrent steps"
    finally:
        for h in handles:
            try:
                h.stop()
            except Exception:
                pass

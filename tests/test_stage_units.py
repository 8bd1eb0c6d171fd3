"""CPU unit tests of the stage-side pieces that the pipeline tests do not reach directly: the offload span,
the role / span validation of the stage loaders and the throughput estimate with its file cache."""
import json
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _tiny(model="tiny-llama", dtype=torch.float32):
    from petals_b200.models.config import PRESETS
    from petals_b200.weights.loader import load_span

    cfg = PRESETS[model]
    w = load_span(model, cfg, 0, cfg.n_layers, need_embed=True, need_head=True, dtype=dtype, device="cpu", seed=3)
    return cfg, w


@pytest.mark.parametrize("model,keep", [("tiny-llama", 0), ("tiny-llama", 1), ("tiny-gpt2", 2), ("tiny-llama", 99)])
def test_offloaded_span_equals_resident_span(model, keep):
    """Streaming blocks through two device slots must not change a single value (reference: the per-forward
    `.to()` shuffling of src/llama_partition.py:161-220); `keep` blocks stay resident, the rest are streamed."""
    from petals_b200.models.torch_ref import TorchSpan
    from petals_b200.runtime.kv_cache import KVCacheManager
    from petals_b200.runtime.offload import OffloadedSpan

    cfg, w = _tiny(model)
    ref = TorchSpan(cfg, w, "cpu", torch.float32)
    off = OffloadedSpan(cfg, dict(w), "cpu", torch.float32, keep_layers_on_gpu=keep)
    assert off.keep == min(keep, cfg.n_layers) and len(off.host) == cfg.n_layers - off.keep
    mgr = KVCacheManager(cfg.n_layers, cfg.n_kv_heads, cfg.head_dim, torch.float32, "cpu")
    kv_a, kv_b = mgr.allocate("a", 32), mgr.allocate("b", 32)
    ids = torch.randint(0, cfg.vocab, (6,), generator=torch.Generator().manual_seed(1))
    ha, hb = ref.forward(ref.embed(ids, 0), kv_a), off.forward(off.embed(ids, 0), kv_b)
    assert torch.equal(ha, hb) and kv_a.length == kv_b.length == 6
    for step in range(2):
        tok = torch.tensor([int(ids[step])])
        ha, hb = ref.forward(ref.embed(tok, kv_a.length), kv_a), off.forward(off.embed(tok, kv_b.length), kv_b)
        assert torch.equal(ha, hb)
    assert torch.equal(ref.head(ha), off.head(hb))


def test_stage_loader_roles_and_spans():
    from src import llama_partition, partition

    full = partition.load_stage_model("tiny-gpt2", "cpu", "stage0", end=2, dtype=torch.float32)
    assert (full.role, full.start, full.end) == ("stage0", 0, 2) and "embed" in full.weights
    seg = partition.load_stage_model("tiny-llama", "cpu", "segment", start=1, end=3, dtype=torch.float32)
    assert (seg.start, seg.end) == (1, 3) and len(seg.weights["layers"]) == 2
    last = partition.load_stage_model("tiny-llama", "cpu", "last", start=3, dtype=torch.float32)
    assert last.end == last.cfg.n_layers and "lm_head" in last.weights
    with pytest.raises(ValueError):
        partition.load_stage_model("tiny-llama", "cpu", "middle", start=1, end=2)
    with pytest.raises(ValueError):
        partition.load_stage_model("tiny-llama", "cpu", "segment", start=3, end=3)
    # the Llama-only module keeps the reference's restriction (src/llama_partition.py:81-83 there)
    with pytest.raises(ValueError):
        llama_partition.load_stage_model("tiny-gpt2", "cpu", "stage0", end=2)
    # a stage object refuses weights of another span
    with pytest.raises(ValueError):
        llama_partition.StageSegment(seg, 0, 3)


def test_throughput_estimate_and_cache(tmp_path, monkeypatch):
    from src import throughput_measurement as tm

    # 100 Mbit/s over one fp16 hidden state of 4096 values = 65536 bits per request
    assert tm.estimate_network_throughput(4096, torch.float16) == pytest.approx(100e6 / 65536)
    assert tm.estimate_network_throughput(4096, torch.float16, 1000.0) == pytest.approx(1e9 / 65536)

    monkeypatch.setattr(tm, "CACHE_PATH", str(tmp_path / "cache" / "throughput.json"))

    class Stage:               # a stage whose measurement fails: the estimate falls back to the network bound
        start, end = 2, 5

        class cfg:
            hidden = 4096

    n = 3   # blocks of the stage
    # measurement fails: the estimate falls back to the network bound, nothing is cached
    calls = []
    monkeypatch.setattr(tm, "measure_compute_throughput", lambda *a, **k: calls.append(1) or (_ for _ in ()).throw(RuntimeError("no")))
    monkeypatch.setattr(tm, "measure_forward_throughput", lambda *a, **k: (_ for _ in ()).throw(RuntimeError("no")))
    thr = tm.get_server_throughput(Stage(), "cpu", torch.float16, network_bandwidth_mbps=100.0, use_cache=True, model_name="m")
    assert thr == pytest.approx(100e6 / 65536 * (1 - tm.NETWORK_OVERHEAD)) and calls == [1]
    assert not os.path.exists(tm.CACHE_PATH)
    # upstream model: per-block rps, (n+1)/2 blocks used on average, min with the network
    monkeypatch.setattr(tm, "measure_compute_throughput", lambda *a, **k: calls.append(2) or 50.0)    # steps/s of the span
    monkeypatch.setattr(tm, "measure_forward_throughput", lambda *a, **k: 4000.0)                      # tokens/s of the span
    info = tm.get_server_throughput_info(Stage(), "cpu", torch.float16, network_bandwidth_mbps=1e6, use_cache=True,
                                         model_name="m")
    assert info["inference_rps"] == pytest.approx(50.0 * n) and info["forward_rps"] == pytest.approx(4000.0 * n)
    assert info["throughput"] == pytest.approx(4000.0 * n / ((n + 1) / 2))
    stored = json.load(open(tm.CACHE_PATH))
    assert list(stored) == ["m|cpu|torch.float16|per_block"]
    # second call is served from the file (the per-block numbers do not depend on the span), nothing is re-measured
    calls.clear()
    again = tm.get_server_throughput(Stage(), "cpu", torch.float16, network_bandwidth_mbps=1e6, use_cache=True, model_name="m")
    assert again == pytest.approx(info["throughput"]) and calls == []
    # network-bound case and the relay penalty
    slow = tm.get_server_throughput_info(Stage(), "cpu", torch.float16, network_bandwidth_mbps=100.0, use_cache=True, model_name="m")
    assert slow["throughput"] == pytest.approx(slow["network_rps"]) == pytest.approx(100e6 / 65536 * 0.8)
    relay = tm.get_server_throughput_info(Stage(), "cpu", torch.float16, network_bandwidth_mbps=100.0, use_cache=True,
                                          model_name="m", reachable_via_relay=True)
    assert relay["throughput"] == pytest.approx(0.2 * slow["throughput"])
    # no bandwidth given, no process group: the loopback RPC itself is timed (not the 100 Mbit/s constant)
    lo = tm.get_server_throughput_info(Stage(), "cpu", torch.float16, use_cache=True, model_name="m")
    assert lo["network_rps"] > 50            # a real measurement of this machine's loopback round trips
    monkeypatch.setattr(tm, "measure_loopback_rpc_rps", lambda *a, **k: 7777.0)
    lo = tm.get_server_throughput_info(Stage(), "cpu", torch.float16, use_cache=True, model_name="m")
    assert lo["network_rps"] == 7777.0 and lo["throughput"] == pytest.approx(4000.0 * n / ((n + 1) / 2))


def test_cache_and_position_helpers():
    """src/utils.py keeps the reference's helper names (there: HF Cache <-> legacy tuples, src/utils.py:5-64)."""
    from petals_b200.runtime.kv_cache import KVCacheManager
    from src import utils

    assert utils.extract_kv_tuple(None) is None and utils.past_length(None) == 0
    assert utils.default_position_ids(None, 3).tolist() == [[0, 1, 2]]
    kv = KVCacheManager(2, 2, 8, torch.float32, "cpu").allocate("s", 16)
    kv.data.normal_()
    kv.length = 5
    k, v = utils.extract_kv_tuple(kv, layer=1)
    assert k.shape == v.shape == (2, 5, 8)
    assert torch.equal(k, kv.data[1, 0, :, :5]) and torch.equal(v, kv.data[1, 1, :, :5])
    assert utils.past_length(kv) == 5 and utils.default_position_ids(kv, 2).tolist() == [[5, 6]]
    assert utils.normalize_cache(kv) is kv
    kv.rollback(3)
    assert utils.past_length(kv) == 3
    with pytest.raises(ValueError):
        kv.rollback(9)

"""The TCP RPC path (cross-host fallback) with sm_100a stages: client Stage0 + two server stages in one
process on cuda:0 must generate the same tokens as the in-box engine."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_rpc_pipeline_with_native_stages_matches_engine():
    from petals_b200.models.config import resolve_config
    from petals_b200.parallel.pipeline import PipelineEngine
    from petals_b200.utils.tokenizer import load_tokenizer
    from src.main import build_parser, parse_splits, run_rank0, run_stage_server_fixed

    model, splits = "tiny-llama-gqa", "1,3"
    dev = torch.device("cuda", 0)

    def args(stage, peers="", **kw):
        a = build_parser().parse_args(["--model", model, "--splits", splits, "--stage", str(stage), "--dtype", "bf16",
                                       "--dht_initial_peers", peers, "--dht_port", "0", "--rpc_port", "0",
                                       "--temperature", "0", "--seed", "5"])
        a.dtype = torch.bfloat16
        for k, v in kw.items():
            setattr(a, k, v)
        return a

    cuts = parse_splits(splits)
    handles, peers = [], ""
    try:
        for st in (1, 2):
            h = run_stage_server_fixed(args(st, peers, kv_capacity=256), dev, cuts, block=False)
            handles.append(h)
            peers = peers or h.dht.get_visible_maddrs()[0]
        assert handles[0].handler.stage_model.backend == "native"
        out, stats = run_rank0(args(0, peers, max_new_tokens=8, kv_capacity=256), dev, cuts)
    finally:
        for h in handles:
            h.stop()
    cfg = resolve_config(model)
    ids = torch.tensor(load_tokenizer(model, cfg).encode("Hello, how are you?"), dtype=torch.long)
    eng = PipelineEngine(model, [], torch.bfloat16, dev, kv_capacity=256, max_sessions=1, seed=5)
    ref = eng.generate([ids], 8)[0][0]
    assert out == ref[: len(out)] and len(out) >= 5

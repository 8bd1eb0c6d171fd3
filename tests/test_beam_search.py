"""Client-side beam search over a served pipeline (hypotheses = sessions, re-ordered on the wire with hypo_ids) against
HF's beam search on the same weights."""
import torch

from tests.test_cpu_pipeline import _args


def _hf_beam(cfg, seed, prompt_ids, n_new, beams):
    import transformers
    from petals_b200.weights.loader import full_hf_state_dict

    hf = transformers.AutoModelForCausalLM.from_config(transformers.AutoConfig.for_model(**cfg.to_hf_dict())).eval()
    hf.load_state_dict(full_hf_state_dict(cfg, seed), strict=False)
    hf.generation_config.eos_token_id = None
    hf.generation_config.pad_token_id = 0
    with torch.no_grad():
        out = hf.generate(torch.tensor([prompt_ids]), num_beams=beams, max_new_tokens=n_new, do_sample=False,
                          length_penalty=1.0, early_stopping=False, num_return_sequences=1, use_cache=True)
    return out[0, len(prompt_ids):].tolist()


def test_beam_search_matches_hf_and_beats_greedy_score():
    from petals_b200.models.config import resolve_config
    from petals_b200.utils.tokenizer import load_tokenizer
    from src.beam_search import beam_search
    from src.main import parse_splits, run_stage_server_fixed
    from src.partition import Stage0, load_stage_model
    from src.rpc_transport import RpcTransport

    model, splits, seed = "tiny-llama-gqa", "1,3", 13
    cuts = parse_splits(splits)
    cfg = resolve_config(model)
    dev = torch.device("cpu")
    handles = []
    try:
        h1 = run_stage_server_fixed(_args(model, splits, 1, "", seed=seed), dev, cuts, block=False)
        handles.append(h1)
        peers = h1.dht.get_visible_maddrs()[0]
        handles.append(run_stage_server_fixed(_args(model, splits, 2, peers, seed=seed), dev, cuts, block=False))
        prompt = load_tokenizer(model, cfg).encode("the quick brown fox")
        full = load_stage_model(model, dev, role="stage0", end=cuts[0], dtype=torch.float32, seed=seed)
        s0 = Stage0(full, cuts[0], kv_capacity=64)
        mgr = s0.new_kv_manager()
        tx = RpcTransport(device=dev, dht_initial_peers=[peers], temperature=0.0,
                          stage_keys=["mini_petals:stage1", "mini_petals:stage2"], request_timeout=30.0)
        B, n_new = 3, 8
        best, score, hyps = beam_search(tx, s0, mgr, prompt, B, n_new)
        assert len(best) == n_new and len(hyps) == B
        # one-beam search is greedy decoding; a wider beam never ends with a worse total log-probability
        greedy, g_score, _ = beam_search(tx, s0, mgr, prompt, 1, n_new)
        assert score >= g_score - 1e-6
        assert not mgr.sessions, "the client's caches are released"
        assert all(not h.handler.kv.sessions for h in handles), "and so are the servers'"
        tx.shutdown()
        ref = _hf_beam(cfg, seed, prompt, n_new, B)
        assert best == ref, "same hypothesis as HF's beam search on the same weights"
    finally:
        for h in handles:
            h.stop()

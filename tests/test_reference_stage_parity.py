"""Same math as the reference's own stage classes: its Stage0 / StageLast (with its OptimizedLlamaDecoderLayer), run
unmodified from baseline/_ref on the weights our loader produces, against our TorchSpan stages - token by token, fp32,
on the CPU.  (Token by token because the reference's live path passes attention_mask=None to every layer - SURVEY B1 -
so its multi-token prefill is not causal; one token at a time the mask does not matter and the two must agree.)"""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


def _reference_modules():
    if not os.path.isdir(os.path.join(REF, "src")):
        if not os.path.isdir("/root/reference/src"):
            pytest.skip("the reference is not available")
        import subprocess

        subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "install_reference.py")], check=False)
    if not os.path.isdir(os.path.join(REF, "src")):
        pytest.skip("the reference could not be installed")
    sys.path.insert(0, ROOT)
    from baseline import reference_harness as H

    H._install_stubs()
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "src" or k.startswith("src.")
             or k == "petals" or k.startswith("petals.")}
    sys.path.insert(0, REF)
    try:
        import importlib

        lp = importlib.import_module("src.llama_partition")
        lp.ReferenceHandler = importlib.import_module("src.rpc_handler").StageConnectionHandler
    finally:
        sys.path.remove(REF)
        for k in [k for k in sys.modules if k == "src" or k.startswith("src.") or k == "petals" or k.startswith("petals.")]:
            sys.modules.pop(k)
        sys.modules.update(saved)
    return lp, H


def test_reference_stages_and_ours_agree_token_by_token():
    import transformers

    from petals_b200.models.config import resolve_config
    from petals_b200.runtime.stages import Stage0, StageLast, StageWeights
    from petals_b200.weights.loader import full_hf_state_dict, load_span

    lp, H = _reference_modules()
    model, seed, cut = "tiny-llama-gqa", 17, 2
    cfg = resolve_config(model)
    sd = full_hf_state_dict(cfg, seed)
    hf = cfg.to_hf_dict()

    def ref_span(start, end, embed, head):
        c = transformers.LlamaConfig(**{k: v for k, v in hf.items() if k not in ("model_type", "architectures")})
        c.num_hidden_layers = end - start
        c.pretraining_tp = 1
        c._attn_implementation = "eager"
        m = transformers.LlamaForCausalLM(c).float().eval()
        part = {}
        for k, v in sd.items():
            if k.startswith("model.layers."):
                li = int(k.split(".")[2])
                if start <= li < end:
                    part[k.replace(f"model.layers.{li}.", f"model.layers.{li - start}.")] = v.float()
            else:
                part[k] = v.float()
        missing, unexpected = m.load_state_dict(part, strict=False)
        assert not unexpected
        return m, c

    torch.manual_seed(0)
    m0, c0 = ref_span(0, cut, True, False)
    m1, c1 = ref_span(cut, cfg.n_layers, False, True)
    r0 = lp.Stage0(m0, cut)
    r1 = lp.StageLast(m1, cut, gpu_device=torch.device("cpu"), keep_layers_on_gpu=cfg.n_layers - cut)
    H._shim(r0, c0)
    H._shim(r1, c1)

    dev = torch.device("cpu")
    w0 = load_span(model, cfg, 0, cut, need_embed=True, need_head=False, dtype=torch.float32, device=dev, seed=seed)
    w1 = load_span(model, cfg, cut, cfg.n_layers, need_embed=False, need_head=True, dtype=torch.float32, device=dev, seed=seed)
    o0 = Stage0(StageWeights(cfg, w0, "stage0", 0, cut, dev, torch.float32), cut, kv_capacity=32)
    o1 = StageLast(StageWeights(cfg, w1, "last", cut, cfg.n_layers, dev, torch.float32), cut, kv_capacity=32)
    kv0 = o0.new_kv_manager().allocate("s", 32)
    kv1 = o1.new_kv_manager().allocate("s", 32)

    ids = torch.randint(0, cfg.vocab, (9,), generator=torch.Generator().manual_seed(3))
    past0 = past1 = None
    with torch.no_grad():
        for t in range(ids.numel()):
            tok = ids[t].view(1, 1)
            pos = torch.tensor([[t]])
            h_ref, past0 = r0(tok, pos, None, past0, use_cache=True)
            out_ref, past1 = r1(h_ref, pos, None, past1, use_cache=True)
            h_our, _ = o0(tok, None, None, kv0)
            out_our, _ = o1(h_our, None, None, kv1)
            assert torch.allclose(h_our.reshape(-1), h_ref.reshape(-1).float(), atol=2e-5, rtol=1e-4), f"hidden, token {t}"
            lr = out_ref.reshape(-1, cfg.vocab)[-1].float()
            lo = out_our.reshape(-1, cfg.vocab)[-1].float()
            assert torch.allclose(lo, lr, atol=5e-5, rtol=1e-4), f"logits, token {t}"
            assert int(lo.argmax()) == int(lr.argmax())


def test_reference_handler_and_ours_answer_the_same_requests():
    """The reference's StageConnectionHandler._run_forward (+ its _sample_token, greedy) and our handler, fed the same
    session - a one-token prefill, then single-token steps with the reference's metadata keys - must return the same
    tokens from the same hidden states."""
    import transformers

    from petals_b200.models.config import resolve_config
    from petals_b200.runtime.stages import StageLast, StageWeights
    from petals_b200.weights.loader import full_hf_state_dict, load_span
    from src.rpc_handler import StageConnectionHandler

    lp, H = _reference_modules()
    _, _, mser = H._install_stubs()
    model, seed, cut = "tiny-llama-gqa", 23, 1
    cfg = resolve_config(model)
    sd = full_hf_state_dict(cfg, seed)
    c = transformers.LlamaConfig(**{k: v for k, v in cfg.to_hf_dict().items() if k not in ("model_type", "architectures")})
    c.num_hidden_layers = cfg.n_layers - cut
    c.pretraining_tp = 1
    c._attn_implementation = "eager"
    m = transformers.LlamaForCausalLM(c).float().eval()
    part = {}
    for k, v in sd.items():
        if k.startswith("model.layers."):
            li = int(k.split(".")[2])
            if li >= cut:
                part[k.replace(f"model.layers.{li}.", f"model.layers.{li - cut}.")] = v.float()
        else:
            part[k] = v.float()
    m.load_state_dict(part, strict=False)
    dev = torch.device("cpu")
    ref_stage = lp.StageLast(m, cut, gpu_device=dev, keep_layers_on_gpu=cfg.n_layers - cut)
    H._shim(ref_stage, c)
    ref = lp.ReferenceHandler(None, ref_stage, dev, request_timeout=30.0, final_stage=True)

    w = load_span(model, cfg, cut, cfg.n_layers, need_embed=False, need_head=True, dtype=torch.float32, device=dev, seed=seed)
    ours = StageConnectionHandler(None, StageLast(StageWeights(cfg, w, "last", cut, cfg.n_layers, dev, torch.float32), cut,
                                                  kv_capacity=64), dev, final_stage=True)
    try:
        g = torch.Generator().manual_seed(5)
        generated = []
        for t in range(10):
            hidden = torch.randn(1, 1, cfg.hidden, generator=g) * 0.5
            md = dict(session_id="parity", seq_len=1, cur_len=t + 1, is_prefill=(t == 0), max_length=64, temperature=0.0,
                      top_p=0.92, top_k=50, generated_tokens=generated[-50:])
            r = ref._run_forward(hidden.clone(), dict(md))
            tok_ref = int(mser.loads(r.metadata)["token_id"])
            _, omd = ours.rpc_forward("k", [hidden.clone()], dict(md))
            assert omd["token_id"] == tok_ref, f"step {t}: ours {omd['token_id']} vs reference {tok_ref}"
            generated.append(tok_ref)
    finally:
        ours.shutdown()

"""NativeSpan (sm_100a kernels + CUDA graph) vs the torch restatement in fp32."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _build(name, dtype, **kw):
    from petals_b200.models.config import PRESETS
    from petals_b200.models.torch_ref import TorchSpan
    from petals_b200.runtime.kv_cache import KVCacheManager
    from petals_b200.runtime.native_span import NativeSpan
    from petals_b200.weights.loader import load_span

    cfg = PRESETS[name]
    dev = torch.device("cuda")
    w16 = load_span(name, cfg, 0, cfg.n_layers, need_embed=True, need_head=True, dtype=dtype, device=dev, seed=3)
    # fp32 oracle runs on the SAME (16-bit rounded) weights
    w32 = {k: (v.float() if torch.is_tensor(v) else v) for k, v in w16.items() if k != "layers"}
    w32["layers"] = [{k: v.float() for k, v in L.items()} for L in w16["layers"]]
    ref = TorchSpan(cfg, w32, dev, torch.float32)
    nat = NativeSpan(cfg, w16, dev, dtype, kv_capacity=256, max_batch=4, max_chunk=64, free_source_weights=False, **kw)
    mref = KVCacheManager(cfg.n_layers, cfg.n_kv_heads, cfg.head_dim, torch.float32, dev)
    mnat = KVCacheManager(cfg.n_layers, cfg.n_kv_heads, cfg.head_dim, dtype, dev)
    return cfg, ref, nat, mref, mnat


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-llama-gqa", "tiny-gpt2"])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("graphs", [True, False])
def test_prefill_then_decode_matches_oracle(name, dtype, graphs):
    cfg, ref, nat, mref, mnat = _build(name, dtype, use_graphs=graphs)
    torch.manual_seed(0)
    ids = torch.randint(0, cfg.vocab, (100,), device="cuda")
    L = 70  # > max_chunk: exercises chunked prefill
    kr, kn = mref.allocate("a", 256), mnat.allocate("a", 256)
    lr = ref.head(ref.forward(ref.embed(ids[:L], 0), kr)[-1:])
    ln, tok = nat.prefill(kn, ids=ids[:L])
    tol = 0.08 if dtype == torch.bfloat16 else 0.02
    scale = lr.abs().max().item()
    assert (ln - lr).abs().max().item() < tol * scale
    assert kn.length == L
    agree = int(tok[0]) == int(lr.argmax())
    for t in range(L, L + 12):
        lr = ref.head(ref.forward(ref.embed(ids[t:t + 1], t), kr))
        ln, tok = nat.decode([kn], ids=ids[t:t + 1])
        assert (ln - lr).abs().max().item() < tol * scale, f"step {t}"
        top2 = lr[0].topk(2).values
        if (top2[0] - top2[1]).item() > 4 * tol * scale:
            assert int(tok[0]) == int(lr.argmax())
    assert kn.length == L + 12


def test_persistent_decode_kernel_matches_multi_kernel_path():
    """The one-launch persistent decode kernel must be bit-identical to the per-op launches."""
    cfg, ref, nat, mref, mnat = _build("tiny-llama-gqa", torch.bfloat16, use_mega=False)
    _, _, nat2, _, mnat2 = _build("tiny-llama-gqa", torch.bfloat16, use_mega=True)
    torch.manual_seed(2)
    ids = torch.randint(0, cfg.vocab, (30,), device="cuda")
    ka, kb = mnat.allocate("a", 256), mnat2.allocate("a", 256)
    _, ta = nat.prefill(ka, ids=ids)
    _, tb = nat2.prefill(kb, ids=ids)
    assert torch.equal(ta, tb)
    for _ in range(6):
        la, ta = nat.decode([ka], ids=ta.clone())
        lb, tb = nat2.decode([kb], ids=tb.clone())
        assert torch.equal(la, lb) and torch.equal(ta, tb)


def test_batched_decode_equals_single(monkeypatch):
    cfg, ref, nat, mref, mnat = _build("tiny-llama-gqa", torch.bfloat16)
    torch.manual_seed(1)
    prompts = [torch.randint(0, cfg.vocab, (n,), device="cuda") for n in (5, 17, 40)]
    kvs = [mnat.allocate(f"s{i}", 256) for i in range(3)]
    nxt = []
    for p, kv in zip(prompts, kvs):
        _, tok = nat.prefill(kv, ids=p)
        nxt.append(tok.clone())
    ids = torch.cat(nxt)
    logits_b, toks_b = nat.decode(kvs, ids=ids)
    logits_b = logits_b.clone()
    # same step one session at a time on fresh copies
    kvs2 = [mnat.allocate(f"t{i}", 256) for i in range(3)]
    for i, (p, kv) in enumerate(zip(prompts, kvs2)):
        nat.prefill(kv, ids=p)
        l1, _ = nat.decode([kv], ids=ids[i:i + 1])
        assert torch.equal(l1[0], logits_b[i]), "batched decode must be bit-identical to single-session decode"


def test_fp8_weight_span_tracks_the_oracle():
    """Opt-in e4m3 weights: logits stay within quantisation error of the fp32 oracle."""
    cfg, ref, nat, mref, mnat = _build("tiny-llama-gqa", torch.bfloat16, weight_format="fp8")
    torch.manual_seed(4)
    ids = torch.randint(0, cfg.vocab, (50,), device="cuda")
    kr, kn = mref.allocate("a", 256), mnat.allocate("a", 256)
    lr = ref.head(ref.forward(ref.embed(ids[:40], 0), kr)[-1:])
    ln, _ = nat.prefill(kn, ids=ids[:40])                 # 40 tokens -> two <=32-token fp8 chunks
    scale = lr.abs().max().item()
    assert (ln - lr).abs().max().item() < 0.15 * scale
    for t in range(40, 46):
        lr = ref.head(ref.forward(ref.embed(ids[t:t + 1], t), kr))
        ln, _ = nat.decode([kn], ids=ids[t:t + 1])
        assert (ln - lr).abs().max().item() < 0.15 * scale
    w16 = sum(P[k].nbytes for P in nat.layers for k in ("qkv", "o", "mlp_in", "down"))
    assert w16 == cfg.n_layers * (cfg.qkv_size * cfg.hidden + cfg.hidden * cfg.q_size + 3 * cfg.hidden * cfg.intermediate)

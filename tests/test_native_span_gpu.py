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


@pytest.mark.parametrize("fmt", ["fp8", "mxfp8"])
def test_fp8_weight_span_tracks_the_oracle(fmt):
    """Opt-in e4m3 weights (one scale per row, or OCP-MX block scales): logits stay within quantisation error of the
    fp32 oracle."""
    cfg, ref, nat, mref, mnat = _build("tiny-llama-gqa", torch.bfloat16, weight_format=fmt)
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
    w8 = sum(P[k].nbytes for P in nat.layers for k in ("qkv", "o", "mlp_in", "down"))
    dense = cfg.n_layers * (cfg.qkv_size * cfg.hidden + cfg.hidden * cfg.q_size + 3 * cfg.hidden * cfg.intermediate)
    assert w8 == dense if fmt == "fp8" else dense < w8 <= dense * 1.07      # MX: + 1 KiB of scales per 16/32 KiB k-block


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-llama-gqa", "tiny-gpt2"])
@pytest.mark.parametrize("graphs", [True, False])
def test_flag_chained_decode_is_bit_identical(name, graphs, monkeypatch):
    """A decode step run as a flag chain (ops poll tagged words instead of waiting for each other's grids) must give
    exactly the logits of the grid-ordered step: same arithmetic, same summation order - for 1..4 sessions per step, over several steps (stale words of an
    earlier step must never be taken for new ones)."""
    from petals_b200.runtime import native_span

    monkeypatch.setattr(native_span, "CHAIN_MAX_TOKENS", 4)      # off by default (measured slower); tested all the same
    cfg, _, chain, _, m1 = _build(name, torch.bfloat16, use_graphs=graphs)
    _, _, plain, _, m2 = _build(name, torch.bfloat16, use_graphs=graphs)
    assert chain.chain_tokens >= 4, "the tiny presets must qualify for the chain"
    plain.chain_tokens = 0
    torch.manual_seed(5)
    prompts = [torch.randint(0, cfg.vocab, (n,), device="cuda") for n in (33, 7, 50, 12)]
    ka = [m1.allocate(f"a{i}", 256) for i in range(4)]
    kb = [m2.allocate(f"b{i}", 256) for i in range(4)]
    toks = []
    for p, x, y in zip(prompts, ka, kb):
        _, t1 = chain.prefill(x, ids=p)
        _, t2 = plain.prefill(y, ids=p)
        assert torch.equal(t1, t2)
        toks.append(t1.clone())
    ids = torch.cat(toks)
    launches0 = chain.kernel_launches
    for step in range(6):
        for B in (1, 2, 4, 3):                      # alternating batch sizes share the word buffers
            la, ta = chain.decode(ka[:B], ids=ids[:B].clone())
            lb, tb = plain.decode(kb[:B], ids=ids[:B].clone())
            assert torch.equal(la, lb), f"step {step} batch {B}: chained logits differ"
            assert torch.equal(ta, tb)
            ids[:B] = ta
    assert chain.kernel_launches > launches0
    # the device-resident greedy loop (positions and the step number advance inside the graph)
    ha = chain.decode_loop(ka[:2], ids[:2].clone(), 5).clone()
    hb = plain.decode_loop(kb[:2], ids[:2].clone(), 5).clone()
    assert torch.equal(ha, hb)


def test_real_shape_layer_elementwise_against_fp32_oracle():
    """One decoder block with Llama-2-7B's real shapes (H = 4096, 32 heads of 128, I = 11008): the native kernels'
    hidden states against the fp32 torch restatement on the same (16-bit rounded) weights, ELEMENT BY ELEMENT
    (rtol / atol, not a max-normalised error) - prefill of a chunk, then single-token steps."""
    from dataclasses import replace

    from petals_b200.models.config import resolve_config
    from petals_b200.models.torch_ref import TorchSpan
    from petals_b200.runtime.kv_cache import KVCacheManager
    from petals_b200.runtime.native_span import NativeSpan
    from petals_b200.weights.loader import load_span

    dtype = torch.float16
    cfg = replace(resolve_config("llama-2-7b"), n_layers=1)
    dev = torch.device("cuda")
    w16 = load_span("llama-2-7b", cfg, 0, 1, need_embed=False, need_head=False, dtype=dtype, device=dev, seed=1)
    w32 = {k: (v.float() if torch.is_tensor(v) else v) for k, v in w16.items() if k != "layers"}
    w32["layers"] = [{k: v.float() for k, v in L.items()} for L in w16["layers"]]
    ref = TorchSpan(cfg, w32, dev, torch.float32)
    nat = NativeSpan(cfg, w16, dev, dtype, kv_capacity=256, max_batch=2, max_chunk=128, free_source_weights=False)
    kr = KVCacheManager(1, cfg.n_kv_heads, cfg.head_dim, torch.float32, dev).allocate("a", 256)
    kn = KVCacheManager(1, cfg.n_kv_heads, cfg.head_dim, dtype, dev).allocate("a", 256)
    torch.manual_seed(3)
    x = (torch.randn(100, cfg.hidden, device=dev) * 0.7).to(dtype)
    want = ref.forward(x[:96].float(), kr)
    got = nat.prefill(kn, h_in=x[:96])
    atol = 0.02 * want.abs().mean().item()
    torch.testing.assert_close(got.float(), want, rtol=2e-2, atol=atol)
    for t in range(96, 100):
        want = ref.forward(x[t:t + 1].float(), kr)
        got = nat.decode([kn], h_in=x[t:t + 1])
        torch.testing.assert_close(got.float(), want, rtol=2e-2, atol=atol)


def test_random_weights_are_the_same_on_cpu_and_gpu():
    """Large random-init tensors come from an integer-only counter generator: bit-identical wherever they are made."""
    from petals_b200.models.config import resolve_config
    from petals_b200.weights.loader import RandomSource

    cfg = resolve_config("llama-2-7b")
    name = "model.layers.0.self_attn.q_proj.weight"            # 16.8M elements: beyond the CPU-generator limit
    a = RandomSource(cfg, 7, "cpu").get(name)
    b = RandomSource(cfg, 7, "cuda").get(name)
    assert b.is_cuda and torch.equal(a, b.cpu())
    assert abs(a.std().item() - 0.02) < 2e-4


@pytest.mark.parametrize("name", ["tiny-llama-gqa", "tiny-gpt2"])
@pytest.mark.parametrize("keep", [0, 1])
def test_offloaded_native_span_equals_resident_span(name, keep):
    """--use_cpu_offload on a B200: packed blocks stream from pinned host memory through two device slots under the
    same sm_100a kernels - bit-identical to the resident span."""
    from petals_b200.models.config import PRESETS
    from petals_b200.runtime.kv_cache import KVCacheManager
    from petals_b200.runtime.native_span import NativeSpan
    from petals_b200.runtime.offload import OffloadedNativeSpan, make_offloaded_native_span
    from petals_b200.weights.loader import load_span

    cfg = PRESETS[name]
    dev = torch.device("cuda")
    dtype = torch.bfloat16
    load = lambda d: load_span(name, cfg, 0, cfg.n_layers, need_embed=True, need_head=True, dtype=dtype, device=d, seed=3)  # noqa: E731
    res = NativeSpan(cfg, load(dev), dev, dtype, kv_capacity=256, max_batch=2, max_chunk=64, use_graphs=False)
    off = make_offloaded_native_span(cfg, load("cpu"), dev, dtype, keep_layers_on_gpu=keep, kv_capacity=256, max_batch=2,
                                     max_chunk=64)
    assert isinstance(off, OffloadedNativeSpan) and off.n_layers == cfg.n_layers and len(off.layers) == keep
    ka = KVCacheManager(cfg.n_layers, cfg.n_kv_heads, cfg.head_dim, dtype, dev).allocate("a", 256)
    kb = KVCacheManager(cfg.n_layers, cfg.n_kv_heads, cfg.head_dim, dtype, dev).allocate("b", 256)
    torch.manual_seed(0)
    ids = torch.randint(0, cfg.vocab, (90,), device=dev)
    la, ta = res.prefill(ka, ids=ids[:80])
    lb, tb = off.prefill(kb, ids=ids[:80])
    assert torch.equal(la, lb) and torch.equal(ta, tb)
    for t in range(80, 86):
        la, ta = res.decode([ka], ids=ids[t:t + 1])
        lb, tb = off.decode([kb], ids=ids[t:t + 1])
        assert torch.equal(la, lb) and torch.equal(ta, tb)
    assert off.bytes_streamed > 0


@pytest.mark.parametrize("name", ["tiny-llama-gqa", "tiny-gpt2"])
def test_subspan_views_chain_to_the_whole_span(name):
    """Views over blocks [0,a) and [a,n) of one span (shared packed weights, own graphs and caches) run back to back
    give bit-identical hidden states and logits to the whole span - the server side of a client that enters a span
    mid-way."""
    from petals_b200.runtime.kv_cache import KVCacheManager

    dtype = torch.bfloat16
    cfg, _, nat, _, mnat = _build(name, dtype)
    n = cfg.n_layers
    a = n // 2
    lo, hi = nat.subspan(0, a), nat.subspan(a, n)
    assert lo.has_embed and not lo.has_head and hi.has_head and not hi.has_embed
    assert lo.layers[0] is nat.layers[0] and hi.layers[-1] is nat.layers[-1], "views share the packed weights"
    dev = torch.device("cuda")
    m_lo = KVCacheManager(a, cfg.n_kv_heads, cfg.head_dim, dtype, dev)
    m_hi = KVCacheManager(n - a, cfg.n_kv_heads, cfg.head_dim, dtype, dev)
    torch.manual_seed(5)
    ids = torch.randint(0, cfg.vocab, (60,), device="cuda")
    L = 40
    kw, kl, kh = mnat.allocate("w", 256), m_lo.allocate("w", 256), m_hi.allocate("w", 256)
    ref_logits, ref_tok = nat.prefill(kw, ids=ids[:L])
    ref_logits, ref_tok = ref_logits.clone(), ref_tok.clone()
    h = lo.prefill(kl, ids=ids[:L]).clone()
    logits, tok = hi.prefill(kh, h_in=h)
    assert torch.equal(logits, ref_logits) and torch.equal(tok, ref_tok)
    for t in range(L, L + 8):
        ref_logits, ref_tok = (x.clone() for x in nat.decode([kw], ids=ids[t:t + 1]))
        h = lo.decode([kl], ids=ids[t:t + 1]).clone()
        logits, tok = hi.decode([kh], h_in=h)
        assert torch.equal(logits, ref_logits) and torch.equal(tok, ref_tok), f"step {t}"
    assert kl.length == kh.length == kw.length == L + 8

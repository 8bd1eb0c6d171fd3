"""tcgen05 stream-K GEMM vs a plain torch fp32 restatement of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(x, w, bias=None, residual=None, epi="none", norm=None, up=None, dtype=torch.bfloat16):
    xf = x.float()
    if norm is not None:
        mode, g, b, eps = norm
        if mode == "rms":
            var = xf.pow(2).mean(-1, keepdim=True)
            xn = (xf * torch.rsqrt(var + eps)).to(dtype).float()
            xf = (g.float() * xn).to(dtype).float()
        else:
            xf = torch.nn.functional.layer_norm(xf, (xf.shape[-1],), g.float(), b.float(), eps).to(dtype).float()
    y = xf @ w.float().t()
    if bias is not None:
        y = y + bias.float()
    if epi == "gelu":
        y = torch.nn.functional.gelu(y.to(dtype).float(), approximate="tanh")
    if epi == "swiglu":
        u = (xf @ up.float().t()).to(dtype).float()
        g_ = y.to(dtype).float()
        y = torch.nn.functional.silu(g_).to(dtype).float() * u
    if residual is not None:
        y = y.to(dtype).float() + residual.float()
    return y


def _close(out, ref, dtype):
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    assert err / scale < tol, f"rel err {err/scale:.4g} (abs {err:.4g}, scale {scale:.4g})"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,N,K", [(1, 256, 128), (1, 4096, 4096), (5, 1000, 512), (16, 512, 11008),
                                   (20, 384, 256), (48, 256, 512), (100, 640, 1024), (300, 512, 768),
                                   (1024, 1024, 512)])
def test_gemm_plain(T, N, K, dtype):
    from petals_b200.ops import gemm, packing

    torch.manual_seed(T * 7 + N)
    dev = "cuda"
    x = (torch.randn(T, K, device=dev) * 0.5).to(dtype)
    w = (torch.randn(N, K, device=dev) / K ** 0.5).to(dtype)
    pw = packing.pack_weight(w)
    out = gemm.linear(x, pw)
    torch.cuda.synchronize()
    _close(out, _ref(x, w, dtype=dtype), dtype)
    # second launch re-uses the self-cleaning counters
    out2 = gemm.linear(x, pw)
    torch.cuda.synchronize()
    assert torch.equal(out, out2), "stream-K result must be bit-reproducible"


@pytest.mark.parametrize("T", [1, 7, 16, 130])
def test_gemm_bias_residual_gelu(T):
    from petals_b200.ops import gemm, packing

    dtype = torch.float16
    torch.manual_seed(3)
    N, K = 3072, 768
    x = (torch.randn(T, K, device="cuda") * 0.5).to(dtype)
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    b = (torch.randn(N, device="cuda") * 0.1).to(dtype)
    r = torch.randn(T, N, device="cuda").to(dtype)
    pw = packing.pack_weight(w, b)
    out = gemm.linear(x, pw, epilogue=gemm.EPI_GELU_TANH)
    _close(out, _ref(x, w, bias=b, epi="gelu", dtype=dtype), dtype)
    out = gemm.linear(x, pw, residual=r)
    _close(out, _ref(x, w, bias=b, residual=r, dtype=dtype), dtype)


@pytest.mark.parametrize("T", [1, 3, 16, 64, 200])
def test_gemm_swiglu_dual(T):
    from petals_b200.ops import gemm, packing

    dtype = torch.bfloat16
    torch.manual_seed(5)
    N, K = 1408, 1024
    x = (torch.randn(T, K, device="cuda") * 0.5).to(dtype)
    wg = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    wu = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    pw = packing.pack_dual(wg, wu)
    out = gemm.linear(x, pw)
    _close(out, _ref(x, wg, epi="swiglu", up=wu, dtype=dtype), dtype)


@pytest.mark.parametrize("mode", ["rms", "layer"])
@pytest.mark.parametrize("T", [1, 4, 16, 29])
def test_gemm_fused_norm(mode, T):
    from petals_b200.ops import gemm, packing

    dtype = torch.bfloat16
    torch.manual_seed(11)
    N, K = 768, 1024
    x = (torch.randn(T, K, device="cuda") * 2 + 0.3).to(dtype)
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    g = (1 + 0.1 * torch.randn(K, device="cuda")).to(dtype)
    b = (0.1 * torch.randn(K, device="cuda")).to(dtype)
    pw = packing.pack_weight(w)
    if mode == "rms":
        out = gemm.linear(x, pw, norm_mode=gemm.NORM_RMS, norm_w=g, norm_eps=1e-5)
        ref = _ref(x, w, norm=("rms", g, None, 1e-5), dtype=dtype)
    else:
        out = gemm.linear(x, pw, norm_mode=gemm.NORM_LAYER, norm_w=g, norm_b=b, norm_eps=1e-5)
        ref = _ref(x, w, norm=("layer", g, b, 1e-5), dtype=dtype)
    _close(out, ref, dtype)


def test_gemm_fp32_logits_and_small_grid():
    from petals_b200.ops import gemm, packing

    dtype = torch.bfloat16
    torch.manual_seed(13)
    N, K, T = 50257, 768, 2
    x = torch.randn(T, K, device="cuda").to(dtype)
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    pw = packing.pack_weight(w)
    out = gemm.linear(x, pw, out_fp32=True)
    assert out.dtype == torch.float32 and out.shape == (T, N)
    _close(out, _ref(x, w, dtype=dtype), dtype)
    out_small = gemm.linear(x, pw, out_fp32=True, grid=7)
    _close(out_small, _ref(x, w, dtype=dtype), dtype)


def _fp8_emulate_act(xn):
    """Per-token e4m3 round trip of already-normalised (minus 1/rms) activations."""
    am = xn.abs().amax(-1, keepdim=True).clamp_min(1e-20)
    q = (xn * (448.0 / am)).to(torch.float8_e4m3fn).float()
    return q * (am / 448.0)


@pytest.mark.parametrize("T", [1, 5, 16, 27])
@pytest.mark.parametrize("norm", [False, True])
def test_gemm_fp8_weights(T, norm):
    from petals_b200.ops import gemm, packing

    dtype = torch.bfloat16
    torch.manual_seed(21 + T)
    N, K = 1024, 2048
    x = (torch.randn(T, K, device="cuda") * 1.5).to(dtype)
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    g = (1 + 0.1 * torch.randn(K, device="cuda")).to(dtype)
    res = torch.randn(T, N, device="cuda").to(dtype)
    pw = packing.pack_weight_fp8(w)
    assert pw.nbytes == N * K                                    # half the bytes of the 16-bit weight
    xf = x.float()
    if norm:
        rstd = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)
        xa = _fp8_emulate_act(xf * g.float()) * rstd
        out = gemm.linear(x, pw, norm_mode=gemm.NORM_RMS, norm_w=g, norm_eps=1e-5, residual=res)
    else:
        xa = _fp8_emulate_act(xf)
        out = gemm.linear(x, pw, residual=res)
    ref = (xa @ packing.dequantize_fp8(w).t()).to(dtype).float() + res.float()
    _close(out, ref, dtype)
    exact = ((xf * g.float() * rstd if norm else xf) @ w.float().t()) + res.float()
    rel = (out.float() - exact).abs().max().item() / exact.abs().max().item()
    assert rel < 0.06, f"fp8 quantisation error {rel:.3f} out of range"


def test_gemm_fp8_swiglu_dual():
    from petals_b200.ops import gemm, packing

    dtype = torch.float16
    torch.manual_seed(33)
    N, K, T = 1408, 1024, 3
    x = (torch.randn(T, K, device="cuda")).to(dtype)
    wg = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    wu = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    pw = packing.pack_dual_fp8(wg, wu)
    out = gemm.linear(x, pw)
    xa = _fp8_emulate_act(x.float())
    gq = (xa @ packing.dequantize_fp8(wg).t()).to(dtype).float()
    uq = (xa @ packing.dequantize_fp8(wu).t()).to(dtype).float()
    ref = torch.nn.functional.silu(gq).to(dtype).float() * uq
    _close(out, ref, dtype)


def _mx_emulate_act(xn):
    """OCP-MX round trip of activations: one power-of-two scale per 32 elements, e4m3 values (the kernel's rule)."""
    from petals_b200.ops import packing

    return packing.mx_dequantize(xn)


@pytest.mark.parametrize("T", [1, 3, 16, 27])
@pytest.mark.parametrize("norm", [False, True])
def test_gemm_mxfp8_block_scaled(T, norm):
    """tcgen05.mma.kind::mxf8f6f4.block_scale: e4m3 weights and activations, UE8M0 scales per 32 elements in tensor
    memory.  Checked against an fp32 matmul of the block-quantised operands (same quantisation rule), and against the
    unquantised product within the quantisation error."""
    from petals_b200.ops import gemm, packing

    dtype = torch.bfloat16
    torch.manual_seed(31 + T)
    N, K = 1024, 2048
    x = (torch.randn(T, K, device="cuda") * 1.5).to(dtype)
    x[:, ::7] *= 8.0                                               # outliers: what block scales are for
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    w[::5] *= 4.0
    g = (1 + 0.1 * torch.randn(K, device="cuda")).to(dtype)
    res = torch.randn(T, N, device="cuda").to(dtype)
    pw = packing.pack_weight_mx(w)
    assert pw.mx and pw.nbytes == (N // 128) * (K // 128) * (16384 + 1024)
    xf = x.float()
    if norm:
        rstd = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)
        xa = _mx_emulate_act(xf * g.float()) * rstd
        out = gemm.linear(x, pw, norm_mode=gemm.NORM_RMS, norm_w=g, norm_eps=1e-5, residual=res)
    else:
        xa = _mx_emulate_act(xf)
        out = gemm.linear(x, pw, residual=res)
    torch.cuda.synchronize()
    ref = (xa @ packing.mx_dequantize(w).t()).to(dtype).float() + res.float()
    _close(out, ref, dtype)
    exact = ((xf * g.float() * rstd if norm else xf) @ w.float().t()) + res.float()
    rel = (out.float() - exact).abs().max().item() / exact.abs().max().item()
    assert rel < 0.06, f"MX quantisation error {rel:.3f} out of range"
    out2 = gemm.linear(x, pw, norm_mode=gemm.NORM_RMS, norm_w=g, norm_eps=1e-5, residual=res) if norm else gemm.linear(x, pw, residual=res)
    torch.cuda.synchronize()
    assert torch.equal(out, out2)


def test_gemm_mxfp8_swiglu_dual():
    from petals_b200.ops import gemm, packing

    dtype = torch.float16
    torch.manual_seed(34)
    N, K, T = 1408, 1024, 3
    x = (torch.randn(T, K, device="cuda")).to(dtype)
    wg = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    wu = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)
    pw = packing.pack_dual_mx(wg, wu)
    out = gemm.linear(x, pw)
    torch.cuda.synchronize()
    xa = _mx_emulate_act(x.float())
    gq = (xa @ packing.mx_dequantize(wg).t()).to(dtype).float()
    uq = (xa @ packing.mx_dequantize(wu).t()).to(dtype).float()
    ref = torch.nn.functional.silu(gq).to(dtype).float() * uq
    _close(out, ref, dtype)


def test_gemm_l2_prefetch_hint_changes_nothing():
    """The optional L2 warm-up ranges are a pure hint: same bits out, odd sizes and offsets accepted."""
    from petals_b200.ops import gemm, packing

    dtype = torch.float16
    torch.manual_seed(5)
    x = (torch.randn(3, 1024, device="cuda") * 0.5).to(dtype)
    w = (torch.randn(768, 1024, device="cuda") / 32).to(dtype)
    nxt = packing.pack_weight((torch.randn(512, 1024, device="cuda") / 32).to(dtype))
    pw = packing.pack_weight(w)
    base = gemm.linear(x, pw)
    hinted = gemm.linear(x, pw, prefetch=[(nxt, 0, nxt.nbytes), (nxt, 4096 + 24, 100_000)])
    past_end = gemm.linear(x, pw, prefetch=[(nxt, nxt.nbytes - 32, 1 << 30)])
    torch.cuda.synchronize()
    assert torch.equal(base, hinted) and torch.equal(base, past_end)
    _close(base, _ref(x, w, dtype=dtype), dtype)


@pytest.mark.parametrize("T", [1, 2, 4, 5, 16])
@pytest.mark.parametrize("shape", [(4096, 4096, False), (1408, 4096, True), (512, 11008, False)])
def test_gemm_ticket_free_fixup_matches_ticket_path_bit_for_bit(T, shape):
    """(value, tag) word exchange (T = 1: the tile's last CTA collects; 2..16: every CTA reduces a share) vs the
    scratch + ticket path: same summation order, so the same bits; the word buffer is left all zero."""
    from petals_b200.ops import gemm, packing

    N, K, dual = shape
    dtype = torch.bfloat16
    torch.manual_seed(T + N)
    x = (torch.randn(T, K, device="cuda") * 0.5).to(dtype)
    res = torch.randn(T, N, device="cuda").to(dtype)
    g = (1 + 0.1 * torch.randn(K, device="cuda")).to(dtype)
    if dual:
        pw = packing.pack_dual((torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype),
                               (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype))
        kw = dict(norm_mode=gemm.NORM_RMS, norm_w=g)
    else:
        pw = packing.pack_weight((torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype))
        kw = dict(residual=res)
    ws_ll = gemm.GemmWorkspace.create("cuda", 16)
    assert ws_ll.ll is not None
    ws_ticket = gemm.GemmWorkspace.create("cuda", 16)
    ws_ticket.ll = None
    a = gemm.linear(x, pw, workspace=ws_ll, **kw)
    b = gemm.linear(x, pw, workspace=ws_ticket, **kw)
    a2 = gemm.linear(x, pw, workspace=ws_ll, **kw)
    torch.cuda.synchronize()
    assert torch.equal(a, b) and torch.equal(a, a2)
    assert int(ws_ll.ll.abs().sum()) == 0 and int(ws_ticket.counters.abs().sum()) == 0


@pytest.mark.parametrize("T", [1, 8, 16])
def test_rms_rstd_kernel_matches_in_kernel_statistics(T):
    from petals_b200.ops import elementwise as E, gemm, packing

    dtype = torch.float16
    torch.manual_seed(T)
    K, N = 4096, 768
    x = (torch.randn(T, K, device="cuda") * 1.7 + 0.2).to(dtype)
    g = (1 + 0.1 * torch.randn(K, device="cuda")).to(dtype)
    pw = packing.pack_weight((torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype))
    rstd = torch.zeros(32, device="cuda", dtype=torch.float32)
    E.rms_rstd(x, 1e-5, rstd)
    ref = torch.rsqrt(x.float().pow(2).mean(-1) + 1e-5)
    assert (rstd[:T] - ref).abs().max().item() < 1e-5 * ref.abs().max().item() + 1e-7
    a = gemm.linear(x, pw, norm_mode=gemm.NORM_RMS, norm_w=g)
    b = gemm.linear(x, pw, norm_mode=gemm.NORM_RMS, norm_w=g, rstd=rstd)
    torch.cuda.synchronize()
    assert torch.equal(a, b)


@pytest.mark.parametrize("T,N,K,dual", [(100, 4096, 4096, False), (128, 4096, 11008, False), (128, 2048, 4096, True),
                                        (200, 4096, 4096, False)])
def test_prefill_gemm_split_k_for_short_prompts(T, N, K, dual, monkeypatch):
    """Few output tiles for 148 SMs: the prefill GEMM splits K over several units per tile and the unit with the
    tile's last ticket adds the partials in k order.  Same result as the unsplit kernel up to fp32 summation order,
    bit-reproducible from launch to launch, counters left clean."""
    from petals_b200.ops import gemm, packing

    dtype = torch.bfloat16
    torch.manual_seed(T + N)
    dev = "cuda"
    x = (torch.randn(T, K, device=dev) * 0.5).to(dtype)
    mk = lambda: (torch.randn(N, K, device=dev) / K ** 0.5).to(dtype)  # noqa: E731
    ws = gemm.GemmWorkspace.create(dev, 256)
    if dual:
        wg, wu = mk(), mk()
        pw = packing.pack_dual(wg, wu)
        out = gemm.linear(x, pw, workspace=ws)
        ref = _ref(x, wg, epi="swiglu", up=wu, dtype=dtype)
    else:
        w = mk()
        res = torch.randn(T, N, device=dev).to(dtype)
        pw = packing.pack_weight(w)
        out = gemm.linear(x, pw, residual=res, workspace=ws)
        ref = _ref(x, w, residual=res, dtype=dtype)
    torch.cuda.synchronize()
    _close(out, ref, dtype)
    assert int(ws.counters.abs().sum()) == 0, "tickets must be handed back"
    out2 = gemm.linear(x, pw, workspace=ws) if dual else gemm.linear(x, pw, residual=res, workspace=ws)
    torch.cuda.synchronize()
    assert torch.equal(out, out2)

"""Attention / elementwise kernels vs plain torch fp32 restatements."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rope_ref(x, pos, inv_freq, dtype):
    # x: [T, heads, D] fp32 ; HF: cos/sin cast to dtype, rotate_half convention
    ang = pos[:, None].float() * inv_freq[None, :].float()            # [T, D/2]
    cos = torch.cat([ang.cos(), ang.cos()], -1).to(dtype).float()[:, None, :]
    sin = torch.cat([ang.sin(), ang.sin()], -1).to(dtype).float()[:, None, :]
    d2 = x.shape[-1] // 2
    rot = torch.cat([-x[..., d2:], x[..., :d2]], -1)
    return x * cos + rot * sin


def _attn_ref(q, k, v, n_rep):
    # q [nh, Tq, D], k/v [nkv, S, D] fp32, causal offsets handled by caller via mask
    k = k.repeat_interleave(n_rep, 0)
    v = v.repeat_interleave(n_rep, 0)
    return q @ k.transpose(1, 2)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("nh,nkv,D,rope", [(8, 8, 128, True), (8, 2, 128, True), (16, 2, 128, True),
                                           (4, 4, 64, False), (8, 1, 128, True),
                                           (32, 32, 128, True)])   # 768 CTAs: the plain-load path (small grids: bulk copies)
@pytest.mark.parametrize("past", [0, 5, 63, 64, 300, 1500])
def test_attn_decode(nh, nkv, D, rope, past, dtype):
    from petals_b200.ops import attention as A

    torch.manual_seed(past + nh)
    dev = "cuda"
    T, s_max, splits = 3, 2048, 8
    inv = A.rope_inv_freq(D, 10000.0, device=dev) if rope else None
    qkv = (torch.randn(T, (nh + 2 * nkv) * D, device=dev)).to(dtype)
    caches, refs = [], []
    pos = torch.tensor([past, past + 1, max(past - 1, 0)], dtype=torch.int32, device=dev)
    kv_ptrs = torch.zeros(T, 2, dtype=torch.int64, device=dev)
    for t in range(T):
        kc = torch.zeros(nkv, s_max, D, device=dev, dtype=dtype)
        vc = torch.zeros(nkv, s_max, D, device=dev, dtype=dtype)
        p = int(pos[t])
        kc[:, :p] = torch.randn(nkv, p, D, device=dev).to(dtype)
        vc[:, :p] = torch.randn(nkv, p, D, device=dev).to(dtype)
        caches.append((kc, vc))
        kv_ptrs[t, 0], kv_ptrs[t, 1] = kc.data_ptr(), vc.data_ptr()
    # reference
    outs = []
    for t in range(T):
        p = int(pos[t])
        row = qkv[t].float()
        q = row[: nh * D].view(1, nh, D)
        k = row[nh * D: (nh + nkv) * D].view(1, nkv, D)
        v = row[(nh + nkv) * D:].view(nkv, D)
        if rope:
            pp = torch.tensor([p], device=dev)
            q = _rope_ref(q, pp, inv, dtype).to(dtype).float()
            k = _rope_ref(k, pp, inv, dtype).to(dtype).float()
        kc, vc = caches[t]
        kfull = torch.cat([kc[:, :p].float(), k[0][:, None, :]], 1)          # [nkv, p+1, D]
        vfull = torch.cat([vc[:, :p].float(), v[:, None, :].to(dtype).float()], 1)
        rep = nh // nkv
        kk = kfull.repeat_interleave(rep, 0)
        vv = vfull.repeat_interleave(rep, 0)
        sc = (q[0][:, None, :] @ kk.transpose(1, 2)) / math.sqrt(D)          # [nh,1,S]
        pr = torch.softmax(sc, -1)
        outs.append((pr @ vv).reshape(nh * D))
        refs.append((k[0], v))
    ref = torch.stack(outs)
    out = torch.empty(T, nh * D, device=dev, dtype=dtype)
    ws = A.DecodeAttnWorkspace(dev, T, nh, nkv, D, splits)
    A.attn_decode(qkv, out, kv_ptrs, pos, inv, ws, nh, nkv, D, s_max)
    torch.cuda.synchronize()
    tol = 2e-2 if dtype == torch.bfloat16 else 3e-3
    assert (out.float() - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item())
    for t in range(T):
        p = int(pos[t])
        kc, vc = caches[t]
        assert (kc[:, p].float() - refs[t][0]).abs().max().item() < tol
        assert torch.equal(vc[:, p], refs[t][1].to(dtype))
    assert int(ws.counters.abs().sum()) == 0
    out2 = torch.empty_like(out)
    A.attn_decode(qkv, out2, kv_ptrs, pos, inv, ws, nh, nkv, D, s_max)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("nh,nkv,D,rope", [(4, 4, 128, True), (8, 2, 128, True), (4, 4, 64, False)])
@pytest.mark.parametrize("T,pos0", [(1, 0), (7, 0), (33, 0), (128, 0), (40, 100), (200, 17), (300, 130), (256, 0)])
def test_rope_append_and_prefill(nh, nkv, D, rope, T, pos0, dtype):
    from petals_b200.ops import attention as A

    torch.manual_seed(T + pos0)
    dev = "cuda"
    s_max = 1024
    inv = A.rope_inv_freq(D, 500000.0, device=dev) if rope else None
    qkv = torch.randn(T, (nh + 2 * nkv) * D, device=dev).to(dtype)
    kc = torch.zeros(nkv, s_max, D, device=dev, dtype=dtype)
    vc = torch.zeros(nkv, s_max, D, device=dev, dtype=dtype)
    kc[:, :pos0] = torch.randn(nkv, pos0, D, device=dev).to(dtype)
    vc[:, :pos0] = torch.randn(nkv, pos0, D, device=dev).to(dtype)
    rowf = qkv.float()
    q = rowf[:, : nh * D].view(T, nh, D)
    k = rowf[:, nh * D: (nh + nkv) * D].view(T, nkv, D)
    v = qkv[:, (nh + nkv) * D:].view(T, nkv, D)
    pos = torch.arange(pos0, pos0 + T, device=dev)
    if rope:
        q = _rope_ref(q, pos, inv, dtype)
        k = _rope_ref(k, pos, inv, dtype)
    q = q.to(dtype).float()
    kfull = torch.cat([kc[:, :pos0].float(), k.to(dtype).float().transpose(0, 1)], 1)
    vfull = torch.cat([vc[:, :pos0].float(), v.float().transpose(0, 1)], 1)
    rep = nh // nkv
    sc = q.transpose(0, 1) @ kfull.repeat_interleave(rep, 0).transpose(1, 2) / math.sqrt(D)   # [nh,T,S]
    S = pos0 + T
    mask = torch.arange(S, device=dev)[None, :] > pos[:, None]
    sc = sc.masked_fill(mask[None], float("-inf"))
    ref = (torch.softmax(sc, -1) @ vfull.repeat_interleave(rep, 0)).transpose(0, 1).reshape(T, nh * D)

    A.rope_append(qkv, kc, vc, inv, nh, nkv, D, s_max, pos0)
    out = torch.empty(T, nh * D, device=dev, dtype=dtype)
    A.attn_prefill(qkv, out, kc, vc, nh, nkv, D, s_max, pos0)
    torch.cuda.synchronize()
    tol = 2e-2 if dtype == torch.bfloat16 else 3e-3
    assert (kc[:, pos0:S].float() - k.to(dtype).float().transpose(0, 1)).abs().max().item() < tol
    assert torch.equal(vc[:, pos0:S], v.transpose(0, 1))
    assert (out.float() - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item())

@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("nh,nkv,D,T,pos0,qscale", [(32, 8, 128, 1100, 77, 1.0), (64, 64, 64, 600, 0, 1.0),
                                                    (8, 8, 128, 700, 300, 1.0), (32, 32, 128, 1024, 0, 6.0),
                                                    (8, 2, 128, 130, 0, 6.0)])
def test_prefill_attention_long(nh, nkv, D, T, pos0, qscale, dtype):
    """Multi-block / two-tiles-per-CTA shapes of the pipelined tcgen05 kernel; the cache tail holds NaNs
    (rows past the valid length must never reach the output) and large logits exercise the lazy rescale."""
    from petals_b200.ops import attention as A

    torch.manual_seed(nh + T)
    dev = "cuda"
    s_max = pos0 + T + 200
    S = pos0 + T
    qkv = torch.randn(T, (nh + 2 * nkv) * D, device=dev).to(dtype)
    qkv[:, : nh * D] *= qscale
    kc = torch.full((nkv, s_max, D), float("nan"), device=dev, dtype=dtype)
    vc = torch.full((nkv, s_max, D), float("nan"), device=dev, dtype=dtype)
    kc[:, :S] = torch.randn(nkv, S, D, device=dev).to(dtype)
    vc[:, :S] = torch.randn(nkv, S, D, device=dev).to(dtype)
    q = qkv[:, : nh * D].float().view(T, nh, D)
    rep = nh // nkv
    pos = torch.arange(pos0, S, device=dev)
    sc = q.transpose(0, 1) @ kc[:, :S].float().repeat_interleave(rep, 0).transpose(1, 2) / math.sqrt(D)
    mask = torch.arange(S, device=dev)[None, :] > pos[:, None]
    sc = sc.masked_fill(mask[None], float("-inf"))
    ref = (torch.softmax(sc, -1) @ vc[:, :S].float().repeat_interleave(rep, 0)).transpose(0, 1).reshape(T, nh * D)
    out = torch.empty(T, nh * D, device=dev, dtype=dtype)
    A.attn_prefill(qkv, out, kc, vc, nh, nkv, D, s_max, pos0)
    torch.cuda.synchronize()
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    assert torch.isfinite(out.float()).all()
    assert (out.float() - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_embed_norm_argmax(dtype):
    from petals_b200.ops import elementwise as E

    torch.manual_seed(0)
    dev = "cuda"
    V, H, T = 1000, 768, 9
    wte = torch.randn(V, H, device=dev).to(dtype)
    wpe = torch.randn(64, H, device=dev).to(dtype)
    ids = torch.randint(0, V, (T,), device=dev)
    out = torch.empty(T, H, device=dev, dtype=dtype)
    E.embed(ids, wte, out)
    assert torch.equal(out, wte[ids])
    E.embed(ids, wte, out, wpe=wpe, pos0=5)
    ref = (wte[ids].float() + wpe[5:5 + T].float()).to(dtype)
    assert torch.equal(out, ref)

    for H2 in (768, 4096, 5120, 8192):
        x = (torch.randn(T, H2, device=dev) * 3 + 1).to(dtype)
        g = (1 + 0.1 * torch.randn(H2, device=dev)).to(dtype)
        b = (0.1 * torch.randn(H2, device=dev)).to(dtype)
        y = E.norm(x, g, None, 1e-5, False)
        xf = x.float()
        r = (g.float() * (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)).to(dtype).float()).to(dtype)
        assert (y.float() - r.float()).abs().max().item() <= (0.07 if dtype == torch.bfloat16 else 0.01)
        y = E.norm(x, g, b, 1e-5, True)
        r = torch.nn.functional.layer_norm(xf, (H2,), g.float(), b.float(), 1e-5)
        assert (y.float() - r).abs().max().item() <= (0.07 if dtype == torch.bfloat16 else 0.01)

    logits = torch.randn(T, 50257, device=dev)
    logits[3, 777] = 100.0
    logits[4, 10] = 50.0
    logits[4, 20000] = 50.0   # tie -> first index
    ids_out = torch.zeros(T, dtype=torch.int64, device=dev)
    ws = E.ArgmaxWorkspace(dev, T)
    E.argmax(logits, ids_out, ws)
    assert torch.equal(ids_out, logits.argmax(-1))
    assert int(ids_out[4]) == 10
    E.argmax(logits, ids_out, ws)
    assert torch.equal(ids_out, logits.argmax(-1))

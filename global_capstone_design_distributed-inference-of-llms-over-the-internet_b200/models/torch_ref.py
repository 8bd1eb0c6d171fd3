"""Plain-torch restatement of the decoder blocks (causal, pre-allocated KV).

Two jobs: (1) the execution path on machines without a B200 (the CPU plumbing
config: GPT-2 2-stage over localhost), and (2) the fp32 oracle the CUDA kernels
are tested against.  Math follows HF ``LlamaDecoderLayer`` / ``GPT2Block``; unlike
the reference's live path (which passes ``attention_mask=None`` to every layer —
SURVEY B1) prefill here is causal.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from ..ops.attention import rope_inv_freq
from ..runtime.kv_cache import SessionKV
from .config import ModelConfig


def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    xf = x.float()
    xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype)
    return w * xn


def rope_cos_sin(pos: torch.Tensor, inv_freq: torch.Tensor, dtype: torch.dtype):
    ang = pos.float()[:, None] * inv_freq.float()[None, :]
    emb = torch.cat([ang, ang], -1)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x: [T, heads, D]; cos/sin: [T, D]."""
    d2 = x.shape[-1] // 2
    rot = torch.cat([-x[..., d2:], x[..., :d2]], -1)
    return x * cos[:, None, :] + rot * sin[:, None, :]


def attention_with_cache(q, k_new, v_new, kv: SessionKV, layer: int, pos0: int, scale: float) -> torch.Tensor:
    """q: [T, nh, D]; k_new/v_new: [T, nkv, D]. Appends in place, attends causally over [0, pos0+T)."""
    T, nh, D = q.shape
    nkv = k_new.shape[1]
    kc, vc = kv.k(layer), kv.v(layer)                       # [nkv, max_len, D]
    kc[:, pos0:pos0 + T] = k_new.transpose(0, 1)
    vc[:, pos0:pos0 + T] = v_new.transpose(0, 1)
    S = pos0 + T
    rep = nh // nkv
    kf = kc[:, :S].repeat_interleave(rep, 0) if rep > 1 else kc[:, :S]
    vf = vc[:, :S].repeat_interleave(rep, 0) if rep > 1 else vc[:, :S]
    scores = torch.matmul(q.transpose(0, 1), kf.transpose(1, 2)) * scale          # [nh, T, S]
    if T > 1:
        qpos = torch.arange(pos0, S, device=q.device)
        mask = torch.arange(S, device=q.device)[None, :] > qpos[:, None]
        scores = scores.masked_fill(mask[None], float("-inf"))
    probs = torch.softmax(scores.float(), dim=-1).to(q.dtype)
    return torch.matmul(probs, vf).transpose(0, 1).reshape(T, nh * D)


class _ScratchKV:
    """Cache stand-in for a forward pass that keeps nothing: fresh K/V buffers per block, sized to the sequence."""

    def __init__(self, n_kv: int, length: int, head_dim: int, dtype, device):
        self._shape, self._dtype, self._device = (n_kv, length, head_dim), dtype, device
        self._bufs: Dict[tuple, torch.Tensor] = {}

    def _buf(self, key):
        if key not in self._bufs:
            self._bufs[key] = torch.zeros(self._shape, dtype=self._dtype, device=self._device)
        return self._bufs[key]

    def k(self, layer: int) -> torch.Tensor:
        return self._buf((layer, 0))

    def v(self, layer: int) -> torch.Tensor:
        return self._buf((layer, 1))


class TorchSpan:
    """Runs decoder blocks [start, end) of a model from canonical weights with torch ops."""

    def __init__(self, cfg: ModelConfig, weights: Dict[str, object], device, dtype: torch.dtype):
        self.cfg, self.w, self.device, self.dtype = cfg, weights, torch.device(device), dtype
        self.layers: List[Dict[str, torch.Tensor]] = weights["layers"]  # type: ignore[assignment]
        self.inv_freq = rope_inv_freq(cfg.head_dim, cfg.rope_theta, cfg.rope_scaling, self.device) if cfg.is_llama else None
        self.scale = 1.0 / math.sqrt(cfg.head_dim)

    def subspan(self, a: int, b: int) -> "TorchSpan":
        """View over blocks [a, b) of this span (shared weights)."""
        import copy

        if not (0 <= a < b <= len(self.layers)):
            raise ValueError(f"blocks [{a},{b}) are not inside this span of {len(self.layers)} blocks")
        v = copy.copy(self)
        v.layers = self.layers[a:b]
        return v

    # ---- stage-0 / last-stage pieces
    def embed(self, ids: torch.Tensor, pos0: int) -> torch.Tensor:
        h = self.w["embed"][ids]                                               # type: ignore[index]
        if not self.cfg.is_llama:
            pos = torch.arange(pos0, pos0 + ids.shape[0], device=ids.device)
            h = h + self.w["pos_embed"][pos]                                   # type: ignore[index]
        return h

    def head(self, h_last: torch.Tensor) -> torch.Tensor:
        """h_last: [t, H] -> logits [t, V] (fp32); callers pass the last position only."""
        c = self.cfg
        if c.is_llama:
            hn = rms_norm(h_last, self.w["final_norm_w"], c.norm_eps)          # type: ignore[arg-type]
        else:
            hn = F.layer_norm(h_last, (c.hidden,), self.w["final_norm_w"], self.w["final_norm_b"], c.norm_eps)
        return F.linear(hn, self.w["lm_head"]).float()                         # type: ignore[arg-type]

    # ---- blocks
    def _llama_layer(self, h, L, kv, li, pos0):
        c = self.cfg
        T = h.shape[0]
        x = rms_norm(h, L["ln1_w"], c.norm_eps)
        qkv = F.linear(x, L["qkv_w"])
        q = qkv[:, : c.q_size].view(T, c.n_heads, c.head_dim)
        k = qkv[:, c.q_size: c.q_size + c.kv_size].view(T, c.n_kv_heads, c.head_dim)
        v = qkv[:, c.q_size + c.kv_size:].view(T, c.n_kv_heads, c.head_dim)
        cos, sin = rope_cos_sin(torch.arange(pos0, pos0 + T, device=h.device), self.inv_freq, h.dtype)
        q, k = apply_rope(q, cos, sin), apply_rope(k, cos, sin)
        a = attention_with_cache(q, k, v, kv, li, pos0, self.scale)
        h = h + F.linear(a, L["o_w"])
        x = rms_norm(h, L["ln2_w"], c.norm_eps)
        m = F.linear(F.silu(F.linear(x, L["gate_w"])) * F.linear(x, L["up_w"]), L["down_w"])
        return h + m

    def _gpt2_layer(self, h, L, kv, li, pos0):
        c = self.cfg
        T = h.shape[0]
        x = F.layer_norm(h, (c.hidden,), L["ln1_w"], L["ln1_b"], c.norm_eps)
        qkv = F.linear(x, L["qkv_w"], L["qkv_b"])
        q, k, v = (t.reshape(T, c.n_heads, c.head_dim) for t in qkv.split(c.hidden, dim=-1))
        a = attention_with_cache(q, k, v, kv, li, pos0, self.scale)
        h = h + F.linear(a, L["o_w"], L["o_b"])
        x = F.layer_norm(h, (c.hidden,), L["ln2_w"], L["ln2_b"], c.norm_eps)
        m = F.linear(F.gelu(F.linear(x, L["fc_w"], L["fc_b"]), approximate="tanh"), L["down_w"], L["down_b"])
        return h + m

    def block_stateless(self, h: torch.Tensor, L: Dict[str, torch.Tensor]) -> torch.Tensor:
        """One block over whole sequences ``h`` [B, T, H] from position 0, causal, nothing cached.  Differentiable:
        this is what the training-side RPCs run (upstream ``run_rpc_forward`` / ``run_rpc_backward``,
        petals/server/block_functions.py:32-139)."""
        fn = self._llama_layer if self.cfg.is_llama else self._gpt2_layer
        outs = []
        for b in range(h.shape[0]):
            kv = _ScratchKV(self.cfg.n_kv_heads, h.shape[1], self.cfg.head_dim, h.dtype, h.device)
            outs.append(fn(h[b], L, kv, 0, 0))
        return torch.stack(outs)

    def forward_stateless(self, h: torch.Tensor) -> torch.Tensor:
        """All blocks of the span over ``h`` [B, T, H], no session (autograd-friendly)."""
        for L in self.layers:
            h = self.block_stateless(h, L)
        return h

    @torch.inference_mode()
    def forward(self, h: torch.Tensor, kv: SessionKV, pos0: Optional[int] = None) -> torch.Tensor:
        """h: [T, H] hidden states of ONE session; appends T tokens at ``pos0`` (default: kv.length)."""
        pos0 = kv.length if pos0 is None else pos0
        T = h.shape[0]
        if pos0 + T > kv.max_len:
            raise ValueError(f"session {kv.session_id}: {pos0}+{T} tokens exceed max_length {kv.max_len}")
        fn = self._llama_layer if self.cfg.is_llama else self._gpt2_layer
        for li, L in enumerate(self.layers):
            h = fn(h, L, kv, li, pos0)
        kv.length = pos0 + T
        kv.touch()
        return h

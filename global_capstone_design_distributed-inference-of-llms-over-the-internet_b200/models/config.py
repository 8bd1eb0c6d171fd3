"""Architecture descriptions for the two supported families (Llama, GPT-2).

The reference accepts any HF model name but only runs Llama-family models
(src/llama_partition.py:81-83) while advertising GPT-2 (README / run_all.py
defaults).  Both are first-class here.  Shapes follow SURVEY §2.7.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field, asdict
from typing import Any, Dict, Optional


@dataclass(frozen=True)
class ModelConfig:
    family: str                       # "llama" | "gpt2"
    n_layers: int
    hidden: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    intermediate: int
    vocab: int
    norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    rope_scaling: Optional[Dict[str, Any]] = None   # {"rope_type":"llama3", factor, low/high_freq_factor, original_max_position_embeddings}
    max_positions: int = 4096
    tie_embeddings: bool = False
    name: str = ""
    bos_token_id: Optional[int] = None
    eos_token_id: Optional[int] = None

    # ---- derived
    @property
    def q_size(self) -> int:
        return self.n_heads * self.head_dim

    @property
    def kv_size(self) -> int:
        return self.n_kv_heads * self.head_dim

    @property
    def qkv_size(self) -> int:
        return self.q_size + 2 * self.kv_size

    @property
    def is_llama(self) -> bool:
        return self.family == "llama"

    def layer_param_count(self) -> int:
        h, i = self.hidden, self.intermediate
        if self.is_llama:
            return self.qkv_size * h + h * self.q_size + 3 * h * i + 2 * h
        return (self.qkv_size * h + self.qkv_size) + (h * h + h) + (h * i + i) + (i * h + h) + 4 * h

    def layer_bytes(self, elem_size: int = 2) -> int:
        return self.layer_param_count() * elem_size

    def kv_bytes_per_token(self, n_layers: int, elem_size: int = 2) -> int:
        return 2 * self.kv_size * n_layers * elem_size

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)

    def to_hf_dict(self) -> Dict[str, Any]:
        """config.json contents HF would accept for this architecture."""
        if self.is_llama:
            d = dict(model_type="llama", architectures=["LlamaForCausalLM"], hidden_size=self.hidden,
                     intermediate_size=self.intermediate, num_hidden_layers=self.n_layers,
                     num_attention_heads=self.n_heads, num_key_value_heads=self.n_kv_heads,
                     head_dim=self.head_dim, vocab_size=self.vocab, rms_norm_eps=self.norm_eps,
                     rope_theta=self.rope_theta, max_position_embeddings=self.max_positions,
                     tie_word_embeddings=self.tie_embeddings, hidden_act="silu",
                     attention_bias=False, mlp_bias=False)
            if self.rope_scaling:
                d["rope_scaling"] = dict(self.rope_scaling)
            return d
        return dict(model_type="gpt2", architectures=["GPT2LMHeadModel"], n_embd=self.hidden,
                    n_inner=self.intermediate, n_layer=self.n_layers, n_head=self.n_heads,
                    vocab_size=self.vocab, layer_norm_epsilon=self.norm_eps,
                    n_positions=self.max_positions, n_ctx=self.max_positions,
                    activation_function="gelu_new")


def from_hf_dict(d: Dict[str, Any], name: str = "") -> ModelConfig:
    mt = str(d.get("model_type", "")).lower()
    if mt == "gpt2":
        h = int(d["n_embd"])
        nh = int(d["n_head"])
        return ModelConfig("gpt2", int(d["n_layer"]), h, nh, nh, h // nh,
                           int(d.get("n_inner") or 4 * h), int(d["vocab_size"]),
                           float(d.get("layer_norm_epsilon", 1e-5)), max_positions=int(d.get("n_positions", 1024)),
                           tie_embeddings=True, name=name,
                           bos_token_id=d.get("bos_token_id"), eos_token_id=d.get("eos_token_id"))
    if any(t in mt for t in ("llama", "mistral")):
        h = int(d["hidden_size"])
        nh = int(d["num_attention_heads"])
        nkv = int(d.get("num_key_value_heads") or nh)
        eos = d.get("eos_token_id")
        if isinstance(eos, (list, tuple)):
            eos = eos[0]
        return ModelConfig("llama", int(d["num_hidden_layers"]), h, nh, nkv,
                           int(d.get("head_dim") or h // nh), int(d["intermediate_size"]),
                           int(d["vocab_size"]), float(d.get("rms_norm_eps", 1e-5)),
                           float(d.get("rope_theta", 10000.0)), d.get("rope_scaling"),
                           int(d.get("max_position_embeddings", 4096)),
                           bool(d.get("tie_word_embeddings", False)), name,
                           d.get("bos_token_id"), eos)
    raise ValueError(f"unsupported model_type {mt!r} (supported: gpt2, llama, mistral)")


def _llama(name, L, H, nh, nkv, I, V, theta=10000.0, maxpos=4096, scaling=None, eps=1e-5):
    return ModelConfig("llama", L, H, nh, nkv, 128 if H >= 2048 else H // nh, I, V, eps, theta, scaling, maxpos, False,
                       name, 1, 2)


_LLAMA3_SCALING = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                   "original_max_position_embeddings": 8192}

PRESETS: Dict[str, ModelConfig] = {
    "gpt2": ModelConfig("gpt2", 12, 768, 12, 12, 64, 3072, 50257, 1e-5, max_positions=1024,
                        tie_embeddings=True, name="gpt2", bos_token_id=50256, eos_token_id=50256),
    "gpt2-medium": ModelConfig("gpt2", 24, 1024, 16, 16, 64, 4096, 50257, 1e-5, max_positions=1024,
                               tie_embeddings=True, name="gpt2-medium", bos_token_id=50256, eos_token_id=50256),
    "llama-2-7b": _llama("llama-2-7b", 32, 4096, 32, 32, 11008, 32000),
    "llama-2-13b": _llama("llama-2-13b", 40, 5120, 40, 40, 13824, 32000),
    "llama-2-70b": _llama("llama-2-70b", 80, 8192, 64, 8, 28672, 32000),
    "llama-3-8b": _llama("llama-3-8b", 32, 4096, 32, 8, 14336, 128256, theta=500000.0, maxpos=8192),
    "llama-3.1-8b": _llama("llama-3.1-8b", 32, 4096, 32, 8, 14336, 128256, theta=500000.0, maxpos=131072,
                           scaling=_LLAMA3_SCALING),
    # tiny shapes for tests (head_dim 128 keeps the native kernels' fast paths exercised)
    "tiny-llama": ModelConfig("llama", 4, 256, 2, 2, 128, 512, 1024, 1e-5, 10000.0, None, 512, False, "tiny-llama", 1, 2),
    "tiny-llama-gqa": ModelConfig("llama", 4, 512, 4, 2, 128, 768, 1024, 1e-5, 500000.0, None, 512, False,
                                  "tiny-llama-gqa", 1, 2),
    # shardable 8 ways (heads, kv heads, 64-wide k-blocks of the MLP): tensor-parallel tests at world 4 / 8
    "tiny-llama-tp8": ModelConfig("llama", 2, 1024, 8, 8, 128, 2048, 1024, 1e-5, 10000.0, None, 512, False,
                                  "tiny-llama-tp8", 1, 2),
    "tiny-gpt2": ModelConfig("gpt2", 4, 128, 2, 2, 64, 512, 1000, 1e-5, max_positions=256, tie_embeddings=True,
                             name="tiny-gpt2", bos_token_id=999, eos_token_id=999),
}

_ALIASES = {
    "gpt2": "gpt2", "openai-community/gpt2": "gpt2", "gpt2-medium": "gpt2-medium",
    "meta-llama/llama-2-7b-hf": "llama-2-7b", "meta-llama/llama-2-7b-chat-hf": "llama-2-7b",
    "meta-llama/llama-2-13b-hf": "llama-2-13b", "meta-llama/llama-2-70b-hf": "llama-2-70b",
    "meta-llama/meta-llama-3-8b": "llama-3-8b", "meta-llama/llama-3-8b": "llama-3-8b",
    "meta-llama/llama-3.1-8b": "llama-3.1-8b", "meta-llama/llama-3.1-8b-instruct": "llama-3.1-8b",
}


def resolve_config(model: str) -> ModelConfig:
    """``model`` is a local HF directory (config.json), a preset, or a known hub id."""
    if os.path.isdir(model) and os.path.exists(os.path.join(model, "config.json")):
        with open(os.path.join(model, "config.json")) as f:
            return from_hf_dict(json.load(f), name=model)
    key = model.lower()
    if key.startswith("random:"):
        key = key[len("random:"):]
    key = _ALIASES.get(key, key)
    if key in PRESETS:
        return PRESETS[key]
    raise ValueError(
        f"unknown model {model!r}: pass a local HF checkpoint directory or one of {sorted(PRESETS)} "
        "(no network: hub downloads are unavailable)"
    )


def local_checkpoint_dir(model: str) -> Optional[str]:
    """Directory holding real weights for ``model`` or None (=> seeded random init)."""
    if os.path.isdir(model):
        for f in ("model.safetensors.index.json", "model.safetensors", "pytorch_model.bin.index.json",
                  "pytorch_model.bin"):
            if os.path.exists(os.path.join(model, f)):
                return model
    return None

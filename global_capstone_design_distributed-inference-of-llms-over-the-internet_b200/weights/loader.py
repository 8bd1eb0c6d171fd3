"""Span-selective weight loading into the framework's canonical layout.

Canonical per-layer tensors (all ``[out, in]`` like torch ``Linear``; GPT-2's
``Conv1D`` weights are transposed once here):

    ln1_w [ln1_b]  qkv_w [qkv_b]  o_w [o_b]  ln2_w [ln2_b]
    llama: gate_w up_w down_w          gpt2: fc_w fc_b down_w down_b

plus ``embed`` (``[V,H]``), ``pos_embed`` (GPT-2), ``final_norm_w/_b`` and
``lm_head`` (tied to ``embed`` for GPT-2) for the first/last stage.

Sources: a local HF checkpoint directory (safetensors or .bin, sharded or not;
only shards holding the requested blocks are opened), or — because the build box
has no network — a seeded random initialisation that every process reproduces
bit-identically from the tensor *name*, so that independently started stages of
one pipeline agree on the model without exchanging files.

Parity: src/llama_partition.py:477-550 (load_stage_model / _prune_layers) and the
per-block loader petals/server/from_pretrained.py:35-224.
"""
from __future__ import annotations

import hashlib
from typing import Callable, Dict, List, Optional

import torch

from ..models.config import ModelConfig, local_checkpoint_dir
from .safetensors_io import CheckpointIndex

TensorGetter = Callable[[str], torch.Tensor]


# ---------------------------------------------------------------- HF naming
def hf_layer_names(cfg: ModelConfig, i: int) -> Dict[str, str]:
    if cfg.is_llama:
        p = f"model.layers.{i}."
        return {"ln1_w": p + "input_layernorm.weight", "q_w": p + "self_attn.q_proj.weight",
                "k_w": p + "self_attn.k_proj.weight", "v_w": p + "self_attn.v_proj.weight",
                "o_w": p + "self_attn.o_proj.weight", "ln2_w": p + "post_attention_layernorm.weight",
                "gate_w": p + "mlp.gate_proj.weight", "up_w": p + "mlp.up_proj.weight",
                "down_w": p + "mlp.down_proj.weight"}
    p = f"h.{i}."
    return {"ln1_w": p + "ln_1.weight", "ln1_b": p + "ln_1.bias", "c_attn_w": p + "attn.c_attn.weight",
            "c_attn_b": p + "attn.c_attn.bias", "c_proj_w": p + "attn.c_proj.weight",
            "c_proj_b": p + "attn.c_proj.bias", "ln2_w": p + "ln_2.weight", "ln2_b": p + "ln_2.bias",
            "fc_w": p + "mlp.c_fc.weight", "fc_b": p + "mlp.c_fc.bias", "proj_w": p + "mlp.c_proj.weight",
            "proj_b": p + "mlp.c_proj.bias"}


def hf_global_names(cfg: ModelConfig) -> Dict[str, str]:
    if cfg.is_llama:
        return {"embed": "model.embed_tokens.weight", "final_norm_w": "model.norm.weight", "lm_head": "lm_head.weight"}
    return {"embed": "wte.weight", "pos_embed": "wpe.weight", "final_norm_w": "ln_f.weight",
            "final_norm_b": "ln_f.bias"}


def hf_shapes(cfg: ModelConfig) -> Dict[str, tuple]:
    """Shape of every HF tensor of the architecture (drives random init and tests)."""
    H, I, V = cfg.hidden, cfg.intermediate, cfg.vocab
    out: Dict[str, tuple] = {}
    for i in range(cfg.n_layers):
        n = hf_layer_names(cfg, i)
        if cfg.is_llama:
            out.update({n["ln1_w"]: (H,), n["q_w"]: (cfg.q_size, H), n["k_w"]: (cfg.kv_size, H),
                        n["v_w"]: (cfg.kv_size, H), n["o_w"]: (H, cfg.q_size), n["ln2_w"]: (H,),
                        n["gate_w"]: (I, H), n["up_w"]: (I, H), n["down_w"]: (H, I)})
        else:
            out.update({n["ln1_w"]: (H,), n["ln1_b"]: (H,), n["c_attn_w"]: (H, 3 * H), n["c_attn_b"]: (3 * H,),
                        n["c_proj_w"]: (H, H), n["c_proj_b"]: (H,), n["ln2_w"]: (H,), n["ln2_b"]: (H,),
                        n["fc_w"]: (H, I), n["fc_b"]: (I,), n["proj_w"]: (I, H), n["proj_b"]: (H,)})
    g = hf_global_names(cfg)
    out[g["embed"]] = (V, H)
    out[g["final_norm_w"]] = (H,)
    if cfg.is_llama:
        if not cfg.tie_embeddings:
            out[g["lm_head"]] = (V, H)
    else:
        out[g["pos_embed"]] = (cfg.max_positions, H)
        out[g["final_norm_b"]] = (H,)
    return out


# ---------------------------------------------------------------- sources
def _name_seed(name: str, seed: int) -> int:
    return int.from_bytes(hashlib.sha256(f"{seed}:{name}".encode()).digest()[:7], "little")


class RandomSource:
    """Deterministic random tensors keyed by (seed, HF tensor name), the SAME values on every device.

    Small tensors come from the CPU generator.  Big ones (a 7B model would take minutes there) are produced on the
    target device from a counter-based generator written with integer tensor ops only (a 32-bit mixing hash of the
    element index and the name seed, four uniforms summed per element), so a CPU oracle, a ``--use_cpu_offload`` stage
    and GPU stages of one pipeline agree on the model bit for bit without exchanging files (torch's own CUDA and CPU
    generators do not produce the same stream).
    """

    CPU_LIMIT = 1 << 22

    def __init__(self, cfg: ModelConfig, seed: int = 0, device: Optional[torch.device] = None):
        self.cfg, self.seed = cfg, seed
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.shapes = hf_shapes(cfg)

    def has(self, name: str) -> bool:
        return name in self.shapes

    @staticmethod
    def _hash32(x: torch.Tensor) -> torch.Tensor:
        """lowbias32-style integer mix on int64 tensors holding 32-bit values."""
        M = 0xFFFFFFFF
        x = (x ^ (x >> 16)) & M
        x = (x * 0x7FEB352D) & M
        x = (x ^ (x >> 15)) & M
        x = (x * 0x846CA68B) & M
        return (x ^ (x >> 16)) & M

    def _counter_randn(self, numel: int, key: int, device) -> torch.Tensor:
        """Approximately normal values from integer arithmetic only (bit-identical on every device): four 24-bit
        uniforms per element from a mixing hash of (element index, key), summed (Irwin-Hall), centred and scaled to
        unit variance.  No transcendental function, so no device-dependent rounding."""
        out = torch.empty(numel, dtype=torch.float32, device=device)
        step = 1 << 24
        k1, k2 = key & 0xFFFFFFFF, (key >> 32) & 0x7FFFFF
        for lo in range(0, numel, step):
            idx = torch.arange(lo, min(numel, lo + step), dtype=torch.int64, device=device)
            h = self._hash32((idx & 0xFFFFFFFF) ^ k1 ^ (((idx >> 32) * 0x9E3779B1) & 0xFFFFFFFF))
            total = torch.zeros_like(idx)
            for j in range(4):
                h = self._hash32((h + 0x6A09E667 + k2 + j) & 0xFFFFFFFF)
                total += h >> 8                                   # 24 random bits
            # sum of four U[0, 2^24): mean 2 * 2^24, variance 4 * 2^48 / 12
            out[lo:lo + idx.numel()] = (total - (2 << 24)).to(torch.float32) * (3.0 ** 0.5 / float(1 << 24))
        return out

    def get(self, name: str) -> torch.Tensor:
        shape = self.shapes[name]
        numel = 1
        for s in shape:
            numel *= s
        is_norm_w = name.endswith(("layernorm.weight", "norm.weight", "ln_1.weight", "ln_2.weight", "ln_f.weight"))
        if numel > self.CPU_LIMIT:
            t = self._counter_randn(numel, _name_seed(name, self.seed), self.device).view(shape)
        else:
            g = torch.Generator(device="cpu")
            g.manual_seed(_name_seed(name, self.seed))
            t = torch.randn(shape, generator=g, dtype=torch.float32)
        if is_norm_w:
            t = 1.0 + 0.05 * t
        else:
            t = 0.02 * t
        return t


class CheckpointSource:
    def __init__(self, directory: str):
        self.index = CheckpointIndex(directory)
        names = set(self.index.names())
        self.prefix = "transformer." if any(n.startswith("transformer.") for n in names) else ""

    def has(self, name: str) -> bool:
        return self.index.has(self.prefix + name) or self.index.has(name)

    def get(self, name: str) -> torch.Tensor:
        if self.index.has(self.prefix + name):
            return self.index.get(self.prefix + name)
        return self.index.get(name)


def open_source(model: str, cfg: ModelConfig, seed: int = 0, device=None):
    d = local_checkpoint_dir(model)
    return CheckpointSource(d) if d else RandomSource(cfg, seed, device)


# ---------------------------------------------------------------- canonical
def load_layer(cfg: ModelConfig, src, i: int, dtype: torch.dtype, device) -> Dict[str, torch.Tensor]:
    n = hf_layer_names(cfg, i)
    mv = lambda t: t.to(device=device, dtype=dtype).contiguous()  # noqa: E731
    if cfg.is_llama:
        qkv = torch.cat([src.get(n["q_w"]).to(device), src.get(n["k_w"]).to(device), src.get(n["v_w"]).to(device)], 0)
        return {"ln1_w": mv(src.get(n["ln1_w"])), "qkv_w": mv(qkv), "o_w": mv(src.get(n["o_w"])),
                "ln2_w": mv(src.get(n["ln2_w"])), "gate_w": mv(src.get(n["gate_w"])),
                "up_w": mv(src.get(n["up_w"])), "down_w": mv(src.get(n["down_w"]))}
    tr = lambda t: mv(t.to(device).t())  # Conv1D [in,out] -> Linear [out,in]  # noqa: E731
    return {"ln1_w": mv(src.get(n["ln1_w"])), "ln1_b": mv(src.get(n["ln1_b"])),
            "qkv_w": tr(src.get(n["c_attn_w"])), "qkv_b": mv(src.get(n["c_attn_b"])),
            "o_w": tr(src.get(n["c_proj_w"])), "o_b": mv(src.get(n["c_proj_b"])),
            "ln2_w": mv(src.get(n["ln2_w"])), "ln2_b": mv(src.get(n["ln2_b"])),
            "fc_w": tr(src.get(n["fc_w"])), "fc_b": mv(src.get(n["fc_b"])),
            "down_w": tr(src.get(n["proj_w"])), "down_b": mv(src.get(n["proj_b"]))}


def load_span(model: str, cfg: ModelConfig, start: int, end: int, *, need_embed: bool, need_head: bool,
              dtype: torch.dtype, device, seed: int = 0, src=None) -> Dict[str, object]:
    """Weights for decoder blocks [start, end) (+ embedding / final norm + head)."""
    if not (0 <= start <= end <= cfg.n_layers):
        raise ValueError(f"bad span [{start}, {end}) for a {cfg.n_layers}-layer model")
    src = src or open_source(model, cfg, seed, device)
    out: Dict[str, object] = {"layers": [load_layer(cfg, src, i, dtype, device) for i in range(start, end)],
                              "start": start, "end": end, "has_embed": bool(need_embed),
                              "has_head": bool(need_head)}
    g = hf_global_names(cfg)
    mv = lambda t: t.to(device=device, dtype=dtype).contiguous()  # noqa: E731
    if need_embed or (need_head and (cfg.tie_embeddings or not src.has(g.get("lm_head", "")))):
        out["embed"] = mv(src.get(g["embed"]))
    if need_embed and not cfg.is_llama:
        out["pos_embed"] = mv(src.get(g["pos_embed"]))
    if need_head:
        out["final_norm_w"] = mv(src.get(g["final_norm_w"]))
        if not cfg.is_llama:
            out["final_norm_b"] = mv(src.get(g["final_norm_b"]))
        if cfg.is_llama and not cfg.tie_embeddings and src.has(g["lm_head"]):
            out["lm_head"] = mv(src.get(g["lm_head"]))
        else:
            out["lm_head"] = out["embed"]
    return out


def full_hf_state_dict(cfg: ModelConfig, seed: int = 0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """Every HF tensor of a randomly initialised model (tests / checkpoint export)."""
    src = RandomSource(cfg, seed)
    return {k: src.get(k).to(dtype) for k in src.shapes}

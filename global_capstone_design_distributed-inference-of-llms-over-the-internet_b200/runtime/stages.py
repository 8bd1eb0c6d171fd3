"""Stage objects: a layer span + its role (first / middle / last) behind one ``forward``.

Mirrors the reference's three classes (src/llama_partition.py:76-474):

    Stage0(full, end)                    embed + blocks [0, end)
    StageSegment(full, start, end, ...)  blocks [start, end)
    StageLast(full, start, ...)          blocks [start, L) + final norm + LM head

``forward(x, position_ids, attention_mask, past_key_values, use_cache) -> (out, past)`` keeps
the reference's argument list, with these deliberate differences: attention is always causal
(``attention_mask`` is accepted and ignored — SURVEY B1), ``past_key_values`` is the session's
pre-allocated :class:`SessionKV` (appended in place and returned as ``past``), and the last
stage computes logits for the final position only (B9).  ``full`` is the :class:`StageWeights`
returned by ``load_stage_model`` — only the requested span is ever read (the reference loads
the whole model and prunes it).

Execution back-end: NativeSpan (sm_100a kernels) for fp16/bf16 on a CUDA device, TorchSpan
otherwise (CPU plumbing config, fp32).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

from ..models.config import ModelConfig, resolve_config
from ..models.torch_ref import TorchSpan
from ..weights.loader import load_span
from .kv_cache import KVCacheManager, SessionKV


@dataclass
class StageWeights:
    cfg: ModelConfig
    weights: Dict[str, object]
    role: str
    start: int
    end: int
    device: torch.device
    dtype: torch.dtype
    use_cpu_offload: bool = False
    model_name: str = ""

    @property
    def config(self) -> ModelConfig:   # reference code reads `.config`
        return self.cfg


def load_stage_weights(model_name: str, device, role: str, *, start: int = 0, end: Optional[int] = None,
                       dtype: torch.dtype = torch.float16, use_cpu_offload: bool = False, seed: int = 0,
                       families=("llama", "gpt2")) -> StageWeights:
    cfg = resolve_config(model_name)
    if cfg.family not in families:
        raise ValueError(f"{model_name}: family {cfg.family!r} is not handled by this module ({families})")
    role = {"stage0": "stage0", "segment": "segment", "last": "last"}.get(role)
    if role is None:
        raise ValueError("role must be one of stage0 / segment / last")
    if role == "stage0":
        start, end = 0, int(end if end is not None else cfg.n_layers)
    elif role == "last":
        start, end = int(start), cfg.n_layers
    else:
        start, end = int(start), int(end if end is not None else cfg.n_layers)
    if role != "stage0" and start >= end:
        raise ValueError(f"empty layer span [{start}, {end}) for {model_name} ({cfg.n_layers} layers)")
    load_dev = torch.device("cpu") if use_cpu_offload else torch.device(device)
    w = load_span(model_name, cfg, start, end, need_embed=(role == "stage0"), need_head=(role == "last"),
                  dtype=dtype, device=load_dev, seed=seed)
    return StageWeights(cfg, w, role, start, end, torch.device(device), dtype, use_cpu_offload, model_name)


class _StageBase(torch.nn.Module):
    role = "segment"

    def __init__(self, full: StageWeights, start: int, end: int, gpu_device=None, keep_layers_on_gpu: int = 0,
                 kv_capacity: int = 2048, max_batch: int = 8, backend: Optional[str] = None,
                 weight_format: str = "native"):
        super().__init__()
        if (start, end) != (full.start, full.end):
            raise ValueError(f"stage span [{start},{end}) does not match loaded weights [{full.start},{full.end})")
        self.cfg, self.start, self.end = full.cfg, start, end
        self.config = full.cfg
        self.device = torch.device(gpu_device) if gpu_device is not None else full.device
        self.dtype = full.dtype
        self.keep_layers_on_gpu = keep_layers_on_gpu
        self.use_cpu_offload = full.use_cpu_offload
        self.kv_capacity = kv_capacity
        native_ok = self.device.type == "cuda" and self.dtype in (torch.float16, torch.bfloat16)
        self.backend = backend or ("native" if native_ok else "torch")
        if self.backend == "native" and full.use_cpu_offload:
            # weights stay in pinned host memory (packed), two device slots stream them under the sm_100a kernels
            from .offload import make_offloaded_native_span

            self.span = make_offloaded_native_span(self.cfg, full.weights, self.device, self.dtype,
                                                   keep_layers_on_gpu=keep_layers_on_gpu, kv_capacity=kv_capacity,
                                                   max_batch=max_batch)
        elif self.backend == "native":
            from .native_span import NativeSpan

            self.span = NativeSpan(self.cfg, full.weights, self.device, self.dtype, kv_capacity=kv_capacity,
                                   max_batch=max_batch, weight_format=weight_format)
        elif full.use_cpu_offload and self.device.type == "cuda":
            from .offload import OffloadedSpan

            self.span = OffloadedSpan(self.cfg, full.weights, self.device, self.dtype,
                                      keep_layers_on_gpu=keep_layers_on_gpu)
        else:
            w = full.weights
            if self.device != full.device or full.use_cpu_offload:
                w = _to_device(w, self.device)
            self.span = TorchSpan(self.cfg, w, self.device, self.dtype)

    def subspan(self, start: int, end: int) -> "_StageBase":
        """Stage object over blocks [start, end) of THIS server's span (absolute block numbers): the same weights, own
        graphs; a view that stops before the model's last block is a segment even on a last-stage server."""
        import copy

        if (start, end) == (self.start, self.end):
            return self
        if not (self.start <= start < end <= self.end) or self.role == "stage0":
            raise ValueError(f"blocks [{start},{end}) are not inside this server's span [{self.start},{self.end})")
        v = copy.copy(self)
        v.span = self.span.subspan(start - self.start, end - self.start)
        v.start, v.end = start, end
        if self.role == "last" and end != self.end:
            v.__class__ = StageSegment
        return v

    # ---- training-side passes: whole sequences, no session cache (upstream rpc_forward / rpc_backward)
    def _dense_block(self, li: int):
        """(helper TorchSpan, dense weight dict) of block ``li``.  A native span keeps only its packed tiles, so the
        block is un-tiled for the moment it is needed - one block's worth of extra memory, never the span's."""
        if self.role == "stage0":
            raise ValueError("stateless passes are served by server stages")
        if isinstance(self.span, TorchSpan):
            return self.span, self.span.layers[li]
        from ..ops import packing

        span = self.span
        if getattr(span, "weight_format", "native") != "native" or getattr(span, "streamed_layers", 0) or \
                getattr(span, "tp_mode", None) is not None:
            raise NotImplementedError("stateless forward / backward need a resident 16-bit single-GPU span")
        if getattr(self, "_dense_helper", None) is None:
            self._dense_helper = TorchSpan(self.cfg, {"layers": []}, self.device, self.dtype)
        P = span.layers[li]
        L = {"ln1_w": P["ln1_w"], "ln2_w": P["ln2_w"], "qkv_w": packing.unpack_weight(P["qkv"]),
             "o_w": packing.unpack_weight(P["o"]), "down_w": packing.unpack_weight(P["down"])}
        if self.cfg.is_llama:
            gu = packing.unpack_weight(P["mlp_in"])
            L["gate_w"], L["up_w"] = gu[0], gu[1]
        else:
            L.update(ln1_b=P["ln1_b"], ln2_b=P["ln2_b"], qkv_b=P["qkv"].bias, o_b=P["o"].bias,
                     fc_w=packing.unpack_weight(P["mlp_in"]), fc_b=P["mlp_in"].bias, down_b=P["down"].bias)
        return self._dense_helper, L

    def _dense_blocks(self):
        for li in range(len(self.span.layers)):
            yield self._dense_block(li)

    @staticmethod
    def _add_prompts(h: torch.Tensor, prompts: Optional[torch.Tensor], li: int) -> torch.Tensor:
        """Deep prompt tuning (upstream block_functions.py:55-61): the client's trainable prefix of block ``li`` is
        ADDED to the first ``prefix_len`` positions of that block's input."""
        if prompts is None:
            return h
        p = prompts[li].to(h.device, h.dtype)
        pre = p.shape[-2]
        return torch.cat([h[:, :pre] + p, h[:, pre:]], dim=1)

    def _check_prompts(self, prompts: Optional[torch.Tensor], h: torch.Tensor) -> Optional[torch.Tensor]:
        if prompts is None:
            return None
        n = len(self.span.layers)
        if prompts.dim() != 4 or prompts.shape[0] != n or prompts.shape[1] != h.shape[0] or \
                prompts.shape[2] > h.shape[1] or prompts.shape[3] != h.shape[2]:
            raise ValueError(f"prompts must be [blocks={n}, batch={h.shape[0]}, prefix<={h.shape[1]}, {h.shape[2]}], "
                             f"got {tuple(prompts.shape)}")
        return prompts

    def forward_stateless(self, hidden: torch.Tensor, prompts: Optional[torch.Tensor] = None) -> torch.Tensor:
        """hidden [B, T, H] -> the span's output [B, T, H]; every sequence starts at position 0, nothing is cached.
        ``prompts`` [blocks, B, prefix, H]: deep prompts, one per block of the span."""
        h = hidden.reshape(-1, hidden.shape[-2], self.cfg.hidden).to(self.device, self.dtype)
        prompts = self._check_prompts(prompts, h)
        with torch.no_grad():
            for li, (helper, L) in enumerate(self._dense_blocks()):
                h = helper.block_stateless(self._add_prompts(h, prompts, li), L)
        return h

    def backward(self, hidden: torch.Tensor, grad_out: torch.Tensor, prompts: Optional[torch.Tensor] = None):
        """Gradient of the span's output w.r.t. its input: re-runs the forward block by block (inputs kept, weights
        un-tiled per block) and back-propagates ``grad_out`` [B, T, H] through it.  With ``prompts`` the result is
        ``(grad_input, grad_prompts)``."""
        h = hidden.reshape(-1, hidden.shape[-2], self.cfg.hidden).to(self.device, self.dtype)
        g = grad_out.reshape(h.shape).to(self.device, self.dtype)
        prompts = self._check_prompts(prompts, h)
        inputs = []
        with torch.no_grad():
            for li, (helper, L) in enumerate(self._dense_blocks()):
                inputs.append(h)
                h = helper.block_stateless(self._add_prompts(h, prompts, li), L)
        grad_prompts = [] if prompts is not None else None
        for li in reversed(range(len(inputs))):
            helper, L = self._dense_block(li)         # (native span: un-tiled again, only this block)
            x = inputs[li].detach().requires_grad_(True)
            with torch.enable_grad():
                y = helper.block_stateless(self._add_prompts(x, prompts, li), L)
            (g,) = torch.autograd.grad(y, x, g)
            if prompts is not None:
                grad_prompts.append(g[:, : prompts.shape[2]].clone())      # d/d(prompt) = d/d(input) on the prefix
        if prompts is not None:
            return g, torch.stack(grad_prompts[::-1])
        return g

    # ---- KV helpers (the handler owns the manager; tests may let the stage own one)
    def new_kv_manager(self, **kw) -> KVCacheManager:
        return KVCacheManager(self.end - self.start, self.cfg.n_kv_heads, self.cfg.head_dim, self.dtype,
                              self.device, default_max_length=self.kv_capacity, **kw)

    @staticmethod
    def _pos0(position_ids, kv: SessionKV) -> int:
        if position_ids is None:
            return kv.length
        return int(position_ids.reshape(-1)[0])

    def _run_blocks(self, h: torch.Tensor, kv: SessionKV, pos0: int) -> torch.Tensor:
        if pos0 != kv.length:
            kv.rollback(pos0)                      # start_from_position semantics
        return self.span.forward(h, kv, pos0)

    @torch.inference_mode()
    def decode_batch(self, hidden: torch.Tensor, kvs, to_cpu: bool = False) -> list:
        """One token for each of several SESSIONS in one step: ``hidden`` [B, H] (row b belongs to ``kvs[b]``) ->
        list of B outputs shaped like :meth:`forward`'s for a single token ([1, 1, H] hidden states, or
        [1, 1, V] fp32 logits on the last stage).  The native engine runs the B sessions as one batched
        kernel sequence; the torch engine falls back to a loop."""
        B = hidden.shape[0]
        if self.role == "stage0":
            raise ValueError("decode_batch serves server stages (hidden-state inputs)")
        self.last_greedy_ids = None
        if self.backend == "native" and B <= self.span.max_batch:
            h = hidden.reshape(B, self.cfg.hidden).to(self.device, self.dtype)
            if self.role == "last":
                logits, ids = self.span.decode(list(kvs), h_in=h, greedy=True)
                out = logits[:B].clone()
                # the step's graph already took the arg-max of every row: one copy for the whole batch instead of one
                # arg-max + sync per request in the handler
                self.last_greedy_ids = [int(t) for t in ids[:B].tolist()]
            else:
                out = self.span.decode(list(kvs), h_in=h).clone()
                if to_cpu:
                    out = out.cpu()                    # one device-to-host copy for the whole batch
            return [out[b].reshape(1, 1, -1) for b in range(B)]
        return [self.forward(hidden[b:b + 1].reshape(1, 1, -1), None, None, kvs[b])[0] for b in range(B)]

    def parameters(self, recurse: bool = True):   # reference probes next(stage.parameters()).dtype
        yield torch.empty(0, dtype=self.dtype, device=self.device)


def _to_device(w, device):
    out = {}
    for k, v in w.items():
        if k == "layers":
            out[k] = [{kk: vv.to(device) for kk, vv in L.items()} for L in v]
        elif torch.is_tensor(v):
            out[k] = v.to(device)
        else:
            out[k] = v
    return out


class Stage0(_StageBase):
    role = "stage0"

    def __init__(self, full: StageWeights, end: int, **kw):
        super().__init__(full, 0, end, **kw)

    @torch.inference_mode()
    def forward(self, input_ids, position_ids=None, attention_mask=None, past_key_values: SessionKV = None,
                use_cache: bool = True) -> Tuple[torch.Tensor, SessionKV]:
        kv = past_key_values
        ids = input_ids.reshape(-1).to(self.device)
        pos0 = self._pos0(position_ids, kv)
        if pos0 != kv.length:
            kv.rollback(pos0)
        if self.backend == "native":
            if self.end == 0:
                raise ValueError("native Stage0 needs at least one block")
            if ids.numel() == 1:
                h = self.span.decode([kv], ids=ids).clone()
            else:
                h = self.span.prefill(kv, ids=ids, pos0=pos0)
        else:
            h = self.span.embed(ids, pos0)
            h = self.span.forward(h, kv, pos0) if self.end > 0 else h
            if self.end == 0:
                kv.length = pos0 + ids.numel()
        return h.unsqueeze(0), kv


class StageSegment(_StageBase):
    role = "segment"

    @torch.inference_mode()
    def forward(self, hidden_states, position_ids=None, attention_mask=None, past_key_values: SessionKV = None,
                use_cache: bool = True) -> Tuple[torch.Tensor, SessionKV]:
        kv = past_key_values
        h = hidden_states.reshape(-1, self.cfg.hidden).to(self.device, self.dtype)
        h = self._run_blocks(h, kv, self._pos0(position_ids, kv))
        return h.unsqueeze(0), kv


class StageLast(_StageBase):
    role = "last"

    def __init__(self, full: StageWeights, start: int, **kw):
        super().__init__(full, start, full.cfg.n_layers, **kw)

    @torch.inference_mode()
    def forward(self, hidden_states, position_ids=None, attention_mask=None, past_key_values: SessionKV = None,
                use_cache: bool = True) -> Tuple[torch.Tensor, SessionKV]:
        """-> (logits [1, 1, V] fp32 of the LAST position, kv)."""
        kv = past_key_values
        h = hidden_states.reshape(-1, self.cfg.hidden).to(self.device, self.dtype)
        pos0 = self._pos0(position_ids, kv)
        if pos0 != kv.length:
            kv.rollback(pos0)
        if self.backend == "native":
            if h.shape[0] == 1:
                logits, _ = self.span.decode([kv], h_in=h, greedy=True)
            else:
                logits, _ = self.span.prefill(kv, h_in=h, pos0=pos0)
            logits = logits[:1].clone()
        else:
            h = self.span.forward(h, kv, pos0)
            logits = self.span.head(h[-1:])
        return logits.unsqueeze(0), kv

"""Tensor-parallel execution of a layer span across the GPUs of one NVSwitch box.

When a (load-balanced) span is wider than one GPU it is split Megatron-style:

    QKV, gate/up  column-parallel  (each rank holds n_heads/R heads, I/R MLP columns; input replicated)
    o_proj, down  row-parallel     (each rank holds a K shard; outputs must be summed)
    KV cache      by kv head       (each rank caches its own heads; nothing to communicate)

The sum after the row-parallel linears is where the reference's (dead) TP path calls an NCCL
all-reduce (petals/server/backend.py:134-136 through the external ``tensor_parallel`` package).
Here it is fused into the GEMM: ``b200_gemm_allreduce`` exchanges every finished output tile through
peer-mapped staging buffers (torch symmetric memory: NVLink P2P stores + flags) from inside the
tcgen05 kernel, so tiles travel while other tiles are still being multiplied and no second launch or
NCCL kernel sits on the per-token critical path.  ``mode="nccl"`` keeps the library baseline
(GEMM, then ``dist.all_reduce``, then the residual add) for comparison.

Every rank computes bit-identical hidden states (partials are added in rank order everywhere).
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..models.config import ModelConfig
from ..ops import gemm as G
from ..ops import native
from ..ops.packing import PackedWeight

MAX_TP_TOKENS = 16


class TpArgs(ctypes.Structure):
    """Mirror of ``b200_tp_args``."""

    _fields_ = [("world", ctypes.c_int32), ("rank", ctypes.c_int32), ("n_max", ctypes.c_int32),
                ("tiles_max", ctypes.c_int32), ("stage", ctypes.c_void_p * 8), ("flags", ctypes.c_void_p * 8),
                ("seq", ctypes.c_void_p), ("exit_counter", ctypes.c_void_p), ("mode", ctypes.c_int32)]


class NvlsArgs(ctypes.Structure):
    """Mirror of ``b200_nvls_args``."""

    _fields_ = [("mc_partial", ctypes.c_void_p), ("mc_out", ctypes.c_void_p), ("residual", ctypes.c_void_p),
                ("bias", ctypes.c_void_p), ("T", ctypes.c_int32), ("N", ctypes.c_int32), ("ldp", ctypes.c_int32),
                ("ldo", ctypes.c_int32), ("ldr", ctypes.c_int32), ("dtype", ctypes.c_int32), ("world", ctypes.c_int32),
                ("rank", ctypes.c_int32), ("flags_local", ctypes.c_void_p), ("flags_peer", ctypes.c_void_p * 8),
                ("seq", ctypes.c_void_p), ("cta_done", ctypes.c_void_p), ("pdl", ctypes.c_int32),
                ("reserved", ctypes.c_int32)]


class NvlsWorkspace:
    """Symmetric, multicast-mapped buffers for the NVLS all-reduce of prefill-size row-parallel outputs
    (``b200_tp_allreduce_nvls``): the partial-output buffer and the two hidden-state buffers of the span."""

    def __init__(self, group: dist.ProcessGroup, device, rows: int, hidden: int, dtype: torch.dtype):
        import torch.distributed._symmetric_memory as symm

        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.partial = symm.empty(rows, hidden, dtype=dtype, device=device)
        self.h = [symm.empty(rows, hidden, dtype=dtype, device=device) for _ in range(2)]
        self.flags = symm.empty(16, dtype=torch.int32, device=device)
        for t in (self.partial, *self.h, self.flags):
            t.zero_()
        hs = [symm.rendezvous(t, group.group_name) for t in (self.partial, *self.h, self.flags)]
        self.ok = all(int(getattr(h, "multicast_ptr", 0) or 0) != 0 for h in hs[:3])
        off = [int(getattr(h, "offset", 0) or 0) for h in hs]
        # multicast address of each buffer's STORAGE base (views are addressed relative to it, see mc_ptr)
        self.mc = {}
        if self.ok:
            for t, h, o in zip((self.partial, *self.h), hs[:3], off[:3]):
                sp = t.untyped_storage().data_ptr()
                self.mc[sp] = int(h.multicast_ptr) + o - (t.data_ptr() - sp)
        self.flag_ptrs = [int(hs[3].buffer_ptrs[r]) + off[3] for r in range(self.world)]
        self.local = torch.zeros(2, dtype=torch.int32, device=device)      # [launches, CTAs done]
        self._handles = hs
        torch.cuda.synchronize(device)
        dist.barrier(group)
        L = native.lib()
        L.b200_tp_allreduce_nvls.argtypes = [ctypes.POINTER(NvlsArgs), ctypes.c_void_p]
        L.b200_tp_allreduce_nvls.restype = ctypes.c_int

    def mc_ptr(self, t: torch.Tensor) -> int:
        """Multicast address of a view into one of the symmetric buffers (rows from the top)."""
        base = t.untyped_storage().data_ptr()
        return self.mc[base] + (t.data_ptr() - base)

    def allreduce(self, T: int, N: int, out: torch.Tensor, residual: Optional[torch.Tensor],
                  bias: Optional[torch.Tensor], pdl: bool = False) -> None:
        a = NvlsArgs()
        a.mc_partial = self.mc_ptr(self.partial)
        a.mc_out = self.mc_ptr(out)
        a.residual = residual.data_ptr() if residual is not None else None
        a.bias = bias.data_ptr() if bias is not None else None
        a.T, a.N = T, N
        a.ldp, a.ldo = self.partial.stride(0), out.stride(0)
        a.ldr = residual.stride(0) if residual is not None else 0
        a.dtype = G.dtype_code(out.dtype)
        a.world, a.rank = self.world, self.rank
        a.flags_local = self.flags.data_ptr()
        for r in range(self.world):
            a.flags_peer[r] = self.flag_ptrs[r]
        a.seq = self.local.data_ptr()
        a.cta_done = self.local.data_ptr() + 4
        a.pdl = 1 if pdl else 0
        native.check(native.lib().b200_tp_allreduce_nvls(ctypes.byref(a), native.current_stream_ptr()),
                     "b200_tp_allreduce_nvls")


class TPWorkspace:
    """Peer-mapped staging buffers + flags for the fused GEMM/all-reduce of one TP group."""

    def __init__(self, group: dist.ProcessGroup, device, n_max: int):
        import torch.distributed._symmetric_memory as symm

        self.group = group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.n_max = (n_max + 127) // 128 * 128
        self.tiles_max = self.n_max // 128
        stage_floats = 2 * 2 * self.world * MAX_TP_TOKENS * self.n_max   # (value, sequence number) pairs
        flag_ints = 2 * self.world * self.tiles_max
        self.stage = symm.empty(stage_floats, dtype=torch.float32, device=device)
        self.flags = symm.empty(flag_ints, dtype=torch.int32, device=device)
        self.stage.zero_()
        self.flags.zero_()
        self.h_stage = symm.rendezvous(self.stage, group.group_name)
        self.h_flags = symm.rendezvous(self.flags, group.group_name)
        self.local = torch.zeros(2, dtype=torch.int32, device=device)       # [seq, exit counter]
        torch.cuda.synchronize(device)
        dist.barrier(group)
        a = TpArgs()
        a.world, a.rank, a.n_max, a.tiles_max = self.world, self.rank, self.n_max, self.tiles_max
        # buffer_ptrs address the allocation block; the tensor may sit at an offset inside it
        off_s = int(getattr(self.h_stage, "offset", 0) or 0)
        off_f = int(getattr(self.h_flags, "offset", 0) or 0)
        for r in range(self.world):
            a.stage[r] = int(self.h_stage.buffer_ptrs[r]) + off_s
            a.flags[r] = int(self.h_flags.buffer_ptrs[r]) + off_f
        if a.stage[self.rank] != self.stage.data_ptr() or a.flags[self.rank] != self.flags.data_ptr():
            raise RuntimeError("symmetric-memory handle does not address the local tensor "
                               f"(stage {a.stage[self.rank]:#x} vs {self.stage.data_ptr():#x})")
        a.seq = self.local.data_ptr()
        a.exit_counter = self.local.data_ptr() + 4
        a.mode = 0
        self.args = a
        d = TpArgs()
        ctypes.memmove(ctypes.byref(d), ctypes.byref(a), ctypes.sizeof(TpArgs))
        d.mode = 1
        self.args_defer = d          # producer pushes only; the next GEMM gathers (linear_gather)
        L = native.lib()
        L.b200_gemm_gather.argtypes = [ctypes.POINTER(native.GemmArgs), ctypes.POINTER(TpArgs), ctypes.c_void_p,
                                       ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]
        L.b200_gemm_gather.restype = ctypes.c_int
        L.b200_gemm_allreduce.argtypes = [ctypes.POINTER(native.GemmArgs), ctypes.POINTER(TpArgs), ctypes.c_void_p]
        L.b200_gemm_allreduce.restype = ctypes.c_int


def linear_allreduce(x: torch.Tensor, w: PackedWeight, tp: TPWorkspace, *, out: torch.Tensor,
                     residual: Optional[torch.Tensor] = None, workspace: Optional[G.GemmWorkspace] = None,
                     pdl: bool = False, add_bias: bool = True, defer: bool = False) -> torch.Tensor:
    """out = sum_over_ranks(x_shard @ W_shard^T) [+ bias] [+ residual], identical on every rank.

    ``defer=True``: only push this rank's partial tiles; ``out`` is NOT written — the next GEMM must be
    issued with :func:`linear_gather`, which waits for the partials, adds them (+ bias, + residual) into
    its input buffer and multiplies."""
    a, out = G.fill_args(x, w, out=out, residual=None if defer else residual, workspace=workspace, pdl=pdl)
    if not add_bias or defer:
        a.bias = None
    tpa = tp.args_defer if defer else tp.args
    rc = native.lib().b200_gemm_allreduce(ctypes.byref(a), ctypes.byref(tpa), native.current_stream_ptr())
    native.check(rc, "b200_gemm_allreduce")
    return out


def linear_gather(x_buf: torch.Tensor, w: PackedWeight, tp: TPWorkspace, *, residual: torch.Tensor,
                  bias: Optional[torch.Tensor] = None, **linear_kw) -> torch.Tensor:
    """GEMM whose input ``x_buf`` is first rebuilt as residual + bias + sum of the partial rows a deferred
    :func:`linear_allreduce` pushed (all-gather -> GEMM in one kernel).  ``x_buf`` holds the reduced hidden
    state afterwards.  Remaining keywords as :func:`petals_b200.ops.gemm.linear`."""
    ws = linear_kw.pop("workspace", None)
    pdl = linear_kw.pop("pdl", False)
    out = linear_kw.pop("out", None)
    a, out = G.fill_args(x_buf, w, out=out, workspace=ws, pdl=pdl, **linear_kw)
    rc = native.lib().b200_gemm_gather(ctypes.byref(a), ctypes.byref(tp.args), residual.data_ptr(), residual.stride(0),
                                       bias.data_ptr() if bias is not None else None, native.current_stream_ptr())
    native.check(rc, "b200_gemm_gather")
    return out


# ----------------------------------------------------------------------------- weight sharding
def shard_layer(cfg: ModelConfig, L: Dict[str, torch.Tensor], rank: int, world: int) -> Dict[str, torch.Tensor]:
    """Canonical layer weights -> this rank's Megatron shard (same key names, smaller tensors)."""
    if cfg.n_heads % world or cfg.n_kv_heads % world or cfg.intermediate % world:
        raise ValueError(f"cannot split {cfg.n_heads}/{cfg.n_kv_heads} heads, {cfg.intermediate} MLP columns {world} ways")
    d = cfg.head_dim
    qh, kh = cfg.n_heads // world, cfg.n_kv_heads // world
    q0, k0, v0 = 0, cfg.q_size, cfg.q_size + cfg.kv_size
    rows = torch.cat([torch.arange(q0 + rank * qh * d, q0 + (rank + 1) * qh * d),
                      torch.arange(k0 + rank * kh * d, k0 + (rank + 1) * kh * d),
                      torch.arange(v0 + rank * kh * d, v0 + (rank + 1) * kh * d)]).to(L["qkv_w"].device)
    out = {k: v for k, v in L.items() if k.startswith("ln")}
    out["qkv_w"] = L["qkv_w"][rows].contiguous()
    if "qkv_b" in L:
        out["qkv_b"] = L["qkv_b"][rows].contiguous()
    out["o_w"] = L["o_w"][:, rank * qh * d:(rank + 1) * qh * d].contiguous()
    if "o_b" in L:
        out["o_b"] = L["o_b"]                         # added once, after the reduction
    ic = cfg.intermediate // world
    sl = slice(rank * ic, (rank + 1) * ic)
    if cfg.is_llama:
        out["gate_w"], out["up_w"] = L["gate_w"][sl].contiguous(), L["up_w"][sl].contiguous()
    else:
        out["fc_w"], out["fc_b"] = L["fc_w"][sl].contiguous(), L["fc_b"][sl].contiguous()
    out["down_w"] = L["down_w"][:, sl].contiguous()
    if "down_b" in L:
        out["down_b"] = L["down_b"]
    return out


def shard_config(cfg: ModelConfig, world: int) -> ModelConfig:
    """Per-rank view: fewer heads and MLP columns, same hidden size."""
    from dataclasses import replace

    return replace(cfg, n_heads=cfg.n_heads // world, n_kv_heads=cfg.n_kv_heads // world,
                   intermediate=cfg.intermediate // world)


def make_tp_span(cfg: ModelConfig, weights: Dict[str, object], device, dtype, group: dist.ProcessGroup, *,
                 mode: str = "fused", **span_kw):
    """NativeSpan over this rank's shard; o_proj / down_proj reduce across ``group``."""
    from ..runtime.native_span import NativeSpan

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    w = dict(weights)
    w["layers"] = [shard_layer(cfg, L, rank, world) for L in weights["layers"]]  # type: ignore[index]
    local_cfg = shard_config(cfg, world)
    if mode == "nccl":
        span_kw["use_graphs"] = False   # NCCL collectives are issued eagerly between the kernels (baseline path)
    span = NativeSpan(local_cfg, w, device, dtype, **span_kw)
    span.full_hidden = cfg.hidden
    span.tp_mode = mode
    span.tp_group = group
    span.tp = TPWorkspace(group, device, cfg.hidden) if mode == "fused" else None
    span.nvls = None
    if mode == "fused" and os.environ.get("B200_TP_NVLS", "1") != "0":
        # prefill-size row-parallel outputs are reduced through the switch: the span's hidden-state buffers move into
        # symmetric, multicast-mapped memory so that the reducing rank can write every rank's copy with multimem.st
        try:
            nv = NvlsWorkspace(group, device, span.h_a.shape[0], cfg.hidden, dtype)
            if nv.ok:
                span.h_a, span.h_b = nv.h
                span.nvls = nv
        except Exception as e:  # noqa: BLE001  (no multicast support: the NCCL path stays)
            import logging

            logging.getLogger(__name__).warning("NVLS workspace unavailable (%s): prefill all-reduce stays on NCCL", e)
    return span

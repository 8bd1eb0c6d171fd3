"""B200 execution of a layer span: packed weights + hand-written sm_100a kernels.

One ``NativeSpan`` owns the decoder blocks ``[start, end)`` of a model on one GPU
(optionally the embedding in front and the final norm + LM head + arg-max
behind) and runs them with the kernels of ``csrc/``:

decode (<= 32 tokens, one per session), per block — 5 launches:
    qkv   = GEMM(norm-prologue(h), Wqkv) [+bias]          gemm.cu (tcgen05, stream-K)
    attn  = RoPE + KV append + split-KV attention         attention.cu
    h2    = GEMM(attn, Wo) [+bias] + h                    residual epilogue
    act   = GEMM(norm-prologue(h2), Wgate|Wup) -> SwiGLU  (GPT-2: Wfc + bias -> GELU-tanh)
    h     = GEMM(act, Wdown) [+bias] + h2
  the whole step (all blocks, plus embedding / head / arg-max on the edge stages)
  is captured once per batch size in a CUDA graph and replayed per token
  (the reference graphs only RoPE and the two RMSNorms: petals/llama/cuda_graphs.py).

prefill (a chunk of one session): standalone norm -> GEMM -> rope_append ->
causal attention against the session cache -> GEMMs, chunked to ``max_chunk``.

Replaces src/llama_partition.py:99-137,222-297,391-474 (Stage forward loops).
"""
from __future__ import annotations

import contextlib
import gc
import math
import os
from typing import Dict, List, Optional, Sequence

import torch

from ..models.config import ModelConfig
from ..ops import attention as A
from ..ops import elementwise as E
from ..ops import gemm as G
from ..ops import native
from ..ops import packing
from ..utils.metrics import nvtx_range
from .kv_cache import SessionKV

MAX_DECODE_BATCH = 32


@contextlib.contextmanager
def _capture_guard():
    """No cyclic garbage collection while a stream is capturing: a collection that happens to run inside the capture
    may finalise dead objects of this thread that own CUDA resources (graphs, streams, raw allocations) - their
    destructors make calls that are illegal during capture and invalidate it ("operation failed due to a previous error
    during capture"; seen when a test suite had left such garbage behind).  Collect first, then hold the collector."""
    gc.collect()
    was = gc.isenabled()
    gc.disable()
    try:
        yield
    finally:
        if was:
            gc.enable()


RSTD_KERNEL_MIN_TOKENS = int(os.environ.get("B200_RSTD_KERNEL_MIN_TOKENS", "8"))
# Decode steps of up to this many tokens run as a flag chain (b200_ll_chain): the kernels of the step hand their
# activations over as self-validating words instead of waiting for each other's grids.  Bit-identical to the
# grid-ordered step, but measured SLOWER on Llama-2-7B batch-1 decode (1xB200): 352 tok/s against 361 tok/s
# grid-ordered (a persistent single-launch variant of the chain reached 259 tok/s and was removed again, see DESIGN
# 2.2) - the per-launch timeline (profiles/decode_timeline_r2_*.txt) shows that what a launch boundary costs is not the
# grid-completion signal but the start-up of the next kernel's CTAs, which cannot be resident before the previous
# kernel's CTAs leave.  So the chain is OFF by default (0); set B200_LL_CHAIN_MAX_TOKENS=4 to experiment.
CHAIN_MAX_TOKENS = int(os.environ.get("B200_LL_CHAIN_MAX_TOKENS", "0"))


def _l2_prefetch_default():
    """B200_L2_PREFETCH_MB="a,b,c,d" (MiB, see _decode_layers); "0" switches the L2 warm-up off.

    Measured on Llama-2-7B decode (1xB200, after the ticket-free fix-up): off 345.9 tok/s, "24,16,16,16" 350.4 /
    351.6, "16,8,8,8" 348.8, "32,24,32,32" 347.2, "48,0,0,0" 344.3, "64,48,64,64" slower than off.  Small ranges
    fill the HBM idle time around the kernel boundaries; large ones compete with the running kernel's own
    stream.  (Before the fix-up change the same hint was neutral.)"""
    v = os.environ.get("B200_L2_PREFETCH_MB", "24,16,16,16")
    if v.strip() in ("0", "", "off"):
        return None
    mb = [int(t) for t in v.split(",")]
    assert len(mb) == 4, "B200_L2_PREFETCH_MB needs four comma-separated MiB values"
    return mb


def _pdl_default() -> bool:
    return os.environ.get("B200_PDL", "1") != "0"


class StepProgram:
    """What a captured decode step does AROUND the blocks (the pipeline engine's in-graph hops).

    ``pre(B, dry)`` runs before the embedding / blocks (e.g. receive the hidden state into ``h_a``, or receive the
    previous token), ``post(B, h, dry)`` after the blocks and the head (e.g. send ``h`` to the next stage).  Both
    are called once with ``dry=True`` outside the capture (warm-up: they must not touch a link then) and once
    inside it.  ``layers=False`` makes a step that only runs ``pre``; ``head=False`` skips the span's own head
    after the blocks (a first stage that only holds a shard of the head for the NEXT token's arg-max)."""

    def __init__(self, key, from_ids: bool, pre=None, post=None, layers: bool = True, head: bool = True):
        self.key, self.from_ids, self.pre, self.post, self.layers, self.head = key, from_ids, pre, post, layers, head


class NativeSpan:
    def __init__(self, cfg: ModelConfig, weights: Dict[str, object], device, dtype: torch.dtype, *,
                 kv_capacity: int = 2048, max_batch: int = 8, max_chunk: int = 512, attn_splits: int = 0,
                 use_graphs: bool = True, pdl: Optional[bool] = None, free_source_weights: bool = True,
                 weight_format: str = "native",
                 head_rows: Optional[tuple] = None, streamed_layers: int = 0):
        assert dtype in (torch.float16, torch.bfloat16), "native path runs fp16/bf16"
        self.cfg, self.device, self.dtype = cfg, torch.device(device), dtype
        self.kv_capacity, self.max_batch, self.max_chunk = kv_capacity, min(max_batch, MAX_DECODE_BATCH), max_chunk
        self.use_graphs = use_graphs
        if weight_format not in ("native", "fp8", "mxfp8"):
            raise ValueError("weight_format must be 'native' (fp16/bf16), 'fp8' (e4m3, one scale per row) or "
                             "'mxfp8' (e4m3 with OCP-MX block scales)")
        if weight_format != "native" and not cfg.is_llama:
            raise ValueError("fp8 weights are implemented for RMSNorm (Llama-family) models")
        self.weight_format = weight_format
        fp8 = weight_format != "native"
        if fp8:
            self.max_chunk = max_chunk = min(max_chunk, G.MAX_FUSED_NORM_TOKENS)   # fp8 GEMM quantises <= 32 tokens per call
        pack_w = {"native": packing.pack_weight, "fp8": packing.pack_weight_fp8, "mxfp8": packing.pack_weight_mx}[weight_format]
        pack_d = {"native": packing.pack_dual, "fp8": packing.pack_dual_fp8, "mxfp8": packing.pack_dual_mx}[weight_format]
        self.pdl = _pdl_default() if pdl is None else pdl
        self.l2_prefetch_mb = _l2_prefetch_default()
        self._prefill_plans = {}
        self._planned_prefill = os.environ.get("B200_PREFILL_PLAN", "1") != "0"
        self.has_embed = bool(weights.get("has_embed", False))
        self.has_head = bool(weights.get("has_head", "lm_head" in weights))
        c = cfg
        with torch.cuda.device(self.device):
            self.layers = []
            for L in weights["layers"]:  # type: ignore[union-attr]
                P = {"ln1_w": L["ln1_w"], "ln1_b": L.get("ln1_b"), "ln2_w": L["ln2_w"], "ln2_b": L.get("ln2_b"),
                     "qkv": pack_w(L["qkv_w"], L.get("qkv_b")),
                     "o": pack_w(L["o_w"], L.get("o_b")),
                     "down": pack_w(L["down_w"], L.get("down_b"))}
                if c.is_llama:
                    P["mlp_in"] = pack_d(L["gate_w"], L["up_w"])
                else:
                    P["mlp_in"] = pack_w(L["fc_w"], L.get("fc_b"))
                self.layers.append(P)
                if free_source_weights:
                    for k in ("qkv_w", "o_w", "down_w", "gate_w", "up_w", "fc_w"):
                        L.pop(k, None)
            self.embed_w = weights.get("embed") if self.has_embed else None
            self.pos_embed_w = weights.get("pos_embed") if self.has_embed else None
            # head_rows = (lo, hi): this span holds only rows [lo, hi) of the LM head (a vocabulary shard; lo must
            # sit on a 128-row tile boundary); its arg-max reports indices with lo added.
            self.head_lo, self.head_hi = (0, c.vocab) if head_rows is None else (int(head_rows[0]), int(head_rows[1]))
            if self.has_head:
                self.final_norm_w = weights["final_norm_w"]
                self.final_norm_b = weights.get("final_norm_b")
                assert self.head_lo % 128 == 0 and 0 <= self.head_lo < self.head_hi <= c.vocab
                hw = weights["lm_head"]
                if head_rows is not None:
                    hw = hw[self.head_lo:self.head_hi]  # type: ignore[index]
                self.lm_head = pack_w(hw)  # type: ignore[arg-type]
            self.inv_freq = (A.rope_inv_freq(c.head_dim, c.rope_theta, c.rope_scaling, self.device)
                             if c.is_llama else None)
            self.scale = 1.0 / math.sqrt(c.head_dim)
            self.norm_mode = G.NORM_RMS if c.is_llama else G.NORM_LAYER
            self.mlp_epi = G.EPI_SWIGLU if c.is_llama else G.EPI_GELU_TANH
            # streamed_layers: blocks that are NOT resident (runtime/offload.py streams them through two device slots)
            self.streamed_layers = int(streamed_layers)
            nl = len(self.layers) + self.streamed_layers
            Tm = max(self.max_chunk, self.max_batch)
            mk = lambda *s, dt=dtype: torch.empty(*s, dtype=dt, device=self.device)  # noqa: E731
            self.h_a, self.h_b, self.xn = mk(Tm, c.hidden), mk(Tm, c.hidden), mk(Tm, c.hidden)
            self.qkv = mk(Tm, c.qkv_size)
            self.attn = mk(Tm, c.q_size)
            # width padded to the packed K of down_proj (64-wide k-blocks); the pad columns stay zero, which is what
            # lets an intermediate size that is not a multiple of 64 (Llama-2-7B split 8 ways: 1376) through
            self.act = torch.zeros(Tm, (c.intermediate + 63) // 64 * 64, dtype=dtype, device=self.device)
            self.ids_in = torch.zeros(Tm, dtype=torch.int64, device=self.device)
            self.rstd = torch.zeros(max(Tm, 32), dtype=torch.float32, device=self.device)
            self.ws = G.GemmWorkspace.create(self.device, Tm)
            self.splits = attn_splits or int(os.environ.get("B200_ATTN_SPLITS", "0")) or max(1, min(16, kv_capacity // 64))
            self.attn_ws = A.DecodeAttnWorkspace(self.device, self.max_batch, c.n_heads, c.n_kv_heads, c.head_dim,
                                                 self.splits)
            # What a captured step reads about "its" sessions lives in ONE device buffer, so that pointing the graph at
            # the sessions of the next step is one host-to-device copy: positions of the step's tokens, the step number
            # the flag chain tags its words with, the sampler rows of the sessions (device-side sampling, see
            # enable_sampling) and, behind them, the table of cache pointers [layer][token][K, V].
            self._make_stage_buffers(nl)
            self.sampler = None
            self._seq_host = 0
            self.chain_tokens = 0
            if (CHAIN_MAX_TOKENS > 0 and not fp8 and self.layers
                    and all(G.chain_capable(P[k]) for P in self.layers for k in ("qkv", "o", "mlp_in", "down"))):
                Tc = self.chain_tokens = min(CHAIN_MAX_TOKENS, self.max_batch, RSTD_KERNEL_MIN_TOKENS - 1)
                ll = lambda width: torch.zeros(Tc, (width // 2 + 3) // 4 * 4, dtype=torch.int64, device=self.device)  # noqa: E731
                self.h_a_ll, self.h_b_ll = ll(c.hidden), ll(c.hidden)
                self.qkv_ll, self.attn_ll, self.act_ll = ll(c.qkv_size), ll(c.q_size), ll(self.act.shape[1])
            self._pos_ring_idx = 0
            self._ptr_tables: Dict[tuple, torch.Tensor] = {}
            if self.has_head:
                self.logits = mk(self.max_batch, self.head_hi - self.head_lo, dt=torch.float32)
                self.ids_out = torch.zeros(self.max_batch, dtype=torch.int64, device=self.device)
                self.argmax_ws = E.ArgmaxWorkspace(self.device, self.max_batch)
        self._graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}
        self._ptr_key = None
        self.kernel_launches = 0   # launches issued or replayed (bench bookkeeping)
        self._launches_per_graph: Dict[tuple, int] = {}

    # ------------------------------------------------------------------ pieces
    @property
    def n_layers(self) -> int:
        return len(self.layers) + self.streamed_layers

    def _make_stage_buffers(self, nl: int) -> None:
        """The per-step staging buffer (positions | step number | sampler rows | cache pointer table) and its pinned
        host ring, for a span of ``nl`` blocks."""
        mb = self.max_batch
        self._n_pos = 2 * mb + 1
        n_head = (self._n_pos + 1) // 2 * 2                     # the int64 table starts 8-byte aligned
        n_tab = max(nl, 1) * mb * 2
        self._stage_dev = torch.zeros(n_head + 2 * n_tab, dtype=torch.int32, device=self.device)
        self._n_stage_head = n_head
        self.pos = self._stage_dev[:self._n_pos]
        self.kv_ptrs = self._stage_dev[n_head:].view(torch.int64).view(max(nl, 1), mb, 2)
        self.step_seq = self.pos[mb:mb + 1]
        self.samp_idx = self.pos[mb + 1:]
        self._pos_ring = [(torch.zeros(n_head + 2 * n_tab, dtype=torch.int32).pin_memory(), torch.cuda.Event())
                          for _ in range(8)]
        self._pos_ring_idx = 0

    def subspan(self, a: int, b: int) -> "NativeSpan":
        """A view that runs only blocks [a, b) of this span: same packed weights and work buffers, its own graphs and
        plans.  Lets a client enter (or leave) a server's span at any block - upstream addresses blocks one by one
        (chained module uids, petals/server/handler.py) - without a second copy of the weights.  Calls on a span and
        its views must not overlap in time (the stage's task pool runs them one at a time)."""
        import copy

        if not (0 <= a < b <= len(self.layers)):
            raise ValueError(f"blocks [{a},{b}) are not inside this span of {len(self.layers)} blocks")
        if self.streamed_layers or self.tp_mode is not None:
            raise NotImplementedError("sub-spans of offloaded or tensor-parallel spans")
        v = copy.copy(self)
        v.layers = self.layers[a:b]
        v.has_embed = self.has_embed and a == 0
        v.has_head = self.has_head and b == len(self.layers)
        v._graphs, v._launches_per_graph, v._prefill_plans, v._ptr_tables = {}, {}, {}, {}
        v._ptr_key = None
        v.__dict__.pop("_graph_out", None)
        with torch.cuda.device(self.device):
            v._make_stage_buffers(b - a)       # its graphs read its own positions and cache pointers
        return v

    def _iter_layers(self):
        """(index, packed weights) of every block in order; a span that streams blocks from the host overrides this."""
        return enumerate(self.layers)

    def weight_bytes(self) -> int:
        n = sum(P[k].nbytes for P in self.layers for k in ("qkv", "o", "mlp_in", "down"))
        if self.has_head:
            n += self.lm_head.nbytes
        return n

    def _count(self, n: int = 1) -> None:
        self.kernel_launches += n

    tp = None          # TPWorkspace when this span is one shard of a tensor-parallel group
    tp_mode = None     # "fused" | "nccl"
    tp_group = None

    def _row_parallel(self, x: torch.Tensor, w, out: torch.Tensor, residual: torch.Tensor, prefetch=None) -> None:
        """o_proj / down_proj: plain on one GPU, reduced across the group when tensor-parallel."""
        if self.tp_mode is None:
            G.linear(x, w, out=out, residual=residual, workspace=self.ws, pdl=self.pdl, prefetch=prefetch)
        elif self.tp_mode == "fused" and x.shape[0] <= 16:
            from ..parallel.tensor_parallel import linear_allreduce

            linear_allreduce(x, w, self.tp, out=out, residual=residual, workspace=self.ws, pdl=self.pdl)
        elif self.tp_mode == "fused" and getattr(self, "nvls", None) is not None and x.shape[0] <= self.nvls.partial.shape[0] \
                and out.untyped_storage().data_ptr() in self.nvls.mc:
            # prefill sizes: this rank's partial goes into the symmetric buffer, then ONE kernel reduces the ranks'
            # partials through the switch (multimem.ld_reduce), adds bias / residual and writes every rank's output
            # (multimem.st) - see csrc/tp_nvls.cu
            T = x.shape[0]
            bias, w.bias = w.bias, None
            try:
                G.linear(x, w, out=self.nvls.partial[:T], workspace=self.ws, pdl=self.pdl)
            finally:
                w.bias = bias
            self.nvls.allreduce(T, w.n, out, residual, bias, pdl=self.pdl)
            self._count()
        else:
            # library baseline: GEMM without bias/residual, NCCL all-reduce, then the epilogue in torch
            import torch.distributed as dist

            bias, w.bias = w.bias, None
            try:
                G.linear(x, w, out=out, workspace=self.ws, pdl=False)
            finally:
                w.bias = bias
            dist.all_reduce(out, group=self.tp_group)
            if bias is not None:
                out.add_(bias)
            out.add_(residual)

    def _decode_layers(self, T: int, h: torch.Tensor, with_head: bool = False) -> torch.Tensor:
        """h: [T, H] view of a static buffer; returns the buffer view holding the result.

        ``with_head``: the caller will run the LM head on the result next."""
        self._head_done = False
        c, pdl = self.cfg, self.pdl
        other = self.h_b if h.data_ptr() == self.h_a.data_ptr() else self.h_a
        deferred = self.tp_mode == "fused" and T <= 16 and getattr(self, "tp_defer", False)   # measured slower than the in-producer exchange
        if deferred:
            return self._decode_layers_tp_deferred(T, h, other)
        # L2 warm-up plan (decode sizes only): every GEMM, once its own weight stream is issued, pulls the head
        # of the following linears' weights into the 126 MB L2, so the HBM does not idle through the fix-up,
        # the attention kernel and the launches.  MiB: [mlp_in from qkv, more mlp_in from o, down from mlp_in,
        # next qkv from down]; o_proj is always taken whole by qkv.
        mb = self.l2_prefetch_mb if T <= 32 else None
        n = len(self.layers)
        M = 1 << 20
        chain = self._chain_on(T) and not self.streamed_layers
        if self.streamed_layers:
            mb = None                       # the next block's weights are not where a prefetch hint could point
        if chain:
            return self._decode_layers_chain(T, h, other, mb)
        for li, P in self._iter_layers():
            qkv, attn, act, h2 = self.qkv[:T], self.attn[:T], self.act[:T], other[:T]
            pf = [None] * 4
            if mb:
                nxt = self.layers[li + 1]["qkv"] if li + 1 < n else (self.lm_head if self.has_head else None)
                pf[0] = [(P["o"], 0, P["o"].nbytes), (P["mlp_in"], 0, mb[0] * M)]
                pf[1] = [(P["mlp_in"], mb[0] * M, mb[1] * M)]
                pf[2] = [(P["down"], 0, mb[2] * M)]
                pf[3] = [(nxt, 0, mb[3] * M)] if nxt is not None else None
            # Many sessions per step: the RMSNorm statistics are computed once by a one-block-per-token kernel
            # instead of once per CTA of the GEMM (~1 us per token pair there); same bits either way.
            rs = None
            if c.is_llama and T >= RSTD_KERNEL_MIN_TOKENS:
                rs = E.rms_rstd(h, c.norm_eps, self.rstd, pdl=pdl)
                self._count()
            G.linear(h, P["qkv"], out=qkv, norm_mode=self.norm_mode, norm_w=P["ln1_w"], norm_b=P["ln1_b"],
                     norm_eps=c.norm_eps, workspace=self.ws, pdl=pdl, prefetch=pf[0], rstd=rs)
            A.attn_decode(qkv, attn, self.kv_ptrs[li], self.pos, self.inv_freq, self.attn_ws, c.n_heads,
                          c.n_kv_heads, c.head_dim, self.kv_capacity, self.scale, pdl=pdl)
            self._row_parallel(attn, P["o"], h2, h, prefetch=pf[1])
            if rs is not None:
                rs = E.rms_rstd(h2, c.norm_eps, self.rstd, pdl=pdl)
                self._count()
            G.linear(h2, P["mlp_in"], out=act, epilogue=self.mlp_epi, norm_mode=self.norm_mode, norm_w=P["ln2_w"],
                     norm_b=P["ln2_b"], norm_eps=c.norm_eps, workspace=self.ws, pdl=pdl, prefetch=pf[2], rstd=rs)
            self._row_parallel(act, P["down"], h, h2, prefetch=pf[3])
            self._count(5)
        return h

    def _chain_on(self, T: int) -> bool:
        return 0 < T <= self.chain_tokens and self.tp_mode is None and self.pdl

    def _link(self, x_ll, x_op, out_ll, out_op, res_ll=None, res_op=0):
        """One link of the flag chain (see b200_ll_chain): who produced the input words, where the output words go."""
        a = native.LLChainArgs()
        a.seq = self.step_seq.data_ptr()
        a.x_ll, a.ldx_ll, a.x_op = x_ll.data_ptr(), x_ll.stride(0), x_op
        if res_ll is not None:
            a.res_ll, a.ldr_ll, a.res_op = res_ll.data_ptr(), res_ll.stride(0), res_op
        if out_ll is not None:
            a.out_ll, a.ldo_ll = out_ll.data_ptr(), out_ll.stride(0)
        a.out_op = out_op
        a.half = self._chain_half
        return a

    def _decode_layers_chain(self, T: int, h: torch.Tensor, other: torch.Tensor, mb) -> torch.Tensor:
        """The blocks of a decode step as one flag chain: 5 launches per block as before, but no launch waits for
        the grid before it - every kernel polls the tagged words it needs (h -> qkv -> attn -> h2 -> act -> h)."""
        c, pdl = self.cfg, True
        n = len(self.layers)
        M = 1 << 20
        h_ll = self.h_a_ll if h.data_ptr() == self.h_a.data_ptr() else self.h_b_ll
        o_ll = self.h_b_ll if h_ll is self.h_a_ll else self.h_a_ll
        qkv, attn, act, h2 = self.qkv[:T], self.attn[:T], self.act[:T], other[:T]
        op = 1
        E.ll_publish(h, h_ll, self.step_seq, op, pdl=pdl)     # head of the chain: waits for whatever produced h
        self._count()
        h_op = op
        self._chain_half = 0

        def gemm(x, w, x_ll, x_op, out, out_ll, res=None, res_ll=None, res_op=0, prefetch=None, **kw):
            nonlocal op
            op += 1
            G.linear(x, w, out=out, residual=res, workspace=self.ws, pdl=pdl, prefetch=prefetch,
                     chain=self._link(x_ll, x_op, out_ll, op, res_ll, res_op), **kw)
            self._chain_half ^= 1
            return op

        for li, P in enumerate(self.layers):
            pf = [None] * 4
            if mb:
                nxt = self.layers[li + 1]["qkv"] if li + 1 < n else (self.lm_head if self.has_head else None)
                pf[0] = [(P["o"], 0, P["o"].nbytes), (P["mlp_in"], 0, mb[0] * M)]
                pf[1] = [(P["mlp_in"], mb[0] * M, mb[1] * M)]
                pf[2] = [(P["down"], 0, mb[2] * M)]
                pf[3] = [(nxt, 0, mb[3] * M)] if nxt is not None else None
            q_op = gemm(h, P["qkv"], h_ll, h_op, qkv, self.qkv_ll, prefetch=pf[0], norm_mode=self.norm_mode,
                        norm_w=P["ln1_w"], norm_b=P["ln1_b"], norm_eps=c.norm_eps)
            op += 1
            a_op = op
            A.attn_decode(qkv, attn, self.kv_ptrs[li], self.pos, self.inv_freq, self.attn_ws, c.n_heads, c.n_kv_heads,
                          c.head_dim, self.kv_capacity, self.scale, pdl=pdl,
                          chain=self._link(self.qkv_ll, q_op, self.attn_ll, a_op))
            o_op = gemm(attn, P["o"], self.attn_ll, a_op, h2, o_ll, res=h, res_ll=h_ll, res_op=h_op, prefetch=pf[1])
            m_op = gemm(h2, P["mlp_in"], o_ll, o_op, act, self.act_ll, prefetch=pf[2], epilogue=self.mlp_epi,
                        norm_mode=self.norm_mode, norm_w=P["ln2_w"], norm_b=P["ln2_b"], norm_eps=c.norm_eps)
            h_op = gemm(act, P["down"], self.act_ll, m_op, h, h_ll, res=h2, res_ll=o_ll, res_op=o_op, prefetch=pf[3])
            self._count(5)
        self._chain_tail = (h_ll, h_op)      # the head GEMM (if any) continues the chain
        return h

    def _decode_layers_tp_deferred(self, T: int, h: torch.Tensor, other: torch.Tensor) -> torch.Tensor:
        """Tensor-parallel decode with the exchange split across kernels: each row-parallel linear only
        PUSHES its partial tiles to the peers; the following column-parallel linear waits for them and adds
        them up while staging its input (all-gather -> GEMM), so the NVLink latency hides behind that
        kernel's launch.  The last down_proj reduces in place (its consumer may be another stage)."""
        from ..parallel.tensor_parallel import linear_allreduce, linear_gather

        c, pdl, tp = self.cfg, self.pdl, self.tp
        pending = None      # (residual buffer, bias) of a pushed-but-not-yet-reduced linear
        n = len(self.layers)
        for li, P in enumerate(self.layers):
            qkv, attn, act, h2 = self.qkv[:T], self.attn[:T], self.act[:T], other[:T]
            kw = dict(out=qkv, norm_mode=self.norm_mode, norm_w=P["ln1_w"], norm_b=P["ln1_b"], norm_eps=c.norm_eps,
                      workspace=self.ws, pdl=pdl)
            if pending is None:
                G.linear(h, P["qkv"], **kw)
            else:
                linear_gather(h, P["qkv"], tp, residual=pending[0], bias=pending[1], **kw)
            A.attn_decode(qkv, attn, self.kv_ptrs[li], self.pos, self.inv_freq, self.attn_ws, c.n_heads,
                          c.n_kv_heads, c.head_dim, self.kv_capacity, self.scale, pdl=pdl)
            linear_allreduce(attn, P["o"], tp, out=h2, workspace=self.ws, pdl=pdl, defer=True)
            linear_gather(h2, P["mlp_in"], tp, residual=h, bias=P["o"].bias, out=act, epilogue=self.mlp_epi,
                          norm_mode=self.norm_mode, norm_w=P["ln2_w"], norm_b=P["ln2_b"], norm_eps=c.norm_eps,
                          workspace=self.ws, pdl=pdl)
            if li + 1 < n:
                linear_allreduce(act, P["down"], tp, out=h, workspace=self.ws, pdl=pdl, defer=True)
                pending = (h2, P["down"].bias)
            else:
                linear_allreduce(act, P["down"], tp, out=h, residual=h2, workspace=self.ws, pdl=pdl)
                pending = None
            self._count(5)
        return h

    def _head(self, h_rows: torch.Tensor, greedy: bool, out_ids: Optional[torch.Tensor] = None,
              init_pair: Optional[torch.Tensor] = None, out_pair: Optional[torch.Tensor] = None,
              gemm_done: bool = False, row_index: Optional[torch.Tensor] = None) -> torch.Tensor:
        """h_rows: [B<=32, H] -> logits [B, V (or this span's vocabulary shard)] fp32 (and ids_out when greedy).

        ``init_pair`` / ``out_pair``: arg-max candidates exchanged between vocabulary shards (see ops.elementwise.argmax)."""
        c = self.cfg
        B = h_rows.shape[0]
        logits = self.logits[:B]
        link = None
        tail = getattr(self, "_chain_tail", None)
        if tail is not None:       # directly behind the blocks of a flag-chained step: poll their output words
            self._chain_tail = None
            link = self._link(tail[0], tail[1], None, tail[1] + 1)
        if not gemm_done:          # (a persistent chain launch has multiplied the head already)
            G.linear(h_rows, self.lm_head, out=logits, norm_mode=self.norm_mode, norm_w=self.final_norm_w,
                     norm_b=self.final_norm_b, norm_eps=c.norm_eps, out_fp32=True, workspace=self.ws, pdl=self.pdl,
                     chain=link)
            self._count()
        if greedy == "sample":
            # device-side sampling: parameters, recent tokens and the Philox stream of every session live on the GPU
            assert self.sampler is not None and self.head_lo == 0 and self.head_hi == c.vocab, \
                "sampling needs enable_sampling() and the whole vocabulary on this span"
            idx = self.samp_idx[:B] if row_index is None else row_index
            self.sampler.sample(logits, idx, self.ids_out[:B] if out_ids is None else out_ids, pdl=self.pdl)
            self._count()
        elif greedy:
            ids = self.ids_out[:B] if out_ids is None else out_ids
            E.argmax(logits, ids, self.argmax_ws, pdl=self.pdl, idx_offset=self.head_lo, init_pair=init_pair,
                     out_pair=out_pair)
            self._count()
        return logits

    def enable_sampling(self, n_rows: int):
        """Device-resident sampling state for up to ``n_rows`` sessions (``decode(..., greedy="sample")``): open a
        row per session with ``span.sampler.open(params, seed)`` and store its index in ``kv.sampler_row``."""
        from ..ops.sampling import SamplerPool

        if self.sampler is None:
            self.sampler = SamplerPool(self.device, n_rows, self.cfg.vocab, self.max_batch)
        return self.sampler

    def _embed(self, T: int, pos0: int, use_pos_buffer: bool) -> torch.Tensor:
        h = self.h_a[:T]
        E.embed(self.ids_in[:T], self.embed_w, h, wpe=self.pos_embed_w,
                pos=self.pos[:T] if use_pos_buffer else None, pos0=pos0, pdl=self.pdl)
        self._count()
        return h

    # ------------------------------------------------------------------ decode
    def _stage_ptrs(self, kvs: Sequence[SessionKV]) -> None:
        """Point the static kv_ptrs / pos buffers (read by the captured graph) at these sessions.

        Pointer tables are cached on the device per session tuple (keyed by the caches' device
        addresses, which fully determine the table) and copied device-to-device; positions go
        through a small ring of pinned staging buffers guarded by events, because a pinned buffer
        must not be rewritten while an earlier asynchronous copy from it is still queued.
        """
        B = len(kvs)
        key = tuple(kv.data.data_ptr() for kv in kvs)
        slot = self._pos_ring_idx
        self._pos_ring_idx = (slot + 1) % len(self._pos_ring)
        host, ev = self._pos_ring[slot]
        ev.synchronize()
        n_copy = self._n_pos
        if key != self._ptr_key:
            tab = self._ptr_tables.get(key)       # host-side table per session tuple (the addresses determine it)
            if tab is None:
                # (a served stage batches whatever sessions are queued, so new tuples are the common case there: the
                # table is assembled from one cached [layers, 2] column per session, not element by element)
                nl = self.n_layers
                tab = torch.zeros(max(nl, 1), self.max_batch, 2, dtype=torch.int64)
                for b, kv in enumerate(kvs):
                    if kv.max_len != self.kv_capacity:
                        raise ValueError(f"session capacity {kv.max_len} != engine kv_capacity {self.kv_capacity}")
                    col = getattr(kv, "_ptr_col", None)
                    if col is None or col.shape[0] != nl or int(col[0, 0]) != kv.data.data_ptr():
                        base = kv.data.data_ptr()
                        per_layer = kv.data.stride(0) * kv.data.element_size()
                        half = kv.data.stride(1) * kv.data.element_size()
                        k_ptr = base + torch.arange(max(nl, 1), dtype=torch.int64) * per_layer
                        col = kv._ptr_col = torch.stack([k_ptr, k_ptr + half], 1)
                    if nl:
                        tab[:, b] = col
                if len(self._ptr_tables) > 256:
                    self._ptr_tables.clear()
                self._ptr_tables[key] = tab
            host[self._n_stage_head:].view(torch.int64).view_as(tab).copy_(tab)
            self._ptr_key = key
            n_copy = host.numel()                 # positions and pointers travel in the same copy
        for b, kv in enumerate(kvs):
            if kv.length + 1 > kv.max_len:
                raise ValueError(f"session {kv.session_id}: max_length {kv.max_len} reached")
            host[b] = kv.length
            host[self.max_batch + 1 + b] = getattr(kv, "sampler_row", 0) or 0
        self._seq_host = self._seq_host % 1000000 + 1        # a new step number for every staged step (never 0)
        host[self.max_batch] = self._seq_host
        self._stage_dev[:n_copy].copy_(host[:n_copy], non_blocking=True)
        ev.record(torch.cuda.current_stream(self.device))

    def prestage(self, kvs: Sequence[SessionKV]) -> None:
        """Queue the NEXT step's positions / cache pointers behind the step that was just launched, so a host loop
        that now blocks on the token does not pay for this staging between two steps."""
        self._stage_ptrs(kvs)
        self._prestaged = tuple((kv.data.data_ptr(), kv.length) for kv in kvs)

    def _decode_body(self, B: int, from_ids: bool, greedy: bool, prog: Optional[StepProgram] = None, dry: bool = False):
        if prog is not None and prog.pre is not None:
            prog.pre(B, dry)
        if prog is not None and not prog.layers:
            return None
        h = self._embed(B, 0, True) if from_ids else self.h_a[:B]
        self._chain_tail = None
        want_head = self.has_head and (prog is None or prog.head)
        h = self._decode_layers(B, h, with_head=want_head)
        if want_head:
            # (the warm-up pass before a capture must not draw: a draw advances the session's random stream)
            self._head(h, False if (dry and greedy == "sample") else greedy,
                       out_pair=getattr(prog, "out_pair", None) if prog is not None else None, gemm_done=self._head_done)
        self._chain_tail = None
        if prog is not None and prog.post is not None:
            prog.post(B, h, dry)
        return h

    def decode(self, kvs: Sequence[SessionKV], h_in: Optional[torch.Tensor] = None,
               ids: Optional[torch.Tensor] = None, greedy: bool = True, prog: Optional[StepProgram] = None):
        """One token for each session in ``kvs``.

        First stage: pass ``ids`` (int64 device tensor [B]) or leave ``ids_in`` pre-filled (e.g. by
        the last stage's arg-max over NVLink).  Other stages: ``h_in`` [B, H].
        Returns hidden states [B, H] (view of a static buffer) — or, on the last stage,
        ``(logits [B, V] fp32, ids_out [B] int64)`` device views.
        """
        with nvtx_range("b200.decode_step"):
            return self._decode_impl(kvs, h_in, ids, greedy, prog)

    def run_program(self, B: int, prog: StepProgram) -> None:
        """Replay a step that touches no session (``prog.layers`` is False), e.g. 'receive the token of slot j'."""
        assert not prog.layers
        key = (B, False, True, prog.key)
        g = self._graphs.get(key) if self.use_graphs else None
        if g is None and self.use_graphs:
            g = self._capture(key, prog)
        if g is not None:
            g.replay()
            self.kernel_launches += self._launches_per_graph[key]
        else:
            self._decode_body(B, False, True, prog)

    def _decode_impl(self, kvs, h_in, ids, greedy, prog=None):
        B = len(kvs)
        assert 1 <= B <= self.max_batch
        from_ids = (self.embed_w is not None and h_in is None) if prog is None else bool(prog.from_ids)
        staged = getattr(self, "_prestaged", None)
        self._prestaged = None
        if staged is None or staged != tuple((kv.data.data_ptr(), kv.length) for kv in kvs):
            self._stage_ptrs(kvs)
        key = (B, from_ids, greedy) if prog is None else (B, from_ids, greedy, prog.key)
        g = None
        if self.use_graphs:
            g = self._graphs.get(key)
            if g is None:
                g = self._capture(key, prog)   # runs the body once on whatever the buffers hold
                self._stage_ptrs(kvs)          # a fresh step number: the warm-up run left chain words tagged with the old one
        if from_ids and ids is not None:
            self.ids_in[:B].copy_(ids, non_blocking=True)
        if not from_ids and h_in is not None:
            self.h_a[:B].copy_(h_in, non_blocking=True)
        if g is not None:
            g.replay()
            self.kernel_launches += self._launches_per_graph[key]
            out_h = self._graph_out[key]
        else:
            out_h = self._decode_body(B, from_ids, greedy, prog)
        for kv in kvs:
            kv.length += 1
            kv.touch()
        if self.has_head and (prog is None or prog.head):
            return self.logits[:B], self.ids_out[:B]
        return out_h

    def _capture(self, key, prog: Optional[StepProgram] = None) -> torch.cuda.CUDAGraph:
        B, from_ids, greedy = key[:3]
        if not hasattr(self, "_graph_out"):
            self._graph_out = {}
        # warm-up outside capture (sets kernel attributes, loads modules)
        before = self.kernel_launches
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            self._decode_body(B, from_ids, greedy, prog, dry=True)
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        self.kernel_launches = before
        # (thread_local: other threads of a server or client process keep making CUDA calls while this one captures)
        with _capture_guard(), torch.cuda.graph(g, stream=s, capture_error_mode="thread_local"):
            out_h = self._decode_body(B, from_ids, greedy, prog)
        per = self.kernel_launches - before
        self.kernel_launches = before  # warm-up + capture passes touched KV slot `pos` twice; harmless: same data
        self._graphs[key] = g
        self._graph_out[key] = out_h
        self._launches_per_graph[key] = per
        return g

    # ------------------------------------------------------------------ device-resident greedy loop
    def decode_loop(self, kvs: Sequence[SessionKV], first_ids: torch.Tensor, n_steps: int) -> torch.Tensor:
        """Greedy-decode ``n_steps`` tokens for every session with no host round trip.

        Only for a span that holds the whole model (embedding and head): the arg-max of
        step i is copied to the embedding input of step i+1 and the positions advance on the
        device, all inside one CUDA graph that is replayed ``n_steps`` times.
        Returns the generated ids ``[n_steps, B]`` (device).
        """
        assert self.has_embed and self.has_head and self.embed_w is not None
        self._prestaged = None
        B = len(kvs)
        for kv in kvs:
            if kv.length + n_steps > kv.max_len:
                raise ValueError(f"session {kv.session_id}: {kv.length}+{n_steps} exceeds max_length {kv.max_len}")
        self._stage_ptrs(kvs)
        key = ("loop", B)
        g = self._graphs.get(key)
        if g is None:
            if not hasattr(self, "hist"):
                self.hist = torch.zeros(4096, self.max_batch, dtype=torch.int64, device=self.device)
                self.step_idx = torch.zeros(1, dtype=torch.int64, device=self.device)

            pos_inc = torch.zeros_like(self.pos)
            pos_inc[:B] = 1
            pos_inc[self.max_batch] = 1          # (the sampler rows behind it stay as they are)

            def body():
                self._decode_body(B, True, True)
                self.hist.index_copy_(0, self.step_idx, self.ids_out[:B].unsqueeze(0) if B == self.max_batch
                                      else torch.nn.functional.pad(self.ids_out[:B], (0, self.max_batch - B)).unsqueeze(0))
                self.ids_in[:B].copy_(self.ids_out[:B])
                self.pos.add_(pos_inc)            # positions of the B sessions and the chain's step number
                self.step_idx.add_(1)

            before = self.kernel_launches
            s = torch.cuda.Stream(device=self.device)
            s.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(s):
                body()
            torch.cuda.current_stream(self.device).wait_stream(s)
            torch.cuda.synchronize(self.device)
            per = self.kernel_launches - before
            g = torch.cuda.CUDAGraph()
            with _capture_guard(), torch.cuda.graph(g, stream=s, capture_error_mode="thread_local"):
                body()
            self.kernel_launches = before
            self._graphs[key] = g
            self._launches_per_graph[key] = per
            self._stage_ptrs(kvs)          # the warm-up advanced pos on the device: restore
        assert n_steps <= self.hist.shape[0]
        self.step_idx.zero_()
        self.ids_in[:B].copy_(first_ids, non_blocking=True)
        for _ in range(n_steps):
            g.replay()
        self.kernel_launches += n_steps * self._launches_per_graph[key]
        for kv in kvs:
            kv.length += n_steps
            kv.touch()
        return self.hist[:n_steps, :B]

    # ------------------------------------------------------------------ prefill
    head_only_tail = False # True on a first stage whose head shard serves the END of the chain (see PipelineEngine)
    last_hidden = None     # [1, H] view: final hidden row of the latest prefill on a span with a head
    head_out_pair = None   # [B, 2] int32: where the head's arg-max leaves its (value, index) candidate (vocabulary shards)

    def prefill(self, kv: SessionKV, h_in: Optional[torch.Tensor] = None, ids: Optional[torch.Tensor] = None,
                pos0: Optional[int] = None, greedy: bool = True):
        """Append T tokens of one session (chunked to ``max_chunk``).

        Returns hidden [T, H] (new tensor) for non-last stages, else (logits [1,V], ids_out[1])
        computed from the last position only.
        """
        c, pdl = self.cfg, self.pdl
        self._prestaged = None
        pos0 = kv.length if pos0 is None else pos0
        T = ids.shape[0] if h_in is None else h_in.shape[0]
        if pos0 + T > kv.max_len:
            raise ValueError(f"session {kv.session_id}: {pos0}+{T} tokens exceed max_length {kv.max_len}")
        if kv.max_len != self.kv_capacity:
            raise ValueError(f"session capacity {kv.max_len} != engine kv_capacity {self.kv_capacity}")
        tail_only = self.head_only_tail     # a first stage that holds a head shard for the chain's tail
        out = None if (self.has_head and not tail_only) else torch.empty(T, c.hidden, dtype=self.dtype, device=self.device)
        last_h = None
        for c0 in range(0, T, self.max_chunk):
            n = min(self.max_chunk, T - c0)
            p0 = pos0 + c0
            if h_in is None:
                self.ids_in[:n].copy_(ids[c0:c0 + n], non_blocking=True)
                h = self._embed(n, p0, False)
            else:
                h = self.h_a[:n]
                h.copy_(h_in[c0:c0 + n], non_blocking=True)
            other = self.h_b
            fused = n <= G.MAX_FUSED_NORM_TOKENS
            ln = not c.is_llama
            if (not fused and self.tp_mode is None and self.weight_format == "native" and self._planned_prefill
                    and not self.streamed_layers):
                self._prefill_chunk_planned(kv, n, p0)
                if out is not None:
                    out[c0:c0 + n].copy_(h, non_blocking=True)
                last_h = h[n - 1:n]
                continue
            for li, P in self._iter_layers():
                qkv, attn, act, h2 = self.qkv[:n], self.attn[:n], self.act[:n], other[:n]
                if fused:
                    G.linear(h, P["qkv"], out=qkv, norm_mode=self.norm_mode, norm_w=P["ln1_w"], norm_b=P["ln1_b"],
                             norm_eps=c.norm_eps, workspace=self.ws, pdl=pdl)
                else:
                    E.norm(h, P["ln1_w"], P["ln1_b"], c.norm_eps, ln, out=self.xn[:n], pdl=pdl)
                    G.linear(self.xn[:n], P["qkv"], out=qkv, workspace=self.ws, pdl=pdl)
                A.rope_append(qkv, kv.k(li), kv.v(li), self.inv_freq, c.n_heads, c.n_kv_heads, c.head_dim,
                              kv.max_len, p0, pdl=pdl)
                A.attn_prefill(qkv, attn, kv.k(li), kv.v(li), c.n_heads, c.n_kv_heads, c.head_dim, kv.max_len, p0,
                               self.scale, pdl=pdl)
                self._row_parallel(attn, P["o"], h2, h)
                if fused:
                    G.linear(h2, P["mlp_in"], out=act, epilogue=self.mlp_epi, norm_mode=self.norm_mode,
                             norm_w=P["ln2_w"], norm_b=P["ln2_b"], norm_eps=c.norm_eps, workspace=self.ws, pdl=pdl)
                else:
                    E.norm(h2, P["ln2_w"], P["ln2_b"], c.norm_eps, ln, out=self.xn[:n], pdl=pdl)
                    G.linear(self.xn[:n], P["mlp_in"], out=act, epilogue=self.mlp_epi, workspace=self.ws, pdl=pdl)
                self._row_parallel(act, P["down"], h, h2)
                self._count(6 if fused else 8)
            if out is not None:
                out[c0:c0 + n].copy_(h, non_blocking=True)
            last_h = h[n - 1:n]
        kv.length = pos0 + T
        kv.touch()
        if self.has_head and not tail_only:
            self.last_hidden = last_h
            ri = None
            if greedy == "sample":
                ri = torch.tensor([getattr(kv, "sampler_row", 0) or 0], dtype=torch.int32, device=self.device)
            self._head(last_h, greedy, out_pair=self.head_out_pair, row_index=ri)
            return self.logits[:1], self.ids_out[:1]
        return out

    # ------------------------------------------------------------------ cached prefill launches
    def _prefill_plan(self, n: int):
        """Argument structs of every launch of an n-token chunk, built once per chunk length.

        A prefill chunk is 8 launches per block; building their arguments through the generic wrappers costs
        ~15 us of Python each.  Only the cache pointers, the start position and the stream change from call to
        call, so everything else is kept as ready ctypes arguments.  Measured on Llama-2-7B: 2048-token prompt
        29.2 -> 27.6 ms, 4096 tokens 60.7 -> 58.5 ms; a 128-token prompt stays at 5.8 ms (GPU-bound: at 128 tokens
        each GEMM has only 32-96 output tiles for 148 SMs and the prefill kernel does not split K)."""
        plan = self._prefill_plans.get(n)
        if plan is not None:
            return plan
        import ctypes as C

        L = native.lib()
        c = self.cfg
        pdl = 1 if self.pdl else 0
        dt = E.dtype_code(self.dtype)
        ln = 0 if c.is_llama else 1
        h, other = self.h_a[:n], self.h_b[:n]
        qkv, attn, act, xn = self.qkv[:n], self.attn[:n], self.act[:n], self.xn[:n]
        ptr = lambda t: t.data_ptr() if t is not None else None   # noqa: E731
        layers = []
        for P in self.layers:
            gemms = []
            for x, w, kw in ((xn, P["qkv"], dict(out=qkv)), (attn, P["o"], dict(out=other, residual=h)),
                             (xn, P["mlp_in"], dict(out=act, epilogue=self.mlp_epi)),
                             (act, P["down"], dict(out=h, residual=other))):
                a, _ = G.fill_args(x, w, workspace=self.ws, pdl=self.pdl, **kw)
                fn = L.b200_gemm_prefill if G.uses_prefill_kernel(n, w) else L.b200_gemm
                gemms.append((fn, a, C.byref(a)))
            norm1 = (ptr(h), ptr(P["ln1_w"]), ptr(P["ln1_b"]), ptr(xn), n, c.hidden, h.stride(0), xn.stride(0),
                     c.norm_eps, ln, dt, pdl)
            norm2 = (ptr(other), ptr(P["ln2_w"]), ptr(P["ln2_b"]), ptr(xn), n, c.hidden, other.stride(0), xn.stride(0),
                     c.norm_eps, ln, dt, pdl)
            layers.append((norm1, gemms, norm2))
        tc = n >= A.TC_PREFILL_MIN_TOKENS and os.environ.get("B200_PREFILL_TC", "1") != "0"
        plan = dict(layers=layers, attn_fn=L.b200_attn_prefill_tc if tc else L.b200_attn_prefill,
                    qkv=ptr(qkv), attn=ptr(attn), ldqkv=qkv.stride(0), ldattn=attn.stride(0),
                    inv_freq=ptr(self.inv_freq), dt=dt, pdl=pdl)
        if len(self._prefill_plans) >= 64:      # one plan per distinct chunk length; keep the table small
            self._prefill_plans.clear()
        self._prefill_plans[n] = plan
        return plan

    def _prefill_chunk_planned(self, kv: SessionKV, n: int, p0: int) -> None:
        """Runs all blocks of one chunk from the cached plan: h_a[:n] in, h_a[:n] out."""
        L = native.lib()
        c = self.cfg
        plan = self._prefill_plan(n)
        stream = native.current_stream_ptr()
        base = kv.data.data_ptr()
        per_layer = kv.data.stride(0) * kv.data.element_size()
        half = kv.data.stride(1) * kv.data.element_size()
        q, o, ldq, ldo, inv, dt, pdl = (plan["qkv"], plan["attn"], plan["ldqkv"], plan["ldattn"], plan["inv_freq"],
                                        plan["dt"], plan["pdl"])
        attn_fn = plan["attn_fn"]
        nh, nkv, d, smax, scale = c.n_heads, c.n_kv_heads, c.head_dim, kv.max_len, self.scale
        check = native.check
        for li, (norm1, gemms, norm2) in enumerate(plan["layers"]):
            kc = base + li * per_layer
            vc = kc + half
            check(L.b200_norm(*norm1, stream), "b200_norm")
            fn, _, ref = gemms[0]
            check(fn(ref, stream), "b200_gemm")
            check(L.b200_rope_append(q, kc, vc, inv, n, nh, nkv, d, smax, ldq, p0, dt, pdl, stream), "b200_rope_append")
            check(attn_fn(q, o, kc, vc, n, nh, nkv, d, smax, ldq, ldo, p0, scale, dt, pdl, stream), "b200_attn_prefill")
            fn, _, ref = gemms[1]
            check(fn(ref, stream), "b200_gemm")
            check(L.b200_norm(*norm2, stream), "b200_norm")
            fn, _, ref = gemms[2]
            check(fn(ref, stream), "b200_gemm")
            fn, _, ref = gemms[3]
            check(fn(ref, stream), "b200_gemm")
        self._count(8 * len(plan["layers"]))

    # ------------------------------------------------------------------ TorchSpan-compatible API
    def forward(self, h: torch.Tensor, kv: SessionKV, pos0: Optional[int] = None) -> torch.Tensor:
        """[T, H] -> [T, H] for one session (decode kernel path when T == 1)."""
        assert not self.has_head
        if h.shape[0] == 1 and (pos0 is None or pos0 == kv.length):
            return self.decode([kv], h_in=h).clone()
        return self.prefill(kv, h_in=h, pos0=pos0)

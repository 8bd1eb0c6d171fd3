"""In-box pipeline engine: one process per GPU, stage-to-stage handoff over NCCL p2p.

The reference relays every hop through the client (star topology) with
GPU->CPU->serialize->libp2p->deserialize->GPU on both legs
(src/rpc_transport.py:736-766, 800-833).  Inside one NVSwitch box the
stages form a true chain instead:

    rank0: embed + blocks[0,s0)  --send[B,H]-->  rank1: blocks[s0,s1)  --> ... -->
    rank_{N-1}: blocks + final norm + LM head + arg-max  --send ids[B]-->  rank0

Hidden states and token ids never leave the devices.  Several sessions are kept in
flight round-robin (one per stage) so every GPU is busy while a session's token is
somewhere else in the chain; each rank's per-session work is one CUDA-graph replay.

The same class runs single-process (world size 1: the whole model on one GPU, with a
device-resident greedy loop) and on CPU with the gloo backend + TorchSpan, which is how
the host-side logic is tested without GPUs.

All ranks call every method with the same arguments (SPMD, like torchrun scripts).
"""
from __future__ import annotations

import time
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from ..models.config import ModelConfig, resolve_config
from ..runtime.kv_cache import KVCacheManager, SessionKV
from ..weights.loader import load_span


def _dist_ready() -> bool:
    return dist.is_available() and dist.is_initialized()


class PipelineEngine:
    def __init__(self, model: str, splits: Sequence[int] = (), dtype: torch.dtype = torch.float16, device=None, *,
                 backend: Optional[str] = None, kv_capacity: int = 2048, max_sessions: int = 8, max_chunk: int = 512,
                 seed: int = 0, use_graphs: bool = True, cfg: Optional[ModelConfig] = None,
                 weight_format: str = "native", tp: int = 1, tp_mode: str = "fused", hop: Optional[str] = None,
                 head_split: Optional[bool] = None, attn_splits: int = 0):
        self.cfg = cfg or resolve_config(model)
        grank = dist.get_rank() if _dist_ready() else 0
        gworld = dist.get_world_size() if _dist_ready() else 1
        # Tensor-parallel spans: `tp` consecutive global ranks form one pipeline stage.  Every rank of a
        # stage holds the same (replicated) hidden state, so stage s rank i simply talks to stage s+1 rank i.
        self.tp = max(1, int(tp))
        if gworld % self.tp:
            raise ValueError(f"world size {gworld} is not a multiple of tp={self.tp}")
        self.tp_rank = grank % self.tp
        self.rank = grank // self.tp            # pipeline stage index
        self.world = gworld // self.tp          # number of pipeline stages
        self.grank, self.gworld = grank, gworld
        self.tp_group = None
        if self.tp > 1:
            for st in range(self.world):       # every rank must create every group
                g = dist.new_group(list(range(st * self.tp, (st + 1) * self.tp)))
                if st == self.rank:
                    self.tp_group = g
        bounds = [0] + [int(s) for s in splits] + [self.cfg.n_layers]
        if len(bounds) - 1 != self.world:
            raise ValueError(f"{len(splits)} split points make {len(bounds) - 1} stages but there are {self.world} "
                             f"pipeline stages (world {gworld} / tp {self.tp})")
        if any(b >= e for b, e in zip(bounds, bounds[1:])):
            raise ValueError(f"empty stage in splits {list(splits)} for a {self.cfg.n_layers}-layer model")
        self.bounds = bounds
        self.start, self.end = bounds[self.rank], bounds[self.rank + 1]
        self.is_first, self.is_last = self.rank == 0, self.rank == self.world - 1
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        self.device = torch.device(device)
        self.dtype = dtype
        self.backend = backend or ("native" if self.device.type == "cuda" else "torch")
        self.kv_capacity = kv_capacity
        self.max_sessions = max_sessions
        # Decode hops: "p2p" = peer stores + polling kernels captured into each stage's graph (parallel/hop.py),
        # "nccl" = send/recv issued by the host around the graph (also the CPU/gloo path).  Prefill chunks always
        # travel over NCCL.  B200_PIPE_HOP overrides the default.
        import os as _os

        if hop is None:
            hop = _os.environ.get("B200_PIPE_HOP", "p2p")
        if self.backend != "native" or self.world == 1 or not use_graphs:
            hop = "nccl"
        if hop not in ("p2p", "nccl"):
            raise ValueError("hop must be 'p2p' or 'nccl'")
        self.hop = hop
        # Vocabulary-sharded LM head: the last stage multiplies rows [0, V0) and hands its arg-max candidate and the
        # final hidden state to stage 0, which multiplies rows [V0, V) at the head of its next step - the head's
        # weight stream is shared by the two ends of the chain instead of making the last stage the slow one.
        if head_split is None:
            head_split = _os.environ.get("B200_HEAD_SPLIT", "1") != "0"
        self.head_split = bool(head_split) and self.hop == "p2p" and self.world > 1
        tiles = (self.cfg.vocab + 127) // 128
        # share of the vocabulary the LAST stage multiplies (the first stage also embeds and merges the candidates,
        # so an even split leaves it the slower of the two; B200_HEAD_SPLIT_LAST_FRAC to measure other splits)
        frac = float(_os.environ.get("B200_HEAD_SPLIT_LAST_FRAC", "0.5"))
        self.v0 = max(1, min(tiles - 1, int(round(tiles * frac)))) * 128
        need_head = self.is_last or (self.is_first and self.head_split)
        head_rows = None
        if self.head_split and need_head:
            head_rows = (0, self.v0) if self.is_last else (self.v0, self.cfg.vocab)
        w = load_span(model, self.cfg, self.start, self.end, need_embed=self.is_first, need_head=need_head,
                      dtype=dtype, device=self.device, seed=seed)
        if self.backend == "native" and self.tp > 1:
            from .tensor_parallel import make_tp_span

            self.span = make_tp_span(self.cfg, w, self.device, dtype, self.tp_group, mode=tp_mode,
                                     kv_capacity=kv_capacity, max_batch=max_sessions, max_chunk=max_chunk,
                                     use_graphs=use_graphs, head_rows=head_rows, attn_splits=attn_splits)
        elif self.backend == "native":
            from ..runtime.native_span import NativeSpan

            self.span = NativeSpan(self.cfg, w, self.device, dtype, kv_capacity=kv_capacity,
                                   max_batch=max_sessions, max_chunk=max_chunk, use_graphs=use_graphs,
                                   weight_format=weight_format, head_rows=head_rows, attn_splits=attn_splits)
        else:
            from ..models.torch_ref import TorchSpan

            self.span = TorchSpan(self.cfg, w, self.device, dtype)
        if self.tp > 1 and self.backend != "native":
            raise ValueError("tensor-parallel spans need the native (CUDA) backend")
        self.kv = KVCacheManager(self.end - self.start, self.cfg.n_kv_heads // self.tp, self.cfg.head_dim, dtype,
                                 self.device, default_max_length=kv_capacity)
        self.sessions: List[SessionKV] = []
        H = self.cfg.hidden
        # per-slot staging buffers: [0] = received from the previous stage, [1] = sent to the next
        self._hid = torch.empty(2, max_sessions, H, dtype=dtype, device=self.device)
        self._tok = torch.zeros(max_sessions, dtype=torch.int64, device=self.device)
        self._tok_send = torch.zeros(max_sessions, dtype=torch.int64, device=self.device)
        # Hidden states flow on the default group, token ids return on a second communicator:
        # with one communicator the ring (last -> first) closes a cycle of in-order NCCL ops.
        self.tok_group = dist.new_group() if self.world > 1 else None
        # Links alternate between two more communicators: a middle stage then receives (link r-1) and sends
        # (link r) on different NCCL streams - on one process group unbatched p2p ops are serialised with each
        # other.  Receives are posted from an otherwise idle side stream, so the next micro-batch's receive does
        # not queue behind the compute that is already enqueued on the main stream.
        self.link_groups = [dist.new_group(), dist.new_group()] if self.world > 1 else [None, None]
        self._recv_stream = torch.cuda.Stream(device=self.device) if (self.device.type == "cuda" and self.world > 1) else None
        self._slot_done = [None] * max_sessions
        self._send_work = [None] * max_sessions
        self._tok_work = [None] * max_sessions
        self._tok_host = torch.zeros(max_sessions, dtype=torch.int64)
        if self.device.type == "cuda":
            self._tok_host = self._tok_host.pin_memory()
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.links = None
        self._greedy = True          # True, or "sample": the last stage samples on the device (see generate)
        self._tail_pending: dict = {}
        if self.hop == "p2p":
            self._init_p2p()

    # ------------------------------------------------------------------ in-graph hops
    def _init_p2p(self) -> None:
        from ..runtime.native_span import MAX_DECODE_BATCH, StepProgram
        from .hop import HopLinks

        sp = self.span
        H = self.cfg.hidden
        t_max = min(self.max_sessions, MAX_DECODE_BATCH)
        self.links = hl = HopLinks(self.device, dist.group.WORLD, H, depth=self.max_sessions + 2, t_max=t_max)
        nxt = self._g(self.rank + 1) if not self.is_last else -1
        first = self._g(0)
        mb = sp.max_batch
        self._pair = torch.zeros(mb, 2, dtype=torch.int32, device=self.device)      # arg-max candidate (value bits, index)
        self._h_tail = torch.zeros(mb, H, dtype=self.dtype, device=self.device)     # final hidden state of the chain
        split = self.head_split
        if split and self.is_first:
            sp.head_only_tail = True
        if split and self.is_last:
            sp.head_out_pair = self._pair

        def recv_h(B, dry):
            if not dry:
                hl.recv_hidden(sp.h_a[:B], pdl=sp.pdl)
                sp._count()

        def send_h(B, h, dry):
            if not dry:
                hl.send_hidden(h, nxt, pdl=sp.pdl)
                sp._count()

        def send_tail(B, h, dry):
            if dry:
                return
            if split:
                hl.send_tail(h, self._pair, 2, B, first, pdl=sp.pdl)
            else:
                hl.send_tail(None, sp.ids_out.view(torch.int32), 2, B, first, pdl=sp.pdl)
            sp._count()

        def recv_tail(B, dry):
            # the new tokens of this micro-batch -> ids_in[:B]
            if split:
                if not dry:
                    hl.recv_tail(self._h_tail[:B], self._pair, 2, B, pdl=sp.pdl)
                    sp._count()
                sp._head(self._h_tail[:B], True, out_ids=sp.ids_in[:B], init_pair=self._pair)
            elif not dry:
                hl.recv_tail(None, sp.ids_in.view(torch.int32), 2, B, pdl=sp.pdl)
                sp._count()

        P = StepProgram
        if self.is_first:
            self._prog_chain = P("p2p_first_chain", True, pre=recv_tail, post=send_h, head=False)
            self._prog_host = P("p2p_first_host", True, post=send_h, head=False)
            self._prog_tail = P("p2p_first_tail", False, pre=recv_tail, layers=False, head=False)
        elif self.is_last:
            self._prog_last = P("p2p_last", False, pre=recv_h, post=send_tail)
            self._prog_last.out_pair = self._pair if split else None
        else:
            self._prog_mid = P("p2p_mid", False, pre=recv_h, post=send_h)

    def _collect_slot(self, j0: int) -> None:
        """Rank 0: receive the tokens the chain produced for the micro-batch starting at slot j0 into _tok."""
        B = self._tail_pending.pop(j0, 0)
        if B:
            self.span.run_program(B, self._prog_tail)
            self._tok[j0:j0 + B].copy_(self.span.ids_in[:B])

    def _round_p2p(self, tokens, sess, to_host: bool, group: int):
        S = len(sess)
        g = max(1, min(int(group), self.links.t_max))
        sp = self.span
        for j0 in range(0, S, g):
            j1 = min(S, j0 + g)
            B = j1 - j0
            kv = [self.sessions[si] for si in sess[j0:j1]]
            if self.is_first:
                if tokens is not None:
                    self._collect_slot(j0)             # a token nobody asked for must still leave the ring
                    self._tok_host[j0:j1] = torch.as_tensor(list(tokens[j0:j1]), dtype=torch.int64)
                    sp.ids_in[:B].copy_(self._tok_host[j0:j1], non_blocking=True)
                    self.h2d_bytes += 8 * B
                    sp.decode(kv, prog=self._prog_host)
                elif self._tail_pending.get(j0) == B:
                    self._tail_pending.pop(j0)
                    sp.decode(kv, prog=self._prog_chain)
                else:
                    self._collect_slot(j0)
                    sp.ids_in[:B].copy_(self._tok[j0:j1])
                    sp.decode(kv, prog=self._prog_host)
                self._tail_pending[j0] = B
            elif self.is_last:
                sp.decode(kv, prog=self._prog_last, greedy=self._greedy)
            else:
                sp.decode(kv, prog=self._prog_mid)
        if self.is_first and to_host:
            for j0 in range(0, S, g):
                self._collect_slot(j0)
            self._tok_host[:S].copy_(self._tok[:S])
            self.d2h_bytes += 8 * S
            return [int(x) for x in self._tok_host[:S]]
        return None

    def _streaming_p2p(self, tokens, sess):
        S = len(sess)
        sp = self.span
        out = [None] * S
        have_prev = getattr(self, "_streaming_live", False)
        for j, si in enumerate(sess):
            kv = [self.sessions[si]]
            if self.is_first:
                if self._tail_pending.pop(j, 0):
                    sp.run_program(1, self._prog_tail)
                    self._tok_host[j:j + 1].copy_(sp.ids_in[:1])          # D2H of session j's result (blocks on it only)
                    self.d2h_bytes += 8
                    out[j] = int(self._tok_host[j])
                else:
                    self._tok_host[j] = int(tokens[j])
                sp.ids_in[:1].copy_(self._tok_host[j:j + 1], non_blocking=True)   # H2D of the step's input
                self.h2d_bytes += 8
                sp.decode(kv, prog=self._prog_host)
                self._tail_pending[j] = 1
            elif self.is_last:
                sp.decode(kv, prog=self._prog_last, greedy=self._greedy)
            else:
                sp.decode(kv, prog=self._prog_mid)
        self._streaming_live = True
        return out if (self.is_first and have_prev) else None

    # ------------------------------------------------------------------ sessions
    def open_sessions(self, n: int) -> List[int]:
        base = len(self.sessions)
        for i in range(n):
            self.sessions.append(self.kv.allocate(f"sess{base + i}", self.kv_capacity))
        return list(range(base, base + n))

    def close_sessions(self) -> None:
        for s in self.sessions:
            self.kv.close(s.session_id)
        self.sessions = []

    # ------------------------------------------------------------------ p2p helpers
    def _g(self, stage: int) -> int:
        """Global rank of this rank's counterpart in pipeline stage ``stage``."""
        return stage * self.tp + self.tp_rank

    def _link(self, src_stage: int):
        """Communicator of the link src_stage -> src_stage + 1."""
        return self.link_groups[src_stage % 2]

    def _post_recv(self, buf: torch.Tensor, slot: int):
        """Asynchronous receive of a micro-batch's hidden state from the previous stage."""
        src = self.rank - 1
        if self._recv_stream is None:
            return dist.irecv(buf, self._g(src), group=self._link(src))
        if self._slot_done[slot] is not None:      # the previous user of this staging slot has consumed it
            self._recv_stream.wait_event(self._slot_done[slot])
        with torch.cuda.stream(self._recv_stream):
            return dist.irecv(buf, self._g(src), group=self._link(src))

    def _mark_consumed(self, slot: int) -> None:
        if self._recv_stream is not None:
            ev = self._slot_done[slot] or torch.cuda.Event()
            ev.record()
            self._slot_done[slot] = ev

    def _send(self, t: torch.Tensor, dst: int, group=None) -> None:
        dist.send(t, self._g(dst), group=group)

    def _recv(self, t: torch.Tensor, src: int, group=None) -> None:
        dist.recv(t, self._g(src), group=group)

    @staticmethod
    def _wait(work) -> None:
        if work is not None:
            work.wait()

    # ------------------------------------------------------------------ per-stage compute
    def _stage_prefill(self, kv: SessionKV, ids: Optional[torch.Tensor], h: Optional[torch.Tensor]):
        sp = self.span
        if self.backend == "native":
            return sp.prefill(kv, h_in=h, ids=ids, greedy=self._greedy)
        if self.is_first:
            h = sp.embed(ids, kv.length)
        h = sp.forward(h, kv)
        if self.is_last:
            logits = sp.head(h[-1:])
            return logits, logits.argmax(-1)
        return h

    def _stage_decode(self, kvs: Sequence[SessionKV], ids: Optional[torch.Tensor], h: Optional[torch.Tensor]):
        sp = self.span
        if self.backend == "native":
            return sp.decode(kvs, h_in=h, ids=ids, greedy=self._greedy)
        outs = []
        for b, kv in enumerate(kvs):
            hb = sp.embed(ids[b:b + 1], kv.length) if self.is_first else h[b:b + 1]
            outs.append(sp.forward(hb, kv))
        hcat = torch.cat(outs, 0)
        if self.is_last:
            logits = sp.head(hcat)
            return logits, logits.argmax(-1)
        return hcat

    # ------------------------------------------------------------------ public API
    def prefill(self, session: int, prompt_ids: torch.Tensor, micro_batch: Optional[int] = None) -> Optional[int]:
        """Run a prompt through the chain; returns the first generated token on rank 0
        (None on the other ranks).  ``prompt_ids``: 1-D int64, same on every rank.

        The prompt travels in micro-batches of ``micro_batch`` tokens (default: the engine's chunk
        size): while stage s+1 works on chunk c, stage s already computes chunk c+1 and the
        hand-off of chunk c rides NCCL in the background — upstream Petals chunks the prompt in time
        only to bound the score matrix (petals/server/backend.py:129-152); here the chunks also fill
        the pipeline."""
        kv = self.sessions[session]
        L = int(prompt_ids.shape[0])
        mb = int(micro_batch or getattr(self.span, "max_chunk", 0) or 512)
        chunks = [(c0, min(mb, L - c0)) for c0 in range(0, L, mb)]
        ids_dev = None
        if self.is_first:
            ids_dev = prompt_ids.to(self.device, non_blocking=True)
            self.h2d_bytes += L * 8
        H = self.cfg.hidden
        tok = self._tok[:1]
        pending = []                       # keep send buffers alive until their isend completed
        out = None
        recv_next = None
        if not self.is_first:
            buf = torch.empty(chunks[0][1], H, dtype=self.dtype, device=self.device)
            recv_next = (buf, dist.irecv(buf, self._g(self.rank - 1)))
        for ci, (c0, n) in enumerate(chunks):
            h = None
            if not self.is_first:
                h, work = recv_next
                work.wait()
                if ci + 1 < len(chunks):   # post the next receive before computing this chunk
                    nb = torch.empty(chunks[ci + 1][1], H, dtype=self.dtype, device=self.device)
                    recv_next = (nb, dist.irecv(nb, self._g(self.rank - 1)))
            out = self._stage_prefill(kv, ids_dev[c0:c0 + n] if self.is_first else None, h)
            if not self.is_last:
                o = out.contiguous()
                pending.append((o, dist.isend(o, self._g(self.rank + 1))))
        for _, w in pending:
            w.wait()
        if self.head_split:
            # vocabulary-sharded head: the last stage's candidate and the final hidden row go to stage 0
            if self.is_last:
                self._send(self.span.last_hidden.contiguous(), 0, self.tok_group)
                self._send(self._pair[:1], 0, self.tok_group)
            if self.is_first:
                self._recv(self._h_tail[:1], self.world - 1, self.tok_group)
                self._recv(self._pair[:1], self.world - 1, self.tok_group)
                self.span._head(self._h_tail[:1], True, out_ids=tok, init_pair=self._pair)
        else:
            if self.is_last:
                tok.copy_(out[1].reshape(-1)[:1])
                if not self.is_first:
                    self._send(tok, 0, self.tok_group)
            if self.is_first and not self.is_last:
                self._recv(tok, self.world - 1, self.tok_group)
        if self.is_first:
            self.d2h_bytes += 8
            return int(tok.cpu()[0])
        return None

    def decode_round(self, tokens: Optional[Sequence[int]] = None, sessions: Optional[Sequence[int]] = None,
                     to_host: bool = True, group: int = 1):
        """One new token for each listed session (default: all), one session per pipeline slot.

        ``tokens``: the previous token of every session (host ints, read on rank 0); when
        None the tokens already resident on the first-stage GPU are used (device loop).
        Returns the new tokens as a list on rank 0 when ``to_host`` (one D2H copy), else None.
        """
        sess = list(range(len(self.sessions))) if sessions is None else list(sessions)
        S = len(sess)
        if self.hop == "p2p":
            return self._round_p2p(tokens, sess, to_host, group)
        if self.world == 1 and tokens is not None and to_host and self.backend == "native":
            # Single-GPU host loop, the latency path: the pinned token buffer is copied straight into the captured
            # graph's input and the result straight out of its output - no device-to-device staging in between.
            kvs = [self.sessions[i] for i in sess]
            for j, t in enumerate(tokens):
                self._tok_host[j] = int(t)
            self.span.ids_in[:S].copy_(self._tok_host[:S], non_blocking=True)
            self.h2d_bytes += 8 * S
            out = self.span.decode(kvs, ids=None, greedy=self._greedy)
            self.span.prestage(kvs)                                   # next step's positions, off the critical path
            self._tok_host[:S].copy_(out[1].reshape(-1)[:S])          # blocking D2H
            self.d2h_bytes += 8 * S
            self._tok_stale = True                                    # the device copy is refreshed on demand
            return [int(x) for x in self._tok_host[:S]]
        if self.is_first and tokens is not None:
            self._tok_host[:S] = torch.as_tensor(list(tokens), dtype=torch.int64)
            self._tok[:S].copy_(self._tok_host[:S], non_blocking=True)
            self.h2d_bytes += 8 * S
        elif self.is_first and getattr(self, "_tok_stale", False):
            self._tok[:S].copy_(self._tok_host[:S], non_blocking=True)
        self._tok_stale = False
        if self.world == 1:
            kvs = [self.sessions[i] for i in sess]
            out = self._stage_decode(kvs, self._tok[:S], None)
            self._tok[:S].copy_(out[1].reshape(-1)[:S])
        else:
            # Asynchronous chain: sends never block the compute stream, receives are waited on right
            # before use, and every slot has its own staging buffers so a send may still be in
            # flight while the next micro-batch is computed.  A micro-batch is `group` sessions that
            # travel together as one [group, H] activation (one batched kernel sequence per stage).
            last = self.world - 1
            g = max(1, int(group))
            starts = list(range(0, S, g))
            pending = None
            if not self.is_first and starts:
                pending = self._post_recv(self._hid[0, starts[0]:min(S, starts[0] + g)], starts[0])
            for bi, j0 in enumerate(starts):
                j1 = min(S, j0 + g)
                kv = [self.sessions[si] for si in sess[j0:j1]]
                slot = j0
                if self.is_first:
                    self._wait(self._tok_work[slot])                  # tokens of this micro-batch from the previous round
                    self._tok_work[slot] = None
                    out = self._stage_decode(kv, self._tok[j0:j1], None)
                    self._wait(self._send_work[slot])
                    self._hid[1, j0:j1].copy_(out[: j1 - j0])
                    self._send_work[slot] = dist.isend(self._hid[1, j0:j1], self._g(1), group=self._link(0))
                    self._tok_work[slot] = dist.irecv(self._tok[j0:j1], self._g(last), group=self.tok_group)
                else:
                    h = self._hid[0, j0:j1]
                    work, pending = pending, None
                    if bi + 1 < len(starts):                          # the next micro-batch's receive overlaps this one's compute
                        n0 = starts[bi + 1]
                        pending = self._post_recv(self._hid[0, n0:min(S, n0 + g)], n0)
                    work.wait()
                    out = self._stage_decode(kv, None, h)
                    self._mark_consumed(slot)
                    self._wait(self._send_work[slot])
                    if self.is_last:
                        self._tok_send[j0:j1].copy_(out[1].reshape(-1)[: j1 - j0])
                        self._send_work[slot] = dist.isend(self._tok_send[j0:j1], self._g(0), group=self.tok_group)
                    else:
                        self._hid[1, j0:j1].copy_(out[: j1 - j0])
                        self._send_work[slot] = dist.isend(self._hid[1, j0:j1], self._g(self.rank + 1),
                                                           group=self._link(self.rank))
            if self.is_first and (to_host or tokens is not None):
                for j in range(S):
                    self._wait(self._tok_work[j])
                    self._tok_work[j] = None
        if self.is_first and to_host:
            self._tok_host[:S].copy_(self._tok[:S])
            self.d2h_bytes += 8 * S
            return [int(x) for x in self._tok_host[:S]]
        return None

    def decode_round_streaming(self, tokens: Optional[Sequence[int]] = None, sessions: Optional[Sequence[int]] = None):
        """Like :meth:`decode_round` with host-visible tokens, but without a barrier between rounds.

        The first stage issues session j's next step as soon as session j's previous token has been
        copied to the host (D2H) and back (H2D from pinned memory) — it does not wait for the other
        sessions of the round, so the chain stays full across rounds.  ``tokens`` seeds the first call;
        afterwards every call returns, on rank 0, the tokens produced by the PREVIOUS call (``None`` the
        first time); call :meth:`flush_streaming` to collect the last ones.
        """
        sess = list(range(len(self.sessions))) if sessions is None else list(sessions)
        S = len(sess)
        if self.world == 1:
            return self.decode_round(tokens, sessions)
        if self.hop == "p2p":
            return self._streaming_p2p(tokens, sess)
        last = self.world - 1
        out = [None] * S
        have_prev = getattr(self, "_streaming_live", False)
        pending = None
        if not self.is_first and S:
            pending = self._post_recv(self._hid[0, 0:1], 0)
        for j, si in enumerate(sess):
            kv = [self.sessions[si]]
            if self.is_first:
                if have_prev:
                    self._wait(self._tok_work[j])
                    self._tok_work[j] = None
                    self._tok_host[j:j + 1].copy_(self._tok[j:j + 1])       # D2H of session j's result (blocks on it only)
                    self.d2h_bytes += 8
                    out[j] = int(self._tok_host[j])
                else:
                    self._tok_host[j] = int(tokens[j])
                self._tok[j:j + 1].copy_(self._tok_host[j:j + 1], non_blocking=True)   # H2D of the step's input
                self.h2d_bytes += 8
                o = self._stage_decode(kv, self._tok[j:j + 1], None)
                self._wait(self._send_work[j])
                self._hid[1, j:j + 1].copy_(o[:1])
                self._send_work[j] = dist.isend(self._hid[1, j:j + 1], self._g(1), group=self._link(0))
                self._tok_work[j] = dist.irecv(self._tok[j:j + 1], self._g(last), group=self.tok_group)
            else:
                h = self._hid[0, j:j + 1]
                work, pending = pending, None
                if j + 1 < S:
                    pending = self._post_recv(self._hid[0, j + 1:j + 2], j + 1)
                work.wait()
                o = self._stage_decode(kv, None, h)
                self._mark_consumed(j)
                self._wait(self._send_work[j])
                if self.is_last:
                    self._tok_send[j:j + 1].copy_(o[1].reshape(-1)[:1])
                    self._send_work[j] = dist.isend(self._tok_send[j:j + 1], self._g(0), group=self.tok_group)
                else:
                    self._hid[1, j:j + 1].copy_(o[:1])
                    self._send_work[j] = dist.isend(self._hid[1, j:j + 1], self._g(self.rank + 1),
                                                    group=self._link(self.rank))
        self._streaming_live = True
        return out if (self.is_first and have_prev) else None

    def flush_streaming(self, sessions: Optional[Sequence[int]] = None):
        """Collect the tokens of the last streaming round (rank 0) and leave streaming mode."""
        S = len(self.sessions) if sessions is None else len(sessions)
        out = None
        if self.hop == "p2p":
            if self.is_first and getattr(self, "_streaming_live", False):
                for j in sorted(self._tail_pending):
                    self._collect_slot(j)
                self._tok_host[:S].copy_(self._tok[:S])
                self.d2h_bytes += 8 * S
                out = [int(x) for x in self._tok_host[:S]]
            self._streaming_live = False
            return out
        if self.is_first and self.world > 1 and getattr(self, "_streaming_live", False):
            for j in range(S):
                self._wait(self._tok_work[j])
                self._tok_work[j] = None
            self._tok_host[:S].copy_(self._tok[:S])
            self.d2h_bytes += 8 * S
            out = [int(x) for x in self._tok_host[:S]]
        self._streaming_live = False
        return out

    def drain(self) -> None:
        """Wait for every outstanding p2p operation (call before reading staging buffers directly)."""
        if self.hop == "p2p":
            if self.is_first:
                for j in sorted(self._tail_pending):      # tokens still in the ring -> _tok
                    self._collect_slot(j)
            return
        for lst in (self._send_work, self._tok_work):
            for j, w in enumerate(lst):
                self._wait(w)
                lst[j] = None

    def generate(self, prompts: Sequence[torch.Tensor], max_new_tokens: int, eos_token_id: Optional[int] = None,
                 streaming: bool = False, sampling=None, seed: int = 0, eos_check_every: int = 8):
        """Generation for a batch of prompts (one session each).  Returns, on rank 0, ``(tokens per prompt, stats)``
        with TTFT / decode timings like the reference client prints (src/main.py:213-225); None elsewhere.

        ``sampling``: a :class:`~petals_b200.runtime.sampler.SamplingParams`; with ``temperature > 0`` the LAST stage
        samples on the device (``b200_sample_stateful``: repetition penalty, temperature, top-k, top-p with the
        reference's rules, src/rpc_handler.py:327-403) from a per-session Philox stream seeded with ``seed`` +
        session index, so the tokens do not depend on the number of stages.  Otherwise greedy.
        ``eos_token_id``: generation stops once every session has produced it; the decision is rank 0's and is
        shared with the other ranks every ``eos_check_every`` rounds (they run ahead in between)."""
        self.close_sessions()
        ids = self.open_sessions(len(prompts))
        sample = sampling is not None and sampling.temperature > 0.0
        if sample:
            if self.backend != "native":
                raise NotImplementedError("device-side sampling runs on the native backend")
            if self.head_split:
                raise ValueError("sampling needs the whole vocabulary on the last stage: build the engine with "
                                 "head_split=False")
            if self.is_last:
                pool = self.span.enable_sampling(4 * self.max_sessions)
                for i in ids:
                    self.sessions[i].sampler_row = pool.open(sampling, seed + i)
        self._greedy = "sample" if sample else True
        t0 = time.perf_counter()
        first = [self.prefill(i, p) for i, p in zip(ids, prompts)]
        ttft = time.perf_counter() - t0
        outs = [[t] for t in first] if self.is_first else None
        cur = first
        t1 = time.perf_counter()
        if streaming and self.world > 1:
            for _ in range(max_new_tokens - 1):
                got = self.decode_round_streaming(cur)
                if self.is_first and got:
                    for o, t in zip(outs, got):
                        o.append(t)
            got = self.flush_streaming()
            if self.is_first and got:
                for o, t in zip(outs, got):
                    o.append(t)
        for step in range(0 if (streaming and self.world > 1) else max_new_tokens - 1):
            cur = self.decode_round(cur)
            if self.is_first:
                for o, t in zip(outs, cur):
                    o.append(t)
            if eos_token_id is not None:
                done = self.is_first and all(eos_token_id in o for o in outs)
                if self.world == 1:
                    if done:
                        break
                elif (step + 1) % max(1, eos_check_every) == 0:
                    # rank 0 decides, everybody learns it at the same round (the chain stays in lock step)
                    flag = torch.tensor([1 if done else 0], dtype=torch.int32, device=self.device)
                    dist.broadcast(flag, src=0)
                    if int(flag.item()):
                        break
        dec = time.perf_counter() - t1
        if sample and self.is_last:
            for i in ids:
                self.span.sampler.close(self.sessions[i].sampler_row)
        self._greedy = True
        if self.is_first:
            if eos_token_id is not None:
                outs = [o[: o.index(eos_token_id) + 1] if eos_token_id in o else o for o in outs]
            return outs, {"ttft_s": ttft, "decode_s": dec, "total_s": ttft + dec,
                          "decode_tok_per_s": (max_new_tokens - 1) * len(prompts) / max(dec, 1e-9)}
        return None

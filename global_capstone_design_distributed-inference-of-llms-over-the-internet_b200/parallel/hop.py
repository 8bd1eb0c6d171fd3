"""Stage-to-stage links of the in-box pipeline that live INSIDE the stages' CUDA graphs.

Each rank owns two receive rings in symmetric (peer-mapped) memory:

    hidden ring   written by the previous stage   [depth][t_max tokens][hidden/4 + 2 cells of 16 B]
    tail ring     written by the LAST stage, read by stage 0: the new token ids, or - when the LM head is
                  sharded over the last and the first stage - the final hidden state plus the last stage's
                  arg-max candidate (value, index)

``send_*`` / ``recv_*`` enqueue ``b200_hop_send`` / ``b200_hop_recv`` (csrc/p2p.cu) on the current stream, so a
stage's whole slot - receive, blocks, send - is one graph replay and the host does nothing else per token.
The NCCL send/recv path of :class:`PipelineEngine` (``hop="nccl"``) is kept next to it for comparison and
for the prefill chunks.

Replaces the reference's hop through the client: src/rpc_transport.py:736-766, src/rpc_handler.py:139-147.
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..ops import native
from ..ops._prototypes import HopArgs

N_EXTRA_MAX = 4


class HopLinks:
    def __init__(self, device, group, hidden: int, depth: int, t_max: int, timeout_ms: int = 30000):
        import torch.distributed._symmetric_memory as symm

        L = native.lib()
        self.device = torch.device(device)
        self.group = group or dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.hidden, self.depth, self.t_max = hidden, int(depth), int(t_max)
        self.pairs = hidden // 4 + N_EXTRA_MAX // 2
        self.timeout_ms = timeout_ms
        ring_bytes = int(L.b200_hop_ring_bytes(self.depth, self.t_max, self.pairs))
        self.ring_bytes = ring_bytes
        # one symmetric allocation: [hidden ring | tail ring]
        self.buf = symm.empty(2 * ring_bytes // 8, dtype=torch.int64, device=self.device)
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, self.group.group_name)
        off = int(getattr(self.handle, "offset", 0) or 0)
        self.base = [int(self.handle.buffer_ptrs[r]) + off for r in range(self.world)]
        if self.base[self.rank] != self.buf.data_ptr():
            raise RuntimeError("symmetric-memory handle does not address the local ring")
        # [sent hidden, received hidden, sent tail, received tail]
        self.ctr = torch.zeros(4, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            native.check(L.b200_hop_init(), "b200_hop_init")
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)

    def _args(self, x, extra, n_extra, ring_ptr, ctr_idx, T, pdl) -> HopArgs:
        a = HopArgs()
        a.x = x.data_ptr() if x is not None else None
        a.extra = extra.data_ptr() if extra is not None else None
        a.ring = ring_ptr
        a.counter = self.ctr.data_ptr() + 4 * ctr_idx
        a.T = T
        a.hidden = self.hidden if x is not None else 0
        a.ldx = x.stride(0) if x is not None else 0
        a.n_extra = n_extra
        a.depth, a.t_max, a.pairs_per_token = self.depth, self.t_max, self.pairs
        a.timeout_ms = self.timeout_ms
        a.pdl = 1 if pdl else 0
        return a

    def _launch(self, fn, name, a) -> None:
        native.check(fn(ctypes.byref(a), native.current_stream_ptr()), name)

    # hidden link: this rank -> `dst` (the same index in the next stage)
    def send_hidden(self, x: torch.Tensor, dst: int, pdl: bool = True) -> None:
        a = self._args(x, None, 0, self.base[dst], 0, x.shape[0], pdl)
        self._launch(native.lib().b200_hop_send, "b200_hop_send", a)

    def recv_hidden(self, x_out: torch.Tensor, pdl: bool = True) -> None:
        a = self._args(x_out, None, 0, self.base[self.rank], 1, x_out.shape[0], pdl)
        self._launch(native.lib().b200_hop_recv, "b200_hop_recv", a)

    # tail link: last stage -> stage 0.  `x` may be None (token ids only, in `extra`).
    def send_tail(self, x, extra: torch.Tensor, n_extra: int, T: int, dst: int, pdl: bool = True) -> None:
        a = self._args(x, extra, n_extra, self.base[dst] + self.ring_bytes, 2, T, pdl)
        self._launch(native.lib().b200_hop_send, "b200_hop_send", a)

    def recv_tail(self, x_out, extra_out: torch.Tensor, n_extra: int, T: int, pdl: bool = True) -> None:
        a = self._args(x_out, extra_out, n_extra, self.base[self.rank] + self.ring_bytes, 3, T, pdl)
        self._launch(native.lib().b200_hop_recv, "b200_hop_recv", a)

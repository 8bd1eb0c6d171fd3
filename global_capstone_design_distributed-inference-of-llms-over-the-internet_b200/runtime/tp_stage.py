"""A served stage that is wider than one GPU: the span runs tensor-parallel over the ranks of one process group.

Upstream Petals lets one server shard its blocks over several local GPUs (``tensor_parallel_devices``,
petals/server/server.py:89,182-187; per-device KV shards petals/server/backend.py:43,88-99) inside ONE process with the
external ``tensor_parallel`` package and an NCCL all-reduce after every row-parallel linear.  Here the stage is one
process per GPU (``torchrun --nproc-per-node k -m src.main --stage s --tp k``):

* rank 0 of the group is the *leader*: it runs the DHT node, the RPC server and the connection handler exactly like a
  one-GPU stage - its stage object is a :class:`TPStage`, whose ``forward`` / ``decode_batch`` first mirror the call
  (a small command + the hidden states, broadcast over NCCL) to the followers and then run the local shard;
* ranks 1..k-1 are *followers* (:func:`serve_follower`): they hold the other shards of the weights and of every
  session's KV cache (split by KV head, nothing to communicate) and repeat the leader's calls.

The shards meet inside the kernels: ``b200_gemm_allreduce`` exchanges every finished output tile of the row-parallel
linears through peer-mapped buffers from inside the tcgen05 GEMM (parallel/tensor_parallel.py), so every rank ends each
step with bit-identical hidden states and only the leader talks to the network.
"""
from __future__ import annotations

import logging
from typing import Optional, Sequence

import torch
import torch.distributed as dist

from ..parallel.tensor_parallel import make_tp_span
from .kv_cache import KVCacheManager, SessionKV
from .stages import StageWeights, _StageBase

logger = logging.getLogger(__name__)

CMD_FORWARD, CMD_BATCH, CMD_CLOSE, CMD_STOP = 1, 2, 3, 4


class TPStage(_StageBase):
    """Leader-side stage object of a tensor-parallel span (same ``forward`` contract as Stage*/StageLast)."""

    def __init__(self, full: StageWeights, start: int, end: int, group, *, final: bool, tp_mode: str = "fused",
                 gpu_device=None, kv_capacity: int = 2048, max_batch: int = 8, keep_layers_on_gpu: int = 0):
        torch.nn.Module.__init__(self)
        if (start, end) != (full.start, full.end):
            raise ValueError(f"stage span [{start},{end}) does not match loaded weights [{full.start},{full.end})")
        self.cfg, self.start, self.end = full.cfg, start, end
        self.config = full.cfg
        self.device = torch.device(gpu_device) if gpu_device is not None else full.device
        self.dtype = full.dtype
        self.kv_capacity = kv_capacity
        self.use_cpu_offload = False
        self.keep_layers_on_gpu = keep_layers_on_gpu
        self.role = "last" if final else "segment"
        self.backend = "native"
        self.group = group
        self.tp = dist.get_world_size(group)
        self.tp_rank = dist.get_rank(group)
        self.leader = dist.get_global_rank(group, 0)
        self.span = make_tp_span(self.cfg, full.weights, self.device, self.dtype, group, mode=tp_mode,
                                 kv_capacity=kv_capacity, max_batch=max_batch)
        self._followers_kv: Optional[KVCacheManager] = None
        self._hdr = torch.zeros(8, dtype=torch.int64, device=self.device)

    # each rank caches its own KV heads
    def new_kv_manager(self, **kw) -> KVCacheManager:
        return KVCacheManager(self.end - self.start, self.cfg.n_kv_heads // self.tp, self.cfg.head_dim, self.dtype,
                              self.device, default_max_length=self.kv_capacity, **kw)

    # ------------------------------------------------------------------ mirroring
    def _send_cmd(self, cmd: int, sids: Sequence[str], T: int, pos0: int, x: Optional[torch.Tensor]) -> None:
        """Leader: header [cmd, n_sessions, T, pos0, name bytes], session names, then the hidden states."""
        names = "\n".join(sids).encode()
        self._hdr[:5] = torch.tensor([cmd, len(sids), T, pos0, len(names)], dtype=torch.int64)
        dist.broadcast(self._hdr, self.leader, group=self.group)
        if names:
            nb = torch.frombuffer(bytearray(names), dtype=torch.uint8).to(self.device)
            dist.broadcast(nb, self.leader, group=self.group)
        if x is not None:
            dist.broadcast(x, self.leader, group=self.group)

    def _local_forward(self, h: torch.Tensor, kv: SessionKV, pos0: int) -> torch.Tensor:
        if pos0 != kv.length:
            kv.rollback(pos0)
        if self.role == "last":
            if h.shape[0] == 1:
                logits, _ = self.span.decode([kv], h_in=h, greedy=True)
            else:
                logits, _ = self.span.prefill(kv, h_in=h, pos0=pos0)
            return logits[:1].clone()
        if h.shape[0] == 1:
            return self.span.decode([kv], h_in=h).clone()
        return self.span.prefill(kv, h_in=h, pos0=pos0)

    @torch.inference_mode()
    def forward(self, hidden_states, position_ids=None, attention_mask=None, past_key_values: SessionKV = None,
                use_cache: bool = True):
        kv = past_key_values
        h = hidden_states.reshape(-1, self.cfg.hidden).to(self.device, self.dtype).contiguous()
        pos0 = self._pos0(position_ids, kv)
        self._send_cmd(CMD_FORWARD, [kv.session_id], h.shape[0], pos0, h)
        out = self._local_forward(h, kv, pos0)
        return out.unsqueeze(0), kv

    @torch.inference_mode()
    def decode_batch(self, hidden: torch.Tensor, kvs, to_cpu: bool = False) -> list:
        B = hidden.shape[0]
        self.last_greedy_ids = None
        h = hidden.reshape(B, self.cfg.hidden).to(self.device, self.dtype).contiguous()
        if B > self.span.max_batch:
            return [self.forward(h[b:b + 1], None, None, kvs[b])[0] for b in range(B)]
        self._send_cmd(CMD_BATCH, [kv.session_id for kv in kvs], B, 0, h)
        if self.role == "last":
            logits, ids = self.span.decode(list(kvs), h_in=h, greedy=True)
            out = logits[:B].clone()
            self.last_greedy_ids = [int(t) for t in ids[:B].tolist()]
        else:
            out = self.span.decode(list(kvs), h_in=h).clone()
            if to_cpu:
                out = out.cpu()
        return [out[b].reshape(1, 1, -1) for b in range(B)]

    def close_session(self, session_id: str) -> None:
        """Leader: free the followers' shards of a session (the leader's own are the handler's business)."""
        self._send_cmd(CMD_CLOSE, [session_id], 0, 0, None)

    def stop_followers(self) -> None:
        self._send_cmd(CMD_STOP, [], 0, 0, None)


def serve_follower(stage: TPStage) -> None:
    """Ranks 1..k-1 of a tensor-parallel stage: repeat every call the leader makes, on this rank's shards."""
    mgr = stage.new_kv_manager()
    hdr = torch.zeros(8, dtype=torch.int64, device=stage.device)
    H = stage.cfg.hidden
    logger.info("tensor-parallel follower %d/%d of blocks [%d:%d) ready", stage.tp_rank, stage.tp, stage.start, stage.end)
    with torch.inference_mode():
        while True:
            dist.broadcast(hdr, stage.leader, group=stage.group)
            cmd, n, T, pos0, name_len = (int(v) for v in hdr[:5].tolist())
            sids = []
            if name_len:
                nb = torch.zeros(name_len, dtype=torch.uint8, device=stage.device)
                dist.broadcast(nb, stage.leader, group=stage.group)
                sids = bytes(nb.cpu().numpy()).decode().split("\n")
            if cmd == CMD_STOP:
                return
            if cmd == CMD_CLOSE:
                for s in sids:
                    mgr.close(s)
                continue
            x = torch.empty(T, H, dtype=stage.dtype, device=stage.device)
            dist.broadcast(x, stage.leader, group=stage.group)
            if cmd == CMD_FORWARD:
                kv = mgr.get(sids[0])
                if kv is None or pos0 == 0 and T > 1:
                    kv = mgr.allocate(sids[0], stage.kv_capacity)
                stage._local_forward(x, kv, pos0)
            elif cmd == CMD_BATCH:
                kvs = [mgr.get(s) for s in sids]
                stage.span.decode(kvs, h_in=x, greedy=True) if stage.role == "last" else stage.span.decode(kvs, h_in=x)

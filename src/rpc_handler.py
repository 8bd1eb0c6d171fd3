"""Server side of the data plane: sessions, KV, forward, sampling (reference: src/rpc_handler.py:43-464).

``StageConnectionHandler`` serves one stage model.  A request carries hidden states
``[1, T, H]`` plus msgpack metadata:

    session_id, seq_len, cur_len, is_prefill, max_length, temperature, top_p, top_k,
    generated_tokens?, is_replay?, repetition_penalty?, start_from_position?, route?

Behaviour kept from the reference: prefill starts a fresh cache (replay prefill clears an
existing one); decode needs an existing session unless ``is_replay``; intermediate stages
answer with hidden states, the final stage with the sampled token id (int64 ``[[id]]`` +
``token_id`` in metadata).

Behaviour fixed: the session cache is pre-allocated to ``max_length`` and the
limit is enforced (B7); a decode step whose ``cur_len`` disagrees with the cached length is an
error instead of a warning — unless ``start_from_position`` asks for a rollback (B4); sessions
can be closed and idle ones expire; no per-hop ``.item()`` range checks; compute runs on a
priority task pool thread, never on the network loop (B11).  Extras from the upstream server
(petals/server/handler.py): ``rpc_info``, ``rpc_close_session`` and server->server forwarding
of the result along ``route`` (the rpc_push idea) so a chain of hosts needs one client round trip.
"""
from __future__ import annotations

import logging
import threading
import time
from typing import Dict, List, Optional, Tuple

import torch

from petals_b200.net.rpc import RpcClient, RpcServer
from petals_b200.runtime.kv_cache import AllocationFailed, KVCacheManager
from petals_b200.runtime.sampler import SamplingParams, sample_token
from petals_b200.runtime.task_pool import PrioritizedTaskPool

logger = logging.getLogger(__name__)

HANDLER_PREFIX = "StageConnectionHandler"


class StageConnectionHandler:
    def __init__(self, dht, stage_model, device, request_timeout: float = 30.0, final_stage: bool = True,
                 attn_cache_tokens: Optional[int] = None, session_ttl: float = 600.0, start_block: Optional[int] = None,
                 end_block: Optional[int] = None, throughput: float = 0.0, seed: Optional[int] = None):
        self.dht = dht
        self.stage_model = stage_model
        self.device = torch.device(device)
        self.request_timeout = request_timeout
        self.final_stage = final_stage
        self.start_block = getattr(stage_model, "start", start_block)
        self.end_block = getattr(stage_model, "end", end_block)
        self.throughput = throughput
        cap = getattr(stage_model, "kv_capacity", 2048)
        self.kv = stage_model.new_kv_manager(max_tokens=attn_cache_tokens, session_ttl=session_ttl)
        self.kv_capacity = cap
        self._default_temperature, self._default_top_p, self._default_top_k = 0.8, 0.9, 0
        self.pool = PrioritizedTaskPool("stage")
        self._fwd_client = RpcClient()
        self._gen = torch.Generator(device="cpu")
        if seed is not None:
            self._gen.manual_seed(seed)
        self.requests = 0
        self.started = time.time()
        self.fail_next = 0   # fault injection (tests): raise on the next N forwards
        self.die_after: Optional[int] = None   # fault injection: "crash" while serving request number N
        self.on_die = None                     # callable that tears the server down (set by the launcher)
        self.die_guard = None                  # optional callable -> bool, consulted right before crashing

    # ------------------------------------------------------------------ registration
    def add_p2p_handlers(self, server: RpcServer) -> None:
        for name in ("rpc_forward", "rpc_forward_stream", "rpc_info", "rpc_close_session", "rpc_push", "rpc_check"):
            server.add_handler(f"{HANDLER_PREFIX}.{name}", getattr(self, name))

    # ------------------------------------------------------------------ core
    def _session_for(self, md: dict, T: int):
        sid = md.get("session_id")
        if sid is None:
            raise ValueError("request.metadata must contain session_id")
        is_replay = bool(md.get("is_replay", False))
        is_prefill = bool(md.get("is_prefill", False))
        cur_len = int(md.get("cur_len", md.get("seq_len", T)))
        max_length = int(md.get("max_length", self.kv_capacity))
        if max_length > self.kv_capacity:
            raise ValueError(f"max_length {max_length} exceeds this server's per-session capacity {self.kv_capacity}")
        kv = self.kv.get(sid)
        if is_prefill:
            if is_replay:
                logger.info("[%s] REPLAY MODE: restoring KV cache (prefill)", str(sid)[:8])
            kv = self.kv.allocate(sid, self.kv_capacity)
            kv.limit = max_length
        elif kv is None:
            if not is_replay:
                raise ValueError(f"Missing past_key_values for session_id={sid}: server restart or cache loss; "
                                 "replay the session with is_replay=True")
            logger.warning("[%s] REPLAY: no cache for a decode step, starting a fresh session", str(sid)[:8])
            kv = self.kv.allocate(sid, self.kv_capacity)
            kv.limit = max_length
        start = md.get("start_from_position")
        if start is not None:
            kv.rollback(int(start))
        expected = cur_len - T
        if not is_prefill and kv.length != expected:
            raise ValueError(f"KV length mismatch for session {str(sid)[:8]}: cached {kv.length} tokens, request "
                             f"implies {expected} (cur_len={cur_len}, T={T})")
        limit = getattr(kv, "limit", self.kv_capacity)
        if kv.length + T > limit:
            raise ValueError(f"session {str(sid)[:8]} would exceed max_length={limit} ({kv.length}+{T})")
        return sid, kv

    def _run_forward(self, hidden_states: torch.Tensor, metadata: Dict) -> Tuple[List[torch.Tensor], Dict]:
        if self.fail_next > 0:
            self.fail_next -= 1
            raise RuntimeError("injected failure")
        if self.die_after is not None and self.requests + 1 >= self.die_after and \
                (self.die_guard is None or self.die_guard()):
            self.die_after = None
            if self.on_die is not None:
                threading.Thread(target=self.on_die, daemon=True).start()
            raise ConnectionError("injected crash")
        T = hidden_states.shape[-2] if hidden_states.dim() >= 2 else 1
        sid, kv = self._session_for(metadata, T)
        pos_ids = torch.arange(kv.length, kv.length + T, dtype=torch.long).unsqueeze(0)
        out, _ = self.stage_model(hidden_states, position_ids=pos_ids, attention_mask=None, past_key_values=kv,
                                  use_cache=True)
        self.requests += 1
        if self.final_stage:
            logits = out.reshape(-1, out.shape[-1])[-1]
            params = SamplingParams(
                float(metadata.get("temperature", self._default_temperature)),
                float(metadata.get("top_p", self._default_top_p)), int(metadata.get("top_k", self._default_top_k)),
                float(metadata.get("repetition_penalty", 1.5)))
            generated = metadata.get("generated_tokens") or []
            if logits.is_cuda and params.temperature > 0.0 and logits.dtype == torch.float32:
                token = self._device_sample(logits, params, generated)          # csrc/sampling.cu
            else:
                token = sample_token(logits, params, generated,
                                     generator=self._gen if logits.device.type == "cpu" else None)
            return [torch.tensor([[token]], dtype=torch.long)], {"token_id": token, "session_id": sid}
        return [out.detach().to("cpu")], {"session_id": sid}

    def _device_sample(self, logits: torch.Tensor, params: SamplingParams, generated) -> int:
        from petals_b200.ops.sampling import DeviceSampler

        if getattr(self, "_dsampler", None) is None or self._dsampler.vocab != logits.shape[-1]:
            self._dsampler = DeviceSampler(logits.device, 1, logits.shape[-1])
        ids = self._dsampler.sample(logits.reshape(1, -1).contiguous(), params, [list(generated)])
        return int(ids[0].item())

    def _submit(self, fn, *args, kind="inference", tokens=0):
        return self.pool.submit(fn, *args, kind=kind, tokens=tokens).result(timeout=self.request_timeout)

    # ------------------------------------------------------------------ RPC methods
    def rpc_forward(self, uid: str, tensors: List[torch.Tensor], metadata: Dict):
        if not tensors:
            raise ValueError("No tensors in request")
        hidden = tensors[0]
        T = hidden.shape[-2] if hidden.dim() >= 2 else 1
        out_t, out_md = self._submit(self._run_forward, hidden, metadata, kind="inference", tokens=T)
        route = metadata.get("route") or []
        if route and not self.final_stage:
            # server->server push: hand the activations to the next hop ourselves and relay its answer
            (host, port, next_uid), rest = route[0], route[1:]
            md = dict(metadata)
            md["route"] = rest
            return self._fwd_client.call((host, int(port)), f"{HANDLER_PREFIX}.rpc_forward", next_uid, out_t, md,
                                         timeout=self.request_timeout)
        return out_t, out_md

    rpc_forward_stream = rpc_forward   # TCP needs no separate streaming path; kept for name parity
    rpc_push = rpc_forward

    def rpc_info(self, uid: str, tensors, metadata: Dict):
        left = self.kv.tokens_left
        return [], {"version": "0.1.0", "start_block": self.start_block, "end_block": self.end_block,
                    "final_stage": self.final_stage, "sessions": len(self.kv.sessions),
                    "cache_tokens_left": left if left is not None else -1, "kv_capacity": self.kv_capacity,
                    "throughput": self.throughput, "requests": self.requests, "queued": self.pool.queued,
                    "uptime_s": time.time() - self.started}

    def rpc_check(self, uid: str, tensors, metadata: Dict):
        """Dial-back probe for joining servers (upstream reachability.py rpc_check)."""
        from petals_b200.net.reachability import check_direct_reachability

        ok, rtt = check_direct_reachability(metadata["maddr"], float(metadata.get("timeout", 2.0)))
        return [], {"reachable": ok, "rtt_s": rtt}

    def rpc_close_session(self, uid: str, tensors, metadata: Dict):
        return [], {"closed": self.kv.close(metadata.get("session_id", ""))}

    def shutdown(self) -> None:
        self.pool.shutdown()
        self._fwd_client.close()

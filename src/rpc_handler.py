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
of the result along ``route`` (the rpc_push idea) so a chain of hosts needs one client round trip;
``rpc_inference`` - one stream per session with the cache allocated once, ``session_timeout`` /
``step_timeout`` and the cache freed when the stream ends (handler.py:132-195, 532-547);
cache admission that waits up to ``alloc_timeout`` for bytes (memory_cache.py:26-225);
``hypo_ids`` beam re-ordering (backend.py:154-158).

Concurrent sessions: single-token decode requests of different sessions that are queued at the
same time are merged into ONE batched step of the stage (up to ``max_batch`` sessions,
task_pool.py:17-177 / backend.py:201-235 upstream) - the decode kernels serve 16 sessions in about
1.4x the time of one.
"""
from __future__ import annotations

import logging
import os
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
                 end_block: Optional[int] = None, throughput: float = 0.0, seed: Optional[int] = None,
                 alloc_timeout: float = 10.0, session_timeout: float = 30 * 60.0, step_timeout: float = 5 * 60.0,
                 max_batch: Optional[int] = None, batch_linger: float = 0.002):
        self.dht = dht
        self.stage_model = stage_model
        self.device = torch.device(device)
        self.request_timeout = request_timeout
        self.final_stage = final_stage
        self.start_block = getattr(stage_model, "start", start_block)
        self.end_block = getattr(stage_model, "end", end_block)
        self.throughput = throughput
        cap = getattr(stage_model, "kv_capacity", 2048)
        self._kv_kw = dict(max_tokens=attn_cache_tokens, session_ttl=session_ttl, alloc_timeout=alloc_timeout)
        self.kv = stage_model.new_kv_manager(**self._kv_kw)
        self._views: Dict[tuple, tuple] = {}
        self._views_lock = threading.Lock()
        self.kv_capacity = cap
        self.session_timeout, self.step_timeout = session_timeout, step_timeout
        # sessions per batched decode step: what the stage's engine takes in one call (1 = never batch)
        span = getattr(stage_model, "span", None)
        self.max_batch = int(max_batch if max_batch is not None else getattr(span, "max_batch", 1) or 1)
        if not hasattr(stage_model, "decode_batch"):
            self.max_batch = 1
        self.streams_open = 0
        self._default_temperature, self._default_top_p, self._default_top_k = 0.8, 0.9, 0
        # sessions that made a decode step in the last 50 ms are expected to make the next one together: the pool
        # waits up to batch_linger for them before it runs a batched step (0 = never wait)
        self._stepping: Dict[str, float] = {}
        self.batch_seconds: Dict[str, float] = {}
        # request-level timing (rpc_info): time inside the server per request, and how long a session takes to come
        # back with its next step after a reply (network + client)
        self._last_reply: Dict[str, float] = {}
        self.think_seconds = self.service_seconds = 0.0
        self.think_count = self.service_count = 0
        self.batch_linger = float(batch_linger)
        self.pool = PrioritizedTaskPool("stage", batch_linger=self.batch_linger if self.max_batch > 1 else 0.0,
                                        expect=self._expected_batch, thread_init=self._bind_device,
                                        linger_max_missing=int(os.environ.get("B200_LINGER_MAX_MISSING", "32")))
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
        # device-to-device hand-off inside one box (petals_b200/net/devshare.py): built on first use
        self._exporter = None
        self._importer = None
        self._box = None
        self.handoff_out = self.handoff_in = 0

    def _view(self, md: dict):
        """(stage, cache manager, is_final) a request runs on: the whole span, or - ``blocks = [b, e)`` in the request,
        a client entering this server's span at a block other than its first - a view over those blocks with its own
        per-session caches (same weights; runtime/stages.py:subspan)."""
        blocks = md.get("blocks")
        if not blocks or (int(blocks[0]), int(blocks[1])) == (self.start_block, self.end_block):
            return self.stage_model, self.kv, self.final_stage
        key = (int(blocks[0]), int(blocks[1]))
        with self._views_lock:
            v = self._views.get(key)
            if v is None:
                if not hasattr(self.stage_model, "subspan"):
                    raise ValueError("this server only serves its whole span")
                stage = self.stage_model.subspan(*key)
                v = self._views[key] = (stage, stage.new_kv_manager(**self._kv_kw),
                                        self.final_stage and key[1] == self.end_block)
                logger.info("serving a sub-span: blocks [%d:%d) of [%s:%s)", key[0], key[1], self.start_block,
                            self.end_block)
        return v

    def _expected_batch(self) -> int:
        now = time.monotonic()
        if len(self._stepping) > 4 * self.max_batch:       # sessions that ended without a close (TTL eviction)
            for sid in [s for s, t in list(self._stepping.items()) if now - t > 60.0]:
                self._stepping.pop(sid, None)
                self._last_reply.pop(sid, None)
        return sum(1 for t in list(self._stepping.values()) if now - t < 0.05)

    # ------------------------------------------------------------------ registration
    def add_p2p_handlers(self, server: RpcServer) -> None:
        for name in ("rpc_forward", "rpc_forward_stream", "rpc_info", "rpc_close_session", "rpc_push", "rpc_check",
                     "rpc_reorder", "rpc_backward", "rpc_backward_stream"):
            server.add_handler(f"{HANDLER_PREFIX}.{name}", getattr(self, name))
        server.add_stream_handler(f"{HANDLER_PREFIX}.rpc_inference", self.rpc_inference)

    # ------------------------------------------------------------------ core
    def _session_for(self, md: dict, T: int, mgr=None):
        mgr = mgr or self.kv
        sid = md.get("session_id")
        if sid is None:
            raise ValueError("request.metadata must contain session_id")
        is_replay = bool(md.get("is_replay", False))
        is_prefill = bool(md.get("is_prefill", False))
        cur_len = int(md.get("cur_len", md.get("seq_len", T)))
        max_length = int(md.get("max_length", self.kv_capacity))
        if max_length > self.kv_capacity:
            raise ValueError(f"max_length {max_length} exceeds this server's per-session capacity {self.kv_capacity}")
        kv = mgr.get(sid)
        if is_prefill:
            if is_replay:
                logger.info("[%s] REPLAY MODE: restoring KV cache (prefill)", str(sid)[:8])
            kv = mgr.allocate(sid, self.kv_capacity)
            kv.limit = max_length
        elif kv is None:
            if not is_replay:
                raise ValueError(f"Missing past_key_values for session_id={sid}: server restart or cache loss; "
                                 "replay the session with is_replay=True")
            logger.warning("[%s] REPLAY: no cache for a decode step, starting a fresh session", str(sid)[:8])
            kv = mgr.allocate(sid, self.kv_capacity)
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

    def _bind_device(self) -> None:
        """CUDA's current device is per host thread and starts at 0: the compute thread of a server on another GPU
        must select its own before it launches anything (kernels take the current stream of the current device)."""
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)

    def _run_forward(self, hidden_states: torch.Tensor, metadata: Dict) -> Tuple[List[torch.Tensor], Dict]:
        self._bind_device()
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
        self._maybe_reorder(metadata)
        stage, mgr, final = self._view(metadata)
        sid, kv = self._session_for(metadata, T, mgr)
        pos_ids = torch.arange(kv.length, kv.length + T, dtype=torch.long).unsqueeze(0)
        out, _ = stage(hidden_states, position_ids=pos_ids, attention_mask=None, past_key_values=kv, use_cache=True)
        self.requests += 1
        return self._finish(out, metadata, sid, final)

    def box(self) -> Optional[str]:
        """The machine id this server hands device tensors off in (None on CPU)."""
        if self._box is None and self.device.type == "cuda":
            from petals_b200.net.devshare import box_id
            self._box = box_id()
        return self._box

    def _gpu_ids(self):
        """(physical index of this server's GPU, physical indices of the GPUs it can address) - what a client needs
        to know before it offers device hand-off."""
        if self.device.type != "cuda":
            return None, []
        from petals_b200.net.devshare import _physical_index, visible_gpus
        return _physical_index(self.device), visible_gpus()

    def _export(self, out: torch.Tensor) -> dict:
        if self._exporter is None:
            from petals_b200.net.devshare import DeviceExporter
            self._exporter = DeviceExporter(self.device)
        self.handoff_out += 1
        return self._exporter.export(out)

    def _import(self, desc: dict) -> torch.Tensor:
        if self._importer is None:
            from petals_b200.net.devshare import DeviceImporter
            self._importer = DeviceImporter(self.device)
        self.handoff_in += 1
        return self._importer.pull(desc)

    def _finish(self, out: torch.Tensor, metadata: Dict, sid, final: Optional[bool] = None) -> Tuple[List[torch.Tensor], Dict]:
        """Stage output of ONE request -> its reply (sampled token on the final stage, hidden states otherwise)."""
        if self.final_stage if final is None else final:
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
            reply = {"token_id": token, "session_id": sid}
            k = int(metadata.get("return_logprobs") or 0)
            if k > 0:
                # beam search on the client: the k most likely next tokens with their log-probabilities
                lp = torch.log_softmax(logits.float(), dim=-1)
                top = torch.topk(lp, min(k, lp.numel()))
                reply["topk_ids"] = [int(i) for i in top.indices.tolist()]
                reply["topk_logprobs"] = [float(v) for v in top.values.tolist()]
            return [torch.tensor([[token]], dtype=torch.long)], reply
        if out.is_cuda and metadata.get("device_handoff") and metadata["device_handoff"] == self.box():
            # the consumer shares this box: the reply names a slot of this GPU's exported arena, no payload
            return [], {"session_id": sid, "device_tensor": self._export(out)}
        return [out.detach() if out.device.type == "cpu" else out.detach().to("cpu")], {"session_id": sid}

    def _maybe_reorder(self, metadata: Dict) -> None:
        """Beam search: ``hypo_ids[i]`` names which hypothesis of ``beam_sessions`` session i continues."""
        hypo = metadata.get("hypo_ids")
        if hypo is None:
            return
        group = metadata.get("beam_sessions")
        if not group:
            raise ValueError("hypo_ids needs beam_sessions (the session ids of the beam, in hypothesis order)")
        self.kv.reorder(list(group), [int(h) for h in hypo])

    def _run_decode_batch(self, items: List[Tuple[torch.Tensor, Dict]]) -> list:
        """One batched decode step for every queued single-token request (different sessions).  Returns one
        reply - or one exception - per request, so a bad request does not fail its neighbours."""
        self._bind_device()
        results: list = [None] * len(items)
        t0 = time.perf_counter()
        ready = []
        seen = set()
        for i, (hidden, md) in enumerate(items):
            try:
                if self.fail_next > 0:
                    self.fail_next -= 1
                    raise RuntimeError("injected failure")
                if self.die_after is not None and self.requests + len(ready) + 1 >= self.die_after and \
                        (self.die_guard is None or self.die_guard()):
                    self.die_after = None
                    if self.on_die is not None:
                        threading.Thread(target=self.on_die, daemon=True).start()
                    raise ConnectionError("injected crash")
                self._maybe_reorder(md)
                sid, kv = self._session_for(md, 1)
                if sid in seen:                       # two steps of ONE session cannot share a batch
                    results[i] = self._run_forward(hidden, md)
                    continue
                seen.add(sid)
                ready.append((i, hidden, md, sid, kv))
            except BaseException as e:  # noqa: BLE001
                results[i] = e
        if ready:
            try:
                H = ready[0][1].shape[-1]
                h = torch.cat([r[1].reshape(1, H) for r in ready], 0)
                t1 = time.perf_counter()
                to_cpu = not self.final_stage and not any(r[2].get("device_handoff") for r in ready)
                outs = self.stage_model.decode_batch(h, [r[4] for r in ready], to_cpu=to_cpu)
                ids = getattr(self.stage_model, "last_greedy_ids", None)
                t2 = time.perf_counter()
                self.requests += len(ready)
                for b, ((i, _, md, sid, _), o) in enumerate(zip(ready, outs)):
                    if ids is not None and float(md.get("temperature", self._default_temperature)) <= 0.0 \
                            and not md.get("return_logprobs"):
                        results[i] = ([torch.tensor([[ids[b]]], dtype=torch.long)], {"token_id": ids[b], "session_id": sid})
                    else:
                        results[i] = self._finish(o, md, sid)
                t3 = time.perf_counter()
                # where a batched step's wall time goes (rpc_info: batch_ms) - prepare / engine step / replies
                for k, v in (("prepare", t1 - t0), ("step", t2 - t1), ("finish", t3 - t2)):
                    self.batch_seconds[k] = self.batch_seconds.get(k, 0.0) + v
            except BaseException as e:  # noqa: BLE001
                for i, *_ in ready:
                    results[i] = e
        return results

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
        desc = metadata.get("device_tensor")
        if desc is not None:
            hidden = self._import(desc)              # one peer copy over NVLink from the producer's arena
            metadata = {k: v for k, v in metadata.items() if k != "device_tensor"}
        elif not tensors:
            raise ValueError("No tensors in request")
        else:
            hidden = tensors[0]
        if metadata.get("stateless") or (metadata.get("session_id") is None and "cur_len" not in metadata
                                         and "is_prefill" not in metadata):
            # no session: upstream's rpc_forward - whole sequences through the span, nothing cached
            return self._stateless(hidden, None, metadata, tensors[1] if len(tensors) > 1 else None)
        T = hidden.shape[-2] if hidden.dim() >= 2 else 1
        t_in = time.perf_counter()
        sid_ = metadata.get("session_id")
        last = self._last_reply.get(sid_)
        if last is not None and T == 1:
            self.think_seconds += t_in - last          # reply sent -> next step of the session received
            self.think_count += 1
        whole = self._view(metadata)[0] is self.stage_model        # sub-span views run unbatched
        if whole and T == 1 and self.max_batch > 1 and not metadata.get("is_prefill") and hidden.numel() == hidden.shape[-1]:
            self._stepping[metadata.get("session_id")] = time.monotonic()
            out_t, out_md = self.pool.submit_batch_item(self._run_decode_batch, (hidden, metadata), kind="inference",
                                                        tokens=1, max_batch=self.max_batch).result(
                                                            timeout=self.request_timeout)
        else:
            out_t, out_md = self._submit(self._run_forward, hidden, metadata, kind="inference", tokens=T)
        t_out = time.perf_counter()
        if T == 1:                                      # (decode steps only: a prompt is followed by client-side work)
            self._last_reply[sid_] = t_out
            self.service_seconds += t_out - t_in
            self.service_count += 1
        route = metadata.get("route") or []
        if route and "token_id" not in out_md:
            # server->server push: hand the activations to the next hop ourselves and relay its answer
            (host, port, next_uid), rest = route[0], route[1:]
            md = dict(metadata)
            md["route"] = rest
            if "device_tensor" in out_md:
                md["device_tensor"] = out_md["device_tensor"]
            return self._fwd_client.call((host, int(port)), f"{HANDLER_PREFIX}.rpc_forward", next_uid, out_t, md,
                                         timeout=self.request_timeout)
        return out_t, out_md

    def _stateless(self, hidden: torch.Tensor, grad_out: Optional[torch.Tensor], metadata: Dict,
                   prompts: Optional[torch.Tensor] = None):
        stage = self._view(metadata)[0]
        if not hasattr(stage, "forward_stateless"):
            raise ValueError("this server only serves inference sessions")
        tokens = int(hidden.numel() // hidden.shape[-1])

        def run():
            self._bind_device()
            if grad_out is None:
                return stage.forward_stateless(hidden, prompts) if prompts is not None else stage.forward_stateless(hidden)
            return stage.backward(hidden, grad_out, prompts) if prompts is not None else stage.backward(hidden, grad_out)

        out = self._submit(run, kind="forward" if grad_out is None else "backward", tokens=tokens)
        self.requests += 1
        outs = list(out) if isinstance(out, tuple) else [out]
        return [o.detach().to("cpu") for o in outs], {"stateless": True}

    def rpc_backward(self, uid: str, tensors: List[torch.Tensor], metadata: Dict):
        """Upstream ``rpc_backward`` (petals/server/handler.py, block_functions.run_rpc_backward): ``tensors`` =
        [inputs, grad_outputs] of this span, both [B, T, H] (+ optional deep prompts [blocks, B, prefix, H]); the answer is
        the gradient w.r.t. the inputs (+ the gradient w.r.t. the prompts).  The
        forward is recomputed here (nothing is kept between the two calls), at forward priority - inference steps of
        other sessions go first."""
        if len(tensors) < 2:
            raise ValueError("rpc_backward needs [inputs, grad_outputs]")
        if tensors[0].shape != tensors[1].shape:
            raise ValueError(f"inputs {tuple(tensors[0].shape)} and grad_outputs {tuple(tensors[1].shape)} differ in shape")
        return self._stateless(tensors[0], tensors[1], metadata, tensors[2] if len(tensors) > 2 else None)

    rpc_backward_stream = rpc_backward
    rpc_forward_stream = rpc_forward   # TCP needs no separate streaming path; kept for name parity
    rpc_push = rpc_forward

    def rpc_reorder(self, uid: str, tensors, metadata: Dict):
        """Beam search between steps: session ``session_ids[i]`` continues hypothesis ``hypo_ids[i]``."""
        self._submit(self.kv.reorder, list(metadata["session_ids"]), [int(h) for h in metadata["hypo_ids"]],
                     kind="inference")
        return [], {"reordered": len(metadata["session_ids"])}

    def rpc_inference(self, uid: str, tensors: List[torch.Tensor], metadata: Dict, stream):
        """One inference SESSION over one connection (upstream handler.py:132-195).

        The opening message names ``session_id`` and ``max_length``; the cache is allocated once, right here,
        waiting up to ``alloc_timeout`` for bytes (AllocationFailed goes back to the client).  Every further
        message is one step - a prompt chunk or a single token - answered in order; the session ends when the
        client says ``close``, when no step arrives for ``step_timeout`` seconds, after ``session_timeout``
        seconds in total, or when the connection drops - and the cache is freed in every one of those cases."""
        sid = metadata.get("session_id")
        if sid is None:
            raise ValueError("rpc_inference: the opening message must carry session_id")
        max_length = int(metadata.get("max_length", self.kv_capacity))
        if max_length > self.kv_capacity:
            raise ValueError(f"max_length {max_length} exceeds this server's per-session capacity {self.kv_capacity}")
        mgr = self._view(metadata)[1]
        kv = mgr.allocate(sid, self.kv_capacity, timeout=metadata.get("alloc_timeout"))
        kv.limit = max_length
        self.streams_open += 1
        deadline = time.monotonic() + float(metadata.get("session_timeout", self.session_timeout))
        step_timeout = float(metadata.get("step_timeout", self.step_timeout))
        stream.send([], {"session_id": sid, "opened": True, "max_length": max_length})
        steps, reason = 0, "closed by client"
        try:
            while True:
                left = deadline - time.monotonic()
                if left <= 0:
                    reason = "session_timeout"
                    break
                try:
                    t_in, md = stream.recv(min(step_timeout, left))
                except TimeoutError:
                    reason = "step_timeout" if step_timeout <= left else "session_timeout"
                    break
                except ConnectionError:
                    reason = "connection lost"
                    break
                if md.get("close"):
                    break
                md = dict(md)
                md.setdefault("session_id", sid)
                md.setdefault("max_length", max_length)
                if metadata.get("blocks"):
                    md.setdefault("blocks", metadata["blocks"])
                if md.get("is_prefill") and kv.length == 0 and steps == 0:
                    md["is_prefill"] = False          # the cache of this stream exists already: append, do not re-allocate
                    md.setdefault("start_from_position", 0)
                try:
                    out_t, out_md = self.rpc_forward(uid, t_in, md)
                    stream.send(out_t, out_md)
                except Exception as e:  # noqa: BLE001  (a failed step is reported, the session stays usable)
                    stream.send([], {"error": f"{type(e).__name__}: {e}"})
                steps += 1
        finally:
            self.streams_open -= 1
            mgr.close(sid)
            logger.info("[%s] inference stream ended after %d steps (%s)", str(sid)[:8], steps, reason)
        return [], {"session_id": sid, "steps": steps, "ended": reason}

    def rpc_info(self, uid: str, tensors, metadata: Dict):
        left = self.kv.tokens_left
        return [], {"version": "0.1.0", "start_block": self.start_block, "end_block": self.end_block,
                    "final_stage": self.final_stage, "sessions": len(self.kv.sessions),
                    "cache_tokens_left": left if left is not None else -1, "kv_capacity": self.kv_capacity,
                    "throughput": self.throughput, "requests": self.requests, "queued": self.pool.queued,
                    "max_batch": self.max_batch, "batches": self.pool.batches,
                    "mean_batch": (self.pool.batched_items / self.pool.batches) if self.pool.batches else 0.0,
                    "streams_open": self.streams_open, "uptime_s": time.time() - self.started,
                    "batch_ms": {k: round(v * 1e3 / max(self.pool.batches, 1), 3) for k, v in self.batch_seconds.items()},
                    "mean_service_ms": round(self.service_seconds * 1e3 / max(self.service_count, 1), 3),
                    "mean_turnaround_ms": round(self.think_seconds * 1e3 / max(self.think_count, 1), 3),
                    "healthy": bool(self.pool._thread.is_alive()) and not self.pool._stop,
                    "box": self.box(), "gpu": self._gpu_ids()[0], "visible_gpus": self._gpu_ids()[1], "handoff_out": self.handoff_out, "handoff_in": self.handoff_in}

    def rpc_check(self, uid: str, tensors, metadata: Dict):
        """Dial-back probe for joining servers (upstream reachability.py rpc_check)."""
        from petals_b200.net.reachability import check_direct_reachability

        ok, rtt = check_direct_reachability(metadata["maddr"], float(metadata.get("timeout", 2.0)))
        return [], {"reachable": ok, "rtt_s": rtt}

    def _close_everywhere(self, sid: str) -> bool:
        self._last_reply.pop(sid, None)
        closed = self.kv.close(sid)
        for _, mgr, _ in self._views.values():
            closed = mgr.close(sid) or closed
        return closed

    def _close(self, sid: str) -> bool:
        self._stepping.pop(sid, None)
        closed = self._close_everywhere(sid)
        mirror = getattr(self.stage_model, "close_session", None)     # tensor-parallel stage: the followers' shards
        if closed and mirror is not None:
            mirror(sid)
        return closed

    def rpc_close_session(self, uid: str, tensors, metadata: Dict):
        sid = metadata.get("session_id", "")
        if getattr(self.stage_model, "close_session", None) is not None:
            return [], {"closed": self._submit(self._close, sid, kind="inference")}    # serialised with the steps
        self._stepping.pop(sid, None)
        return [], {"closed": self._close_everywhere(sid)}

    def shutdown(self) -> None:
        stop = getattr(self.stage_model, "stop_followers", None)
        if stop is not None:
            try:
                self._submit(stop, kind="inference")
            except Exception:  # noqa: BLE001
                pass
        self.pool.shutdown()
        self._fwd_client.close()
        for side in (self._importer, self._exporter):
            if side is not None:
                try:
                    side.close()
                except Exception:  # noqa: BLE001
                    pass
        self._importer = self._exporter = None

"""Client side of the data plane: discover, route, call, journal, recover
(reference: src/rpc_transport.py:45-863).

``RpcTransport`` sends the client's hidden states through the server stages and returns the
sampled token.  Two routing modes, as in the reference:

* ``"stage"``  — fixed keys ``mini_petals:stage1..N``; any live peer under a key may serve it
  (one of the 5 freshest is picked at random).
* ``"module"`` — load-balanced: starting at ``start_block`` pick, per hop, the announced span
  that covers the current block and reaches furthest (ties: higher throughput), until the
  final stage.

Fault tolerance: every tensor sent to a hop is journaled on the CPU; when a call fails the peer
is excluded, a replacement is discovered, the journal is replayed into it to rebuild its KV
cache and the call is retried (at most 3 attempts).  Fixes over the reference:
the in-flight step is NOT part of the replay (the reference replays it and then sends it
again, so the new server sees the token twice), and in module mode the whole route suffix is
recomputed from the failed hop's first block and replayed end to end, so span boundaries stay
aligned.  With ``push=True`` servers forward activations to the next hop themselves (chain
topology; one client round trip per token instead of one per stage).
"""
from __future__ import annotations

import logging
import random
import statistics
import time
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from petals_b200.net.dht import DHT, parse_maddr
from petals_b200.net.rpc import RpcClient, RpcError
from petals_b200.utils import metrics

from .dht_utils import get_block_entries, get_stage_entries
from .rpc_handler import HANDLER_PREFIX

logger = logging.getLogger(__name__)

MAX_ATTEMPTS = 3
FRESHEST = 5


class Hop:
    """One server on the route: where it is and which journal feeds it."""

    def __init__(self, key: str, entry: dict, final: bool, start_block: Optional[int] = None,
                 end_block: Optional[int] = None):
        self.key, self.entry, self.final = key, entry, final
        self.peer_id = str(entry.get("peer_id"))
        host, port, _ = parse_maddr((entry.get("p2p_maddrs") or [""])[0])
        self.addr = (host, port)
        self.start_block, self.end_block = start_block, end_block
        self.box: Optional[str] = None        # machine id of the server, asked on first use (device hand-off)
        self.partial = False                  # module routing: the session enters this server's span mid-way


class RpcTransport:
    def __init__(self, device=None, stage: int = 0, dht_initial_peers: Optional[Sequence[str]] = None,
                 dht_port: int = 0, rpc_port: int = 0, temperature: float = 1.0, top_p: float = 0.92, top_k: int = 50,
                 stage_keys: Optional[Sequence[str]] = None, routing: str = "stage", model_name: str = "default",
                 total_blocks: int = 32, start_block: int = 0, request_timeout: float = 30.0,
                 repetition_penalty: Optional[float] = None, dht: Optional[DHT] = None, push: bool = False,
                 stream: bool = False, session_timeout: Optional[float] = None, step_timeout: Optional[float] = None,
                 device_handoff: bool = False, device_handoff_min_bytes: int = 128 << 10):
        self.device = device
        # (measured on Llama-2-13B, 4xB200: for a single token's 10 KB the arena copy, the peer copy and their
        # synchronisations cost more than the frame bytes they replace - 6.0 vs 4.7 ms per token - so only payloads
        # from device_handoff_min_bytes upwards, i.e. prompts, go device to device)
        self.device_handoff_min_bytes = int(device_handoff_min_bytes)
        # device_handoff: hops that live in this box exchange hidden states GPU to GPU (net/devshare.py); the TCP
        # messages to and from them carry a descriptor instead of the tensor
        self.device_handoff = bool(device_handoff) and device is not None and torch.device(device).type == "cuda"
        self._exporter = self._importer = None
        self._box = None
        self.handoff_hops = 0
        if self.device_handoff:
            from petals_b200.net.devshare import DeviceExporter, DeviceImporter, box_id
            self._exporter, self._importer, self._box = DeviceExporter(device), DeviceImporter(device), box_id()
        self.temperature, self.top_p, self.top_k = temperature, top_p, top_k
        self.repetition_penalty = repetition_penalty
        self.stage_keys = list(stage_keys or [])
        self.routing = routing
        self.model_name, self.total_blocks, self.start_block = model_name, total_blocks, start_block
        self.request_timeout = request_timeout
        self.push = push
        # stream=True: one ``rpc_inference`` stream per (session, server) instead of one unary call per step - the
        # server allocates the session's cache once when the stream opens and frees it when the stream ends
        self.stream = bool(stream) and not push
        self.session_timeout, self.step_timeout = session_timeout, step_timeout
        self._streams: Dict[Tuple[str, str], object] = {}
        peers = [p for p in (dht_initial_peers or []) if p]
        for p in peers:
            if ":" in p and "/p2p/" not in p and not p.startswith("/"):
                raise ValueError(f"initial peer {p!r} must be a full multiaddr (/ip4/../tcp/../p2p/..)")
        self._own_dht = dht is None
        self.dht = dht or DHT(start=True, initial_peers=peers, client_mode=True)
        self.client = RpcClient()
        self.failed_peers: Dict[str, set] = {}
        self.session_routes: Dict[str, List[Hop]] = {}
        # journal[session][hop key] = tensors sent to that hop, in order
        self.client_cache: Dict[str, Dict[str, List[torch.Tensor]]] = {}
        self.session_max_length: Dict[str, int] = {}
        self.last_token: Optional[int] = None
        self.last_reply: dict = {}            # metadata of the final stage's latest answer (token, top-k log-probs ...)
        self.extra_metadata: dict = {}        # merged into every step's request (e.g. return_logprobs for beam search)
        self.hop_times: List[Tuple[str, float]] = []
        self.last_prefill_total: Optional[float] = None
        self.last_decode_total: Optional[float] = None
        self.decode_totals: List[float] = []
        self.recoveries = 0
        logger.info("RpcTransport initialized")

    # ------------------------------------------------------------------ discovery / routing
    def _pick_stage_peer(self, key: str, exclude: set) -> Optional[dict]:
        stage = int(key.rsplit("stage", 1)[-1])
        entries = [e for e in get_stage_entries(self.dht, stage) if str(e.get("peer_id")) not in exclude
                   and e.get("p2p_maddrs")]
        if not entries:
            return None
        entries.sort(key=lambda e: e.get("timestamp", 0), reverse=True)
        return random.choice(entries[:FRESHEST])

    def _discover_stage(self, key: str, exclude: set, tries: int = 10, delay: float = 0.5) -> dict:
        for _ in range(tries):
            e = self._pick_stage_peer(key, exclude)
            if e is not None:
                return e
            time.sleep(delay)
        if exclude:
            # every announced peer has failed at least once: failures may have been transient, so
            # forgive them rather than giving up while live records are still being refreshed
            logger.warning("all peers under %s are excluded; clearing the exclusion list", key)
            exclude.clear()
            e = self._pick_stage_peer(key, exclude)
            if e is not None:
                return e
        raise ConnectionError(f"no live peer under {key} (excluded {len(exclude)})")

    def _module_route(self, from_block: int, exclude: set) -> List[Hop]:
        hops: List[Hop] = []
        b = from_block
        while b < self.total_blocks:
            cands = [e for e in get_block_entries(self.dht, b, self.model_name)
                     if str(e.get("peer_id")) not in exclude and e.get("p2p_maddrs")
                     and e.get("state", "online") != "offline"
                     and e.get("start_block") is not None and int(e["start_block"]) <= b < int(e["end_block"])]
            if not cands:
                raise ConnectionError(f"no live server holds block {b} of {self.model_name}")
            # prefer a server whose span STARTS at b (it runs its whole span: batched steps, one graph); any server
            # that covers b will do otherwise - it is asked for blocks [b, its end) only (md["blocks"], served from
            # a view over its span with caches of its own)
            aligned = [e for e in cands if int(e["start_block"]) == b]
            best = max(aligned or cands, key=lambda e: (int(e["end_block"]), float(e.get("throughput") or 0.0)))
            end = int(best["end_block"])
            hop = Hop(f"block_{b}", best, final=bool(best.get("final_stage")) or end >= self.total_blocks,
                      start_block=b, end_block=end)
            hop.partial = int(best["start_block"]) != b
            hops.append(hop)
            b = end
        return hops

    def _get_route(self, session_id: str) -> List[Hop]:
        route = self.session_routes.get(session_id)
        if route is None:
            if self.routing == "module":
                route = self._module_route(self.start_block, set())
            else:
                route = []
                for i, key in enumerate(self.stage_keys):
                    e = self._discover_stage(key, self.failed_peers.get(key, set()))
                    route.append(Hop(key, e, final=(i == len(self.stage_keys) - 1)))
            self.session_routes[session_id] = route
        return route

    # ------------------------------------------------------------------ calls
    def _metadata(self, session_id: str, T: int, cur_len: int, is_prefill: bool, generated: Optional[Sequence[int]],
                  is_replay: bool = False) -> dict:
        md = {"session_id": session_id, "seq_len": int(T), "cur_len": int(cur_len), "is_prefill": bool(is_prefill),
              "max_length": int(self.session_max_length.get(session_id, 0) or 0) or None,
              "temperature": self.temperature, "top_p": self.top_p, "top_k": self.top_k}
        if md["max_length"] is None:
            del md["max_length"]
        if generated is not None and not is_prefill:
            md["generated_tokens"] = [int(t) for t in list(generated)[-50:]]
        if is_replay:
            md["is_replay"] = True
        if self.repetition_penalty is not None:
            md["repetition_penalty"] = float(self.repetition_penalty)
        if self.extra_metadata and not is_replay:
            md.update(self.extra_metadata)
        return md

    def _in_box(self, hop: Hop) -> bool:
        """Can this hop's server and the client hand tensors over device to device: same machine, and each side can
        address the other's GPU (asked once per server)?"""
        if not self.device_handoff:
            return False
        if hop.box is None:
            try:
                from petals_b200.net.devshare import _physical_index, visible_gpus
                _, info = self.client.call(hop.addr, f"{HANDLER_PREFIX}.rpc_info", hop.key, [], {}, timeout=5.0)
                mine = _physical_index(torch.device(self.device))
                ok = (info.get("box") == self._box and mine in (info.get("visible_gpus") or [])
                      and info.get("gpu") in visible_gpus())
                hop.box = self._box if ok else ""
            except Exception:  # noqa: BLE001
                hop.box = ""
        return hop.box == self._box

    def _call(self, hop: Hop, tensor: torch.Tensor, md: dict, desc: Optional[dict] = None,
              consumer_in_box: bool = False):
        """One step on one hop.  ``desc``: the producer's hand-off descriptor of ``tensor`` (passed through untouched
        when the hop can pull it itself); ``consumer_in_box``: whoever reads this hop's output shares the box."""
        tensors = [tensor]
        if hop.partial:
            md = dict(md)
            md["blocks"] = [int(hop.start_block), int(hop.end_block)]
        big = tensor.numel() * tensor.element_size() >= self.device_handoff_min_bytes
        consumer_in_box = consumer_in_box and big          # a hop's output has its input's shape
        if tensor.is_cuda:
            if big and self._in_box(hop):
                md = dict(md)
                md["device_tensor"] = desc if desc is not None else self._exporter.export(tensor)
                tensors = []
                self.handoff_hops += 1
            else:
                tensors = [tensor.to("cpu")]
        if consumer_in_box and self._in_box(hop):
            md = dict(md)
            md["device_handoff"] = self._box
        if self.stream:
            out, omd = self._call_stream(hop, tensors, md)
        else:
            out, omd = self.client.call(hop.addr, f"{HANDLER_PREFIX}.rpc_forward", hop.key, tensors, md,
                                        timeout=self.request_timeout)
        return out, omd

    def _call_stream(self, hop: Hop, tensors: List[torch.Tensor], md: dict):
        sid = md["session_id"]
        key = (sid, hop.peer_id)
        st = self._streams.get(key)
        try:
            if st is None:
                open_md = {"session_id": sid, "max_length": md.get("max_length") or self.session_max_length.get(sid)}
                if open_md["max_length"] is None:
                    del open_md["max_length"]
                if md.get("blocks"):
                    open_md["blocks"] = md["blocks"]
                if self.session_timeout is not None:
                    open_md["session_timeout"] = self.session_timeout
                if self.step_timeout is not None:
                    open_md["step_timeout"] = self.step_timeout
                st = self.client.open_stream(hop.addr, f"{HANDLER_PREFIX}.rpc_inference", hop.key, [], open_md,
                                             timeout=self.request_timeout)
                self._streams[key] = st
            return st.call(tensors, md, timeout=self.request_timeout)
        except (TimeoutError, ConnectionError):
            self._drop_stream(key)
            raise

    def _drop_stream(self, key) -> None:
        st = self._streams.pop(key, None)
        if st is not None:
            try:
                st.close(timeout=1.0)
            except Exception:  # noqa: BLE001
                pass

    def reorder_hypotheses(self, session_ids: Sequence[str], hypo_ids: Sequence[int]) -> None:
        """Beam search: on every server of the beam's route, session ``session_ids[i]`` continues the hypothesis
        that ``session_ids[hypo_ids[i]]`` held (upstream ``hypo_ids``, petals/server/backend.py:154-158).  All
        sessions of a beam share one route (the first session's)."""
        route = self._get_route(session_ids[0])
        for sid in session_ids[1:]:
            self.session_routes.setdefault(sid, route)
        for hop in route:
            self.client.call(hop.addr, f"{HANDLER_PREFIX}.rpc_reorder", hop.key, [],
                             {"session_ids": list(session_ids), "hypo_ids": [int(h) for h in hypo_ids]},
                             timeout=self.request_timeout)
        # the journals follow the hypotheses, so that a later replay rebuilds the re-ordered caches
        snap = {sid: {k: list(v) for k, v in self.client_cache.get(sid, {}).items()} for sid in session_ids}
        for i, h in enumerate(hypo_ids):
            if int(h) != i:
                self.client_cache[session_ids[i]] = {k: list(v) for k, v in snap[session_ids[int(h)]].items()}

    def _journal(self, session_id: str, key: str) -> List[torch.Tensor]:
        return self.client_cache.setdefault(session_id, {}).setdefault(key, [])

    def _replay(self, session_id: str, hops: List[Hop], first_key: str) -> None:
        """Rebuild the KV caches of ``hops`` (a route suffix) from the journal of its first hop."""
        past = list(self._journal(session_id, first_key))
        if not past:
            return
        logger.info("replaying %d journaled steps into %d hop(s)", len(past), len(hops))
        for k in [h.key for h in hops[1:]]:
            self.client_cache[session_id][k] = []        # re-recorded below
        cur_len = 0
        for i, t in enumerate(past):
            T = t.shape[-2]
            cur_len += T
            x = t
            for j, hop in enumerate(hops):
                if hop.final:
                    md = self._metadata(session_id, T, cur_len, i == 0, None, is_replay=True)
                    md["temperature"] = 0.0              # the token of a replayed step is discarded
                    self._call(hop, x, md)
                    break
                if j > 0:
                    self._journal(session_id, hop.key).append(x.clone())
                md = self._metadata(session_id, T, cur_len, i == 0, None, is_replay=True)
                out, omd = self._call(hop, x, md, consumer_in_box=j + 1 < len(hops) and self._in_box(hops[j + 1]))
                x, _ = self._take(out, omd)

    def _take(self, out, omd) -> Tuple[torch.Tensor, Optional[dict]]:
        """A hop's output: the tensor of the reply, or - handed off on the device - this GPU's pull of the slot the
        reply names (that copy is what gets journaled; the descriptor itself travels on to the next hop)."""
        desc = omd.get("device_tensor") if isinstance(omd, dict) else None
        if desc is not None:
            return self._importer.pull(desc), desc
        return out[0], None

    def _send(self, session_id: str, hidden: torch.Tensor, cur_len: int, is_prefill: bool,
              generated: Optional[Sequence[int]]) -> int:
        big = hidden.numel() * hidden.element_size() >= self.device_handoff_min_bytes
        cur = hidden.detach() if self.device_handoff and hidden.is_cuda and big else hidden.detach().to("cpu")
        cur_desc = None
        T = cur.shape[-2]
        route = self._get_route(session_id)
        t_all = time.perf_counter()
        i = 0
        while i < len(route):
            hop = route[i]
            md = self._metadata(session_id, T, cur_len, is_prefill, generated)
            if self.push and not hop.final:
                md["route"] = [[h.addr[0], h.addr[1], h.key] for h in route[i + 1:]]
            attempt = 0
            while True:
                t0 = time.perf_counter()
                try:
                    if self.push:
                        chained = not hop.final and all(self._in_box(h) for h in route[i:])
                    else:
                        chained = i + 1 < len(route) and self._in_box(route[i + 1])
                    out, omd = self._call(hop, cur, md, desc=cur_desc if attempt == 0 else None,
                                          consumer_in_box=chained)
                    break
                except (TimeoutError, ConnectionError, RpcError, RuntimeError, ValueError) as e:
                    attempt += 1
                    logger.warning("hop %s (%s) failed: %s", hop.key, hop.peer_id[:12], e)
                    if attempt >= MAX_ATTEMPTS:
                        raise
                    if self.push:
                        # the error may come from any server down the chain: find the first dead one
                        # and continue hop by hop from where the chain is known to be intact
                        dead = self._find_dead_hop(route, i)
                        self.push_fallbacks = getattr(self, "push_fallbacks", 0) + 1
                        md.pop("route", None)
                        if dead is not None and dead > i:
                            # hops [i, dead) consumed the step already: rebuild them too by treating
                            # hop i as the recovery point (replay is idempotent: prefill resets caches)
                            self.failed_peers.setdefault(route[dead].key, set()).add(route[dead].peer_id)
                            e2 = None
                            if self.routing == "module":
                                excl = set().union(*self.failed_peers.values())
                                route = route[:i] + self._module_route(hop.start_block, excl)
                            else:
                                e2 = self._discover_stage(route[dead].key, self.failed_peers[route[dead].key], tries=5)
                                route = route[:dead] + [Hop(route[dead].key, e2, route[dead].final)] + route[dead + 1:]
                            self.session_routes[session_id] = route
                            self._replay(session_id, route[i:], route[i].key)
                            hop = route[i]
                            self.recoveries += 1
                            self.push = False          # finish this session hop by hop
                            time.sleep(0.2)
                            continue
                        if dead is None:
                            # every server of the chain answers: the error was transient and came from some hop
                            # at or after i, but hops [i, failing) have already appended this step.  Nobody is to
                            # blame; rebuild the whole suffix from hop i's journal (a replayed prefill resets each
                            # cache) and resend the step hop by hop.
                            self._replay(session_id, route[i:], hop.key)
                            self.recoveries += 1
                            self.push = False
                            metrics.emit("recovery", session=session_id, hop=hop.key, peer=hop.peer_id[:12],
                                         error="transient downstream error: " + str(e)[:160])
                            time.sleep(0.2)
                            continue
                    self.failed_peers.setdefault(hop.key, set()).add(hop.peer_id)
                    logger.info("Marked peer %s as failed for %s", hop.peer_id[:12], hop.key)
                    metrics.emit("recovery", session=session_id, hop=hop.key, peer=hop.peer_id[:12], error=str(e)[:200])
                    route, hop = self._recover(session_id, route, i)
                    if self.push and not hop.final:
                        md["route"] = [[h.addr[0], h.addr[1], h.key] for h in route[i + 1:]]
                    time.sleep(0.2)
            # journal only what the server has actually consumed
            self._journal(session_id, hop.key).append(cur.clone())
            self.hop_times.append((hop.key, time.perf_counter() - t0))
            metrics.emit("hop", session=session_id, hop=hop.key, peer=hop.peer_id[:12], seconds=self.hop_times[-1][1],
                         tokens=int(T), prefill=bool(is_prefill))
            if "token_id" in omd or hop.final or (self.push and "token_id" in omd):
                self.last_token = int(omd.get("token_id", out[0].reshape(-1)[0].item()))
                self.last_reply = dict(omd)
                break
            cur, cur_desc = self._take(out, omd)
            i += 1
        total = time.perf_counter() - t_all
        if is_prefill:
            self.last_prefill_total = total
        else:
            self.last_decode_total = total
            self.decode_totals.append(total)
        return self.last_token

    def _find_dead_hop(self, route: List[Hop], i: int) -> Optional[int]:
        for j in range(i, len(route)):
            try:
                self.client.call(route[j].addr, f"{HANDLER_PREFIX}.rpc_info", route[j].key, [], {}, timeout=2.0)
            except Exception:  # noqa: BLE001
                return j
        return None

    def _recover(self, session_id: str, route: List[Hop], i: int) -> Tuple[List[Hop], Hop]:
        self.recoveries += 1
        hop = route[i]
        if self.routing == "module":
            excl = set().union(*self.failed_peers.values()) if self.failed_peers else set()
            suffix = self._module_route(hop.start_block, excl)
            route = route[:i] + suffix
            self.session_routes[session_id] = route
            self._replay(session_id, suffix, suffix[0].key)
        else:
            e = self._discover_stage(hop.key, self.failed_peers.setdefault(hop.key, set()), tries=2)
            new = Hop(hop.key, e, hop.final)
            route = route[:i] + [new] + route[i + 1:]
            self.session_routes[session_id] = route
            self._replay(session_id, [new] if not new.final else [new], new.key)
        return route, route[i]

    # ------------------------------------------------------------------ training-side passes (no sessions)
    def _stateless_hop(self, route: List[Hop], i: int, method: str, tensors: List[torch.Tensor]):
        """One stateless call on hop i; a dead server is replaced and the call repeated (nothing to replay)."""
        for attempt in range(MAX_ATTEMPTS):
            hop = route[i]
            md = {"stateless": True}
            if hop.partial:
                md["blocks"] = [int(hop.start_block), int(hop.end_block)]
            try:
                out, _ = self.client.call(hop.addr, f"{HANDLER_PREFIX}.{method}", hop.key, tensors, md,
                                          timeout=self.request_timeout)
                return out[0]
            except (TimeoutError, ConnectionError, RpcError) as e:
                if attempt == MAX_ATTEMPTS - 1:
                    raise
                logger.warning("stateless %s on %s (%s) failed: %s", method, hop.key, hop.peer_id[:12], e)
                self.failed_peers.setdefault(hop.key, set()).add(hop.peer_id)
                self.recoveries += 1
                if self.routing == "module":
                    excl = set().union(*self.failed_peers.values())
                    route[i:] = self._module_route(hop.start_block, excl)
                else:
                    route[i] = Hop(hop.key, self._discover_stage(hop.key, self.failed_peers[hop.key], tries=5), hop.final)

    def forward_stateless(self, hidden: torch.Tensor) -> Tuple[torch.Tensor, list]:
        """Whole sequences ``hidden`` [B, T, H] through every server span of a fresh route (upstream
        ``sequential_forward``): returns the last span's output and what :meth:`backward_stateless` needs."""
        route = list(self._get_route(f"stateless-{time.time_ns()}"))
        x = hidden.detach().to("cpu")
        saved = []
        i = 0
        while i < len(route):
            y = self._stateless_hop(route, i, "rpc_forward", [x])
            saved.append(x)
            x = y
            i += 1
        return x, [route, saved]

    def backward_stateless(self, ctx: list, grad_out: torch.Tensor) -> torch.Tensor:
        """Gradient w.r.t. the input of :meth:`forward_stateless` (upstream ``sequential_backward``): every span
        recomputes its forward from the saved input and back-propagates."""
        route, saved = ctx
        g = grad_out.detach().to("cpu")
        for i in reversed(range(len(saved))):
            g = self._stateless_hop(route, i, "rpc_backward", [saved[i], g])
        return g

    # ------------------------------------------------------------------ public API (reference names)
    def send_prefill(self, L: int, hidden: torch.Tensor, session_id: str, max_length: int) -> None:
        # a prefill starts the session over on every server (fresh cache), so the journal starts over too:
        # a later replay must not rebuild "old session + new prompt"
        self.client_cache.pop(session_id, None)
        self.session_max_length[session_id] = int(max_length)
        self._send(session_id, hidden, L, True, None)

    def send_decode_step(self, cur_len: int, hidden: torch.Tensor, session_id: str, max_length: int,
                         generated_tokens: Optional[Sequence[int]] = None) -> None:
        self.session_max_length.setdefault(session_id, int(max_length))
        self._send(session_id, hidden, cur_len, False, generated_tokens)

    def recv_token(self) -> int:
        if self.last_token is None:
            raise RuntimeError("no token received yet")
        return self.last_token

    def close_session(self, session_id: str) -> None:
        for key in [k for k in self._streams if k[0] == session_id]:
            self._drop_stream(key)                       # ending the stream frees the server-side cache
        for hop in ([] if self.stream else self.session_routes.get(session_id, [])):
            try:
                self.client.call(hop.addr, f"{HANDLER_PREFIX}.rpc_close_session", hop.key, [],
                                 {"session_id": session_id}, timeout=5.0)
            except Exception:  # noqa: BLE001 - best effort
                pass
        self.session_routes.pop(session_id, None)
        self.client_cache.pop(session_id, None)
        self.session_max_length.pop(session_id, None)

    def server_info(self, session_id: str) -> List[dict]:
        return [self.client.call(h.addr, f"{HANDLER_PREFIX}.rpc_info", h.key, [], {}, timeout=5.0)[1]
                for h in self._get_route(session_id)]

    def latency_summary(self) -> Dict[str, float]:
        d = self.decode_totals
        if not d:
            return {}
        s = sorted(d)
        return {"decode_p50_ms": statistics.median(s) * 1e3, "decode_p99_ms": s[min(len(s) - 1, int(0.99 * len(s)))] * 1e3,
                "decode_mean_ms": sum(s) / len(s) * 1e3, "steps": len(s)}

    def shutdown(self) -> None:
        for key in list(self._streams):
            self._drop_stream(key)
        self.client.close()
        for side in (self._importer, self._exporter):
            if side is not None:
                try:
                    side.close()
                except Exception:  # noqa: BLE001
                    pass
        self._importer = self._exporter = None
        if self._own_dht:
            self.dht.shutdown()

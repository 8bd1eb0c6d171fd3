"""Command-line entry point: ``python -m src.main --model M --splits a,b,c --stage N ...``

Same flags and roles as the reference (src/main.py:775-838):

* ``--stage 0``  client ("rank 0"): embeds the prompt, runs blocks ``[0, splits[0])`` locally, sends
  hidden states through the server stages and prints the generated text with TTFT / decode /
  total times (``run_rank0``).
* ``--stage k``  server for blocks ``[splits[k-1], splits[k])`` (the last one adds final norm + LM
  head + sampling).  Any number of stages works: ``--splits 6`` is client + one final stage,
  ``--splits 8,16,24`` the reference's four (the reference hard-codes exactly three servers, B3).
* ``--use_load_balancing``  the server ignores ``--stage``/``splits[1:]``, asks the DHT where the
  swarm is weakest, serves ``--num_blocks`` blocks there and periodically re-evaluates
  (``run_stage_server_with_load_balancing``).

Log lines other tools scrape are kept verbatim: ``DHT visible multiaddrs (use for
--dht_initial_peers): [...]``, ``Stage{N} handlers registered, waiting for requests...``,
``Load Balancing server ready, blocks=..., throughput=... rps``.
"""
from __future__ import annotations

import argparse
import logging
import os
import sys
import random
import threading
import time
from typing import List, Optional
from uuid import uuid4

import torch

from petals_b200.models.config import resolve_config
from petals_b200.net.dht import DHT, PeerID, get_dht_time
from petals_b200.net.rpc import RpcServer
from petals_b200.utils.tokenizer import load_tokenizer

from .dht_utils import (announce_offline, declare_model, get_remote_module_infos, register_blocks_on_dht,
                        register_server_on_dht, register_stage_on_dht)
from .load_balancing import choose_best_blocks, should_choose_other_blocks
from .partition import Stage0, StageLast, StageSegment, load_stage_model
from .rpc_handler import StageConnectionHandler
from .rpc_transport import RpcTransport
from .throughput_measurement import get_server_throughput

logger = logging.getLogger(__name__)
logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s %(name)s: %(message)s")

HEARTBEAT_PERIOD = 15.0
RECORD_TTL = 45.0


def parse_splits(splits_str: str) -> List[int]:
    return [int(x.strip()) for x in splits_str.split(",") if x.strip()]


def _format_initial_peers(dht_initial_peers: str) -> List[str]:
    return [p.strip() for p in (dht_initial_peers or "").split(",") if p.strip()]


def _get_local_ip() -> str:
    import socket

    try:
        s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        s.connect(("8.8.8.8", 80))
        ip = s.getsockname()[0]
        s.close()
        return ip
    except OSError:
        return "127.0.0.1"


def stage_span(splits: List[int], stage: int, n_layers: int):
    """Blocks and role of server stage ``stage`` (1-based) for cut points ``splits``."""
    if not 1 <= stage <= len(splits):
        raise ValueError(f"--stage must be in 1..{len(splits)} for --splits {splits}")
    start = splits[stage - 1]
    final = stage == len(splits)
    end = n_layers if final else splits[stage]
    return start, end, final


# --------------------------------------------------------------------------- client
@torch.inference_mode()
def run_rank0(args, device, splits, dht: Optional[DHT] = None):
    """Client generation loop; returns ``(generated ids, stats)`` after printing like the reference."""
    cfg = resolve_config(args.model)
    stage0_end = splits[0]
    full = load_stage_model(args.model, device, role="stage0", end=stage0_end, dtype=args.dtype,
                            seed=getattr(args, "seed", 0))
    max_new = args.max_new_tokens
    tok = load_tokenizer(args.model, cfg)
    prompt = args.prompt
    input_ids = torch.tensor([tok.encode(prompt)], dtype=torch.long)
    L = input_ids.size(1)
    max_length = L + max_new
    s0 = Stage0(full, stage0_end, kv_capacity=max(getattr(args, "kv_capacity", 0) or 0, max_length))
    kv0 = s0.new_kv_manager().allocate("client", s0.kv_capacity)
    total_blocks = args.total_blocks or cfg.n_layers
    n_server_stages = len(splits)
    tx = RpcTransport(device=device, stage=0, dht_initial_peers=_format_initial_peers(args.dht_initial_peers),
                      dht_port=args.dht_port, rpc_port=args.rpc_port, temperature=args.temperature, top_p=args.top_p,
                      top_k=args.top_k, stage_keys=[f"mini_petals:stage{i}" for i in range(1, n_server_stages + 1)],
                      routing=("module" if args.use_load_balancing else "stage"), model_name=args.model,
                      total_blocks=total_blocks, start_block=stage0_end, request_timeout=args.request_timeout,
                      repetition_penalty=getattr(args, "repetition_penalty", None), dht=dht,
                      push=getattr(args, "push", False), stream=getattr(args, "stream", False),
                      device_handoff=getattr(args, "device_handoff", False),
                      device_handoff_min_bytes=int(getattr(args, "device_handoff_min_kb", 128)) << 10)
    session_id = str(uuid4())
    t_prefill_start = time.perf_counter()
    hidden, _ = s0(input_ids, None, None, kv0)
    tx.send_prefill(L, hidden, session_id=session_id, max_length=max_length)
    next_id = tx.recv_token()
    generated = [next_id]
    on_token = getattr(args, "on_token", None)        # optional observer (benchmarks / fault injection scripts)
    if on_token is not None:
        on_token(0, next_id)
    prefill_time = time.perf_counter() - t_prefill_start
    logger.info("Prefill completed in %.3fs", prefill_time)
    logger.info("Generated token: %d (%s)", next_id, tok.decode([next_id]))
    cur_len = L + 1
    t_decode_start = time.perf_counter()
    eos = tok.eos_token_id
    repeat, last = 0, None
    for _ in range(max_new - 1):
        hidden, _ = s0(torch.tensor([[next_id]], dtype=torch.long), None, None, kv0)
        tx.send_decode_step(cur_len, hidden, session_id=session_id, max_length=max_length, generated_tokens=generated)
        next_id = tx.recv_token()
        if on_token is not None:
            on_token(len(generated), next_id)
        if eos is not None and next_id == eos:
            logger.info("EOS token generated, stopping generation")
            break
        if next_id == last:
            repeat += 1
            if repeat >= 5 and not getattr(args, "no_repeat_stop", False):
                logger.warning("Consecutive repetition detected (token %d), stopping generation", next_id)
                break
        else:
            repeat, last = 0, next_id
        generated.append(next_id)
        cur_len += 1
    decode_time = time.perf_counter() - t_decode_start
    total_time = time.perf_counter() - t_prefill_start
    text = tok.decode(generated)
    print(f"\n{'=' * 80}\nPROMPT: {prompt}\nGENERATED: {text}\n{'=' * 80}\n")
    logger.info("Decode completed in %.3fs", decode_time)
    logger.info("Total time: %.3fs", total_time)
    logger.info("TTFT (Time to First Token): %.3fs", prefill_time)
    stats = {"ttft_s": prefill_time, "decode_s": decode_time, "total_s": total_time,
             "decode_tok_per_s": max(len(generated) - 1, 0) / max(decode_time, 1e-9), "recoveries": tx.recoveries,
             **tx.latency_summary()}
    if stats.get("decode_p50_ms"):
        logger.info("Decode per-token latency p50 %.2f ms, p99 %.2f ms", stats["decode_p50_ms"], stats["decode_p99_ms"])
    tx.close_session(session_id)
    tx.shutdown()
    return generated, stats


# --------------------------------------------------------------------------- servers
class ServerHandle:
    """A running stage server (DHT node + RPC server + handler + heartbeat)."""

    def __init__(self, dht, rpc, handler, peer_id, maddrs, stop_event, threads, blocks, model_name):
        self.dht, self.rpc, self.handler, self.peer_id, self.maddrs = dht, rpc, handler, peer_id, maddrs
        self._stop, self._threads, self.blocks, self.model_name = stop_event, threads, blocks, model_name
        self.rebalance_requested = threading.Event()

    def stop(self, announce: bool = True) -> None:
        self._stop.set()
        if announce and self.blocks:
            try:
                announce_offline(self.dht, self.peer_id, self.blocks, self.model_name, p2p_maddrs=self.maddrs)
            except Exception:  # noqa: BLE001
                pass
        self.rpc.stop()
        self.handler.shutdown()
        self.dht.shutdown()

    def wait(self, timeout: Optional[float] = None) -> bool:
        """Blocks until stopped or a re-balance is requested; True => re-place this server."""
        t0 = time.time()
        while not self._stop.is_set():
            if self.rebalance_requested.wait(0.5):
                return True
            if timeout is not None and time.time() - t0 > timeout:
                return False
        return False


def _make_dht(args, port_offset: int = 0) -> DHT:
    peers = _format_initial_peers(args.dht_initial_peers)
    announce_ip = args.public_ip or None
    public_port = args.public_dht_port if args.public_dht_port is not None else args.dht_port
    host = "0.0.0.0" if args.public_ip else "127.0.0.1"
    dht = DHT(start=True, initial_peers=peers, host_maddrs=[f"/ip4/{host}/tcp/{args.dht_port}"],
              announce_maddrs=[f"/ip4/{announce_ip}/tcp/{public_port}"] if announce_ip else None)
    logger.info("DHT visible multiaddrs (use for --dht_initial_peers): %s", [str(m) for m in dht.get_visible_maddrs()])
    return dht


def start_stage_server(args, device, stage_model, final_stage: bool, stage_index: Optional[int] = None,
                       lb_blocks: Optional[List[int]] = None, throughput: float = 0.0,
                       dht: Optional[DHT] = None) -> ServerHandle:
    dht = dht or _make_dht(args)
    host = "0.0.0.0" if args.public_ip else "127.0.0.1"
    rpc = RpcServer(host, args.rpc_port).start()
    peer_id = PeerID.generate()
    ip = args.public_ip or "127.0.0.1"
    port = args.public_rpc_port if args.public_rpc_port is not None else rpc.port
    maddrs = [f"/ip4/{ip}/tcp/{port}/p2p/{peer_id}"]
    handler = StageConnectionHandler(dht, stage_model, device, request_timeout=args.request_timeout,
                                     final_stage=final_stage, throughput=throughput,
                                     seed=getattr(args, "seed", None),
                                     attn_cache_tokens=_attn_cache_tokens(args, stage_model),
                                     alloc_timeout=getattr(args, "alloc_timeout", 10.0),
                                     session_timeout=getattr(args, "session_timeout", 30 * 60.0),
                                     step_timeout=getattr(args, "step_timeout", 5 * 60.0),
                                     max_batch=getattr(args, "max_batch", None),
                                     batch_linger=getattr(args, "batch_linger_ms", 2.0) * 1e-3)
    # many handler threads + one compute thread: hand the GIL over faster than CPython's 5 ms default
    sys.setswitchinterval(float(os.environ.get("B200_GIL_SWITCH_US", "200")) * 1e-6)
    stop = threading.Event()
    model_name = args.model
    handler.on_die = lambda: (stop.set(), rpc.stop())   # scripted kill (tests / kill_stage.py --soft)

    def next_pings():
        """RTTs to servers that start where this one ends (upstream ModuleAnnouncerThread, server.py:721-735)."""
        if not lb_blocks:
            return {}
        from petals_b200.net.reachability import check_direct_reachability
        from .dht_utils import get_block_entries

        out = {}
        for e in get_block_entries(dht, lb_blocks[-1] + 1, model_name)[:5]:
            if e.get("p2p_maddrs") and str(e.get("peer_id")) != str(peer_id):
                ok, rtt = check_direct_reachability(e["p2p_maddrs"][0], 1.0)
                if ok:
                    out[str(e["peer_id"])] = rtt
        return out

    def publish():
        exp = get_dht_time() + RECORD_TTL
        cfg_ = getattr(stage_model, "cfg", None)
        declare_model(dht, model_name, peer_id, {
            "n_layers": getattr(cfg_, "n_layers", None), "hidden": getattr(cfg_, "hidden", None),
            "blocks": [handler.start_block, handler.end_block], "final_stage": final_stage,
            "dtype": str(getattr(stage_model, "dtype", "")).replace("torch.", ""), "p2p_maddrs": maddrs,
            "stage": stage_index, "throughput": handler.throughput}, expiration_time=exp)
        if stage_index is not None:
            extra = {}
            if lb_blocks:
                extra = {"blocks": [lb_blocks[0], lb_blocks[-1] + 1], "throughput": throughput}
            register_stage_on_dht(dht, stage_index, peer_id, maddrs, expiration_time=exp, **extra)
        if lb_blocks:
            register_server_on_dht(dht, peer_id, lb_blocks[0], lb_blocks[-1] + 1, handler.throughput, model_name,
                                   server_address=maddrs[0], p2p_maddrs=maddrs, final_stage=final_stage,
                                   expiration_time=exp)
            left = handler.kv.tokens_left
            register_blocks_on_dht(dht, peer_id, lb_blocks, model_name, p2p_maddrs=maddrs, start_block=lb_blocks[0],
                                   end_block=lb_blocks[-1] + 1, throughput=handler.throughput,
                                   final_stage=final_stage, expiration_time=exp,
                                   cache_tokens_left=-1 if left is None else left, next_pings=next_pings())

    def heartbeat():
        while not stop.wait(HEARTBEAT_PERIOD):
            try:
                publish()
            except Exception as e:  # noqa: BLE001
                logger.warning("heartbeat failed: %s", e)

    publish()
    handler.add_p2p_handlers(rpc)
    t = threading.Thread(target=heartbeat, name="heartbeat", daemon=True)
    t.start()
    logger.info("Stage%s handlers registered, waiting for requests...", stage_index if stage_index is not None else "")
    return ServerHandle(dht, rpc, handler, peer_id, maddrs, stop, [t], lb_blocks or [], model_name)


def _attn_cache_tokens(args, stage_model) -> Optional[int]:
    """KV budget of a server in tokens (upstream --attn_cache_tokens, petals/server/server.py:204-209): the
    upstream defaults, but never less than what ``max_batch`` full-length sessions need on this stage."""
    from petals_b200.runtime.block_utils import default_limits

    v = getattr(args, "attn_cache_tokens", None)
    if v == 0:
        return None
    if v is None:
        v = default_limits(stage_model.cfg)[1]
        v = max(v, (getattr(args, "max_batch", None) or 1) * stage_model.kv_capacity)
    return max(int(v), stage_model.kv_capacity)


def _default_num_blocks(args, cfg, device) -> int:
    """--num_blocks when the user gave none: as many blocks as this GPU holds next to its KV budget (upstream
    ``_choose_num_blocks``, petals/server/server.py:275-326); 4 on a CPU-only host."""
    from petals_b200.runtime.block_utils import choose_num_blocks, default_limits

    if device.type != "cuda":
        return min(4, cfg.n_layers)
    tokens = getattr(args, "attn_cache_tokens", None) or default_limits(cfg)[1]
    n = choose_num_blocks(cfg, device, args.dtype if isinstance(args.dtype, torch.dtype) else torch.bfloat16,
                          weight_format=getattr(args, "weight_format", "native") or "native", attn_cache_tokens=tokens)
    logger.info("num_blocks not given: this GPU holds %d of %d blocks next to a %d-token cache", n, cfg.n_layers, tokens)
    return n


def _build_stage(args, device, start: int, end: int, final: bool, n_layers: int):
    role = "last" if final else "segment"
    full = load_stage_model(args.model, device, role=role, start=start, end=end, dtype=args.dtype,
                            use_cpu_offload=args.use_cpu_offload, seed=getattr(args, "seed", 0))
    kw = dict(gpu_device=device, keep_layers_on_gpu=args.keep_layers_on_gpu)
    if getattr(args, "kv_capacity", None):
        kw["kv_capacity"] = args.kv_capacity
    if getattr(args, "max_batch", None):
        kw["max_batch"] = args.max_batch
    wf = getattr(args, "weight_format", "native") or "native"
    if wf != "native":
        if device.type != "cuda" or _tp_width(args) > 1 or args.use_cpu_offload:
            raise ValueError("--weight_format fp8/mxfp8 needs a resident single-GPU span on CUDA")
        kw["weight_format"] = wf
    if _tp_width(args) > 1:
        # a span wider than one GPU: this process holds one tensor-parallel shard of it (runtime/tp_stage.py)
        import torch.distributed as dist

        from petals_b200.runtime.tp_stage import TPStage

        kw.pop("keep_layers_on_gpu", None)
        return TPStage(full, start, end, dist.group.WORLD, final=final, **kw)
    return StageLast(full, start, **kw) if final else StageSegment(full, start, end, **kw)


def _tp_width(args) -> int:
    return int(getattr(args, "tp", 1) or 1) if not getattr(args, "inbox", False) else 1


def _tp_setup(args, device) -> int:
    """Served stage with --tp k: one process per GPU (torchrun), rank 0 serves, the others follow.  Returns the rank."""
    import torch.distributed as dist

    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    if dist.get_world_size() != _tp_width(args):
        raise SystemExit(f"--tp {args.tp} needs exactly {args.tp} processes (torchrun --nproc-per-node {args.tp}); "
                         f"got {dist.get_world_size()}")
    return dist.get_rank()


def run_stage_server_fixed(args, device, splits, block: bool = True):
    cfg = resolve_config(args.model)
    start, end, final = stage_span(splits, args.stage, cfg.n_layers)
    tp_rank = _tp_setup(args, device) if _tp_width(args) > 1 else 0
    stage_model = _build_stage(args, device, start, end, final, cfg.n_layers)
    if tp_rank > 0:
        from petals_b200.runtime.tp_stage import serve_follower

        serve_follower(stage_model)        # returns when the leader stops
        return None
    handle = start_stage_server(args, device, stage_model, final, stage_index=args.stage)
    if block:
        handle.wait()
    return handle


def run_stage_server_with_load_balancing(args, device, splits, num_blocks=None, total_blocks=None,
                                         max_placements: Optional[int] = None, block: bool = True):
    """Join where the swarm is weakest, serve, re-evaluate every ~``mean_balance_check_period`` s."""
    cfg = resolve_config(args.model)
    total_blocks = total_blocks or args.total_blocks or cfg.n_layers
    num_blocks = num_blocks or args.num_blocks or _default_num_blocks(args, cfg, device)
    num_blocks = min(num_blocks, total_blocks - (splits[0] if splits else 0))
    min_block = splits[0] if splits else 0
    tp = _tp_width(args)
    tp_rank = _tp_setup(args, device) if tp > 1 else 0
    if tp > 1:
        import torch.distributed as dist

        # the leader looks at the swarm and decides; the followers take its word (one placement: re-placing a
        # tensor-parallel server would mean reloading every shard, so it keeps its blocks)
        max_placements = 1
        span = [0, 0]
        if tp_rank > 0:
            dist.broadcast_object_list(span, src=0)
            from petals_b200.runtime.tp_stage import serve_follower

            stage_model = _build_stage(args, device, span[0], span[1], span[1] >= total_blocks, cfg.n_layers)
            serve_follower(stage_model)
            return None
    dht = _make_dht(args)
    placements = 0
    handle = None
    while True:
        # Servers started together (one script, one box) would all see the same empty swarm and pick the same
        # blocks: spread their decisions over time (upstream server.py:407-411).
        sel_delay = float(getattr(args, "mean_block_selection_delay", 0.0) or 0.0)
        if sel_delay > 0:
            time.sleep(random.random() * 2 * sel_delay)
        infos, delay = [], 1.0
        for _ in range(3):
            infos = get_remote_module_infos(dht, args.model, total_blocks)
            if any(i.server_info is not None for i in infos):
                break
            time.sleep(delay)
            delay *= 1.5
        blocks = choose_best_blocks(num_blocks, infos, total_blocks=total_blocks, min_block=min_block)
        start, end = blocks[0], min(blocks[-1] + 1, total_blocks)
        blocks = list(range(start, end))
        final = end >= total_blocks
        logger.info("Selected blocks [%d:%d) (final_stage=%s)", start, end, final)
        if tp > 1:
            import torch.distributed as dist

            dist.broadcast_object_list([start, end], src=0)
        stage_model = _build_stage(args, device, start, end, final, cfg.n_layers)
        try:
            thr = get_server_throughput(stage_model, device, args.dtype, cfg.hidden,
                                        network_bandwidth_mbps=args.network_bandwidth_mbps, model_name=args.model,
                                        use_cache=not getattr(args, "no_throughput_cache", False))
        except Exception as e:  # noqa: BLE001
            logger.warning("throughput measurement failed (%s); using fallback 10.0 rps", e)
            thr = 10.0
        handle = start_stage_server(args, device, stage_model, final, stage_index=None, lb_blocks=blocks,
                                    throughput=thr, dht=dht)
        logger.info("Load Balancing server ready, blocks=[%d:%d), throughput=%.2f rps", start, end, thr)
        placements += 1

        def rebalance_check(h=handle, sm=stage_model):
            period = args.mean_balance_check_period
            while not h._stop.is_set():
                if h._stop.wait(random.uniform(0, 2 * period)):
                    return
                try:
                    h.handler.throughput = get_server_throughput(
                        sm, device, args.dtype, cfg.hidden, network_bandwidth_mbps=args.network_bandwidth_mbps,
                        model_name=args.model, use_cache=not getattr(args, "no_throughput_cache", False))
                    mi = get_remote_module_infos(dht, args.model, total_blocks)
                    if should_choose_other_blocks(str(h.peer_id), mi, args.balance_quality, total_blocks, min_block):
                        logger.info("Swarm is imbalanced: re-placing this server")
                        h.rebalance_requested.set()
                        return
                except Exception as e:  # noqa: BLE001
                    logger.warning("rebalance check failed: %s", e)

        threading.Thread(target=rebalance_check, name="rebalance", daemon=True).start()
        if not block or (max_placements is not None and placements >= max_placements):
            return handle
        if not handle.wait():
            return handle
        # re-place: leave (OFFLINE announce), keep the DHT node, load other blocks
        handle._stop.set()
        announce_offline(dht, handle.peer_id, handle.blocks, args.model, p2p_maddrs=handle.maddrs)
        handle.rpc.stop()
        handle.handler.shutdown()
        del stage_model
        if torch.cuda.is_available():
            torch.cuda.empty_cache()


def run_stage_server(args, device, splits):
    if getattr(args, "use_load_balancing", False):
        return run_stage_server_with_load_balancing(args, device, splits, args.num_blocks, args.total_blocks)
    return run_stage_server_fixed(args, device, splits)


# --------------------------------------------------------------------------- in-box NCCL pipeline
def run_inbox(args, device, splits):
    """All stages inside one NVSwitch box: one process per GPU (torchrun), NCCL p2p chain.

    ``--splits a,b`` with 3 x tp processes puts blocks [0,a) on stage 0, [a,b) on stage 1, [b,L) on stage 2
    (the first stage also embeds, the last one samples); rank 0 prints what the reference client prints."""
    import torch.distributed as dist

    from petals_b200.parallel.pipeline import PipelineEngine

    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = "nccl" if device.type == "cuda" else "gloo"
        kw = {"device_id": device} if backend == "nccl" else {}
        dist.init_process_group(backend, **kw)
    cfg = resolve_config(args.model)
    tok = load_tokenizer(args.model, cfg)
    ids = torch.tensor(tok.encode(args.prompt), dtype=torch.long)
    cap = max(getattr(args, "kv_capacity", None) or 0, ((ids.numel() + args.max_new_tokens + 255) // 256) * 256)
    from petals_b200.runtime.sampler import SamplingParams

    sampling = None
    if args.temperature > 0:      # same knobs as the TCP path; the last stage samples on the device
        rp = getattr(args, "repetition_penalty", None)
        sampling = SamplingParams(args.temperature, args.top_p, args.top_k, 1.5 if rp is None else rp)
    eng = PipelineEngine(args.model, splits, args.dtype, device, kv_capacity=cap, max_sessions=1,
                         seed=getattr(args, "seed", 0), tp=args.tp, head_split=False if sampling else None)
    res = eng.generate([ids], args.max_new_tokens, eos_token_id=tok.eos_token_id, sampling=sampling,
                       seed=getattr(args, "seed", 0))
    if dist.get_rank() == 0:
        outs, stats = res
        print(f"\n{'=' * 80}\nPROMPT: {args.prompt}\nGENERATED: {tok.decode(outs[0])}\n{'=' * 80}\n")
        logger.info("Decode completed in %.3fs", stats["decode_s"])
        logger.info("Total time: %.3fs", stats["total_s"])
        logger.info("TTFT (Time to First Token): %.3fs", stats["ttft_s"])
        logger.info("Decode throughput: %.1f tok/s", stats["decode_tok_per_s"])
    dist.barrier()
    return res


# --------------------------------------------------------------------------- CLI
def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Mini Petals on B200: distributed pipeline inference")
    p.add_argument("--model", type=str, required=True)
    p.add_argument("--splits", type=str, required=True, help="comma-separated cut points, e.g. 8,16,24 (or 6)")
    p.add_argument("--dtype", type=str, default="fp16", choices=["fp16", "bf16", "fp32"])
    p.add_argument("--max_new_tokens", type=int, default=64)
    p.add_argument("--prompt", type=str, default="Hello, how are you?")
    p.add_argument("--dht_initial_peers", type=str, default="")
    p.add_argument("--public_ip", type=str, default="")
    p.add_argument("--public_dht_port", type=int, default=None)
    p.add_argument("--public_rpc_port", type=int, default=None)
    p.add_argument("--dht_port", type=int, default=8000)
    p.add_argument("--rpc_port", type=int, default=8001)
    p.add_argument("--stage", type=int, required=True, help="0 = client, k >= 1 = server stage k")
    p.add_argument("--request_timeout", type=float, default=30.0)
    p.add_argument("--temperature", type=float, default=1.0)
    p.add_argument("--top_p", type=float, default=0.92)
    p.add_argument("--top_k", type=int, default=50)
    p.add_argument("--use_cpu_offload", action="store_true")
    p.add_argument("--keep_layers_on_gpu", type=int, default=0)
    p.add_argument("--use_load_balancing", action="store_true")
    p.add_argument("--num_blocks", type=int, default=None)
    p.add_argument("--total_blocks", type=int, default=None)
    p.add_argument("--balance_quality", type=float, default=0.75)
    p.add_argument("--mean_balance_check_period", type=float, default=120.0)
    p.add_argument("--network_bandwidth_mbps", type=float, default=None)
    p.add_argument("--mean_block_selection_delay", type=float, default=5.0,
                   help="load-balanced server: mean of the random wait before choosing blocks (join-race avoidance)")
    # additions (not in the reference)
    p.add_argument("--repetition_penalty", type=float, default=None, help="override the server default 1.5 (B8)")
    p.add_argument("--seed", type=int, default=0, help="seed of the random-init weights when no checkpoint is local")
    p.add_argument("--kv_capacity", type=int, default=None, help="per-session KV capacity in tokens")
    p.add_argument("--push", action="store_true", help="servers forward activations to the next hop themselves")
    p.add_argument("--stream", action="store_true",
                   help="client: one rpc_inference stream per session and server (cache allocated once per stream)")
    p.add_argument("--attn_cache_tokens", type=int, default=None,
                   help="server: KV budget in tokens over all sessions (default: the upstream rule, 16384 for GQA "
                        "models and 4096 otherwise, never less than one session); 0 = unlimited")
    p.add_argument("--alloc_timeout", type=float, default=10.0,
                   help="server: how long a new session may wait for KV cache bytes before AllocationFailed")
    p.add_argument("--session_timeout", type=float, default=30 * 60.0, help="server: lifetime of an inference stream")
    p.add_argument("--step_timeout", type=float, default=5 * 60.0, help="server: idle time of an inference stream")
    p.add_argument("--max_batch", type=int, default=16,
                   help="server: single-token requests of different sessions merged into one decode step (1 = off)")
    p.add_argument("--weight_format", type=str, default="native", choices=["native", "fp8", "mxfp8"],
                   help="server: weights of the span as loaded (fp16/bf16), e4m3 with one scale per output row, or "
                        "OCP-MX e4m3 with one UE8M0 scale per 32 elements (upstream: --quant_type)")
    p.add_argument("--batch_linger_ms", type=float, default=2.0,
                   help="server: how long a batched decode step waits for the other sessions that have been stepping "
                        "lately (0 = take what is queued)")
    p.add_argument("--no_throughput_cache", action="store_true", help="server: re-measure the throughput on every start")
    p.add_argument("--device_handoff", action="store_true",
                   help="client: servers on GPUs of this machine exchange hidden states device to device (CUDA IPC + "
                        "NVLink peer copies); the RPC messages to them carry a descriptor instead of the tensor")
    p.add_argument("--device_handoff_min_kb", type=int, default=128,
                   help="client: smallest payload that goes device to device (single tokens are cheaper in the frame)")
    p.add_argument("--no_repeat_stop", action="store_true",
                   help="client: keep generating when one token repeats 5 times (random-weight benchmarks)")
    p.add_argument("--gpu", type=int, default=None, help="CUDA device index of this process (default: LOCAL_RANK or 0)")
    p.add_argument("--inbox", action="store_true",
                   help="in-box mode: launch with torchrun (one process per GPU); stages hand hidden states over "
                        "NCCL p2p instead of TCP RPC; --splits gives the stage boundaries, --stage is ignored")
    p.add_argument("--tp", type=int, default=1,
                   help="tensor-parallel width: with --inbox of every pipeline stage; for a served stage (--stage k or "
                        "--use_load_balancing) the span runs on this many GPUs - launch with torchrun "
                        "--nproc-per-node <tp>, rank 0 serves, the other ranks hold the other shards")
    return p


def main(argv=None):
    args = build_parser().parse_args(argv)
    local_rank = int(os.environ.get("LOCAL_RANK", "0")) if args.gpu is None else args.gpu
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank)
        device = torch.device("cuda", local_rank)
    else:
        device = torch.device("cpu")
    logger.info("Using device: %s", device)
    args.dtype = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[args.dtype]
    if device.type == "cpu" and args.dtype != torch.float32:
        logger.info("CPU device: computing in fp32")
        args.dtype = torch.float32
    splits = parse_splits(args.splits)
    if args.inbox:
        return run_inbox(args, device, splits)
    if args.stage == 0:
        run_rank0(args, device, splits)
    else:
        run_stage_server(args, device, splits)


if __name__ == "__main__":
    main()

"""Server throughput for load balancing (reference: src/throughput_measurement.py:15-263).

``throughput = min(compute_rps, network_rps * (1 - overhead))`` with the reference's fallbacks
(network only, then 10.0).  What is measured differs:

* compute: the reference times 10 forwards of a random ``[1,1,H]`` input with no KV cache.  Here
  the real decode step is timed — prefill a scratch session to a representative KV length, then
  CUDA-event-time (or perf_counter on CPU) single-token steps against that cache;
* network: the reference assumes 100 Mbit/s.  Inside a box the hop is NVLink: when
  ``torch.distributed`` is initialised a p2p ping-pong of one hidden state is timed; otherwise
  the given ``--network_bandwidth_mbps`` (default 100, as the reference) is used.

* the announced number follows upstream (petals/server/throughput.py:96-103): per-block
  ``inference_rps`` / ``forward_rps``, the ``(n+1)/2`` average-blocks-used correction, a relay penalty.

A file-locked JSON cache keyed by model/device/dtype (upstream petals/server/throughput.py:55-92)
avoids re-measuring on every restart; it is on by default.
"""
from __future__ import annotations

import fcntl
import json
import logging
import os
import time
from typing import Dict, Optional

import torch

logger = logging.getLogger(__name__)

DEFAULT_NETWORK_MBPS = 100.0
NETWORK_OVERHEAD = 0.2
FALLBACK_RPS = 10.0
CACHE_PATH = os.path.join(os.path.expanduser(os.environ.get("PETALS_B200_CACHE", "~/.cache/petals_b200")),
                          "throughput.json")


def measure_compute_throughput(stage_model, device, dtype: torch.dtype = torch.float16, hidden_size: Optional[int] = None,
                               num_warmup: int = 2, num_iterations: int = 10, kv_length: int = 128) -> float:
    """Decode steps per second of this stage against a ``kv_length``-token session cache."""
    cfg = stage_model.cfg
    H = hidden_size or cfg.hidden
    dev = torch.device(device)
    mgr = stage_model.new_kv_manager()
    kv_length = max(1, min(kv_length, stage_model.kv_capacity - num_warmup - num_iterations - 1))
    kv = mgr.allocate("__throughput_probe__", stage_model.kv_capacity)
    is_first = stage_model.role == "stage0"

    def make(T):
        if is_first:
            return torch.randint(0, cfg.vocab, (1, T), device=dev)
        return (torch.randn(1, T, H, device=dev) * 0.1).to(dtype)

    with torch.inference_mode():
        stage_model(make(kv_length), None, None, kv)
        for _ in range(num_warmup):
            stage_model(make(1), None, None, kv)
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(num_iterations):
                stage_model(make(1), None, None, kv)
            e1.record()
            torch.cuda.synchronize(dev)
            elapsed = e0.elapsed_time(e1) / 1e3
        else:
            t0 = time.perf_counter()
            for _ in range(num_iterations):
                stage_model(make(1), None, None, kv)
            elapsed = time.perf_counter() - t0
    mgr.close(kv.session_id)
    rps = num_iterations / max(elapsed, 1e-9)
    logger.info("Compute throughput: %.2f rps (%d decode steps, KV length %d)", rps, num_iterations, kv_length)
    return rps


def measure_p2p_bandwidth_mbps(hidden_size: int, dtype: torch.dtype, device, iters: int = 20) -> Optional[float]:
    """NCCL/gloo ping-pong of one hidden state with the neighbour rank (None if not distributed)."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
        return None
    rank, world = dist.get_rank(), dist.get_world_size()
    peer = rank ^ 1
    if peer >= world:
        return None
    t = torch.zeros(256, hidden_size, dtype=dtype, device=device)   # a 256-token prefill chunk
    nbytes = t.numel() * t.element_size()

    def once():
        if rank < peer:
            dist.send(t, peer); dist.recv(t, peer)
        else:
            dist.recv(t, peer); dist.send(t, peer)

    for _ in range(3):
        once()
    if t.is_cuda:
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        once()
    if t.is_cuda:
        torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / (2 * iters)
    return nbytes * 8 / dt / 1e6


def estimate_network_throughput(hidden_size: int, dtype: torch.dtype = torch.float16,
                                bandwidth_mbps: Optional[float] = None) -> float:
    """Requests/s the link sustains for one ``[1,1,H]`` hidden state per request."""
    mbps = DEFAULT_NETWORK_MBPS if bandwidth_mbps is None else float(bandwidth_mbps)
    bits = hidden_size * torch.empty((), dtype=dtype).element_size() * 8
    return mbps * 1e6 / bits


def _cache_key(model_name: str, device, dtype, start: int, end: int) -> str:
    name = torch.cuda.get_device_name(device) if torch.device(device).type == "cuda" else "cpu"
    return f"{model_name}|{name}|{dtype}|{start}:{end}"


def _cache_load() -> Dict[str, float]:
    try:
        with open(CACHE_PATH) as f:
            fcntl.flock(f, fcntl.LOCK_SH)
            return json.load(f)
    except (OSError, ValueError):
        return {}


def _cache_store(key: str, value: float) -> None:
    try:
        os.makedirs(os.path.dirname(CACHE_PATH), exist_ok=True)
        with open(CACHE_PATH, "a+") as f:
            fcntl.flock(f, fcntl.LOCK_EX)
            f.seek(0)
            try:
                data = json.load(f)
            except ValueError:
                data = {}
            data[key] = value
            f.seek(0)
            f.truncate()
            json.dump(data, f)
    except OSError:
        pass


def measure_forward_throughput(stage_model, device, dtype: torch.dtype = torch.float16, hidden_size: Optional[int] = None,
                               n_tokens: int = 256, num_iterations: int = 3) -> float:
    """Tokens per second of this stage on ``n_tokens``-token forward passes (upstream ``forward_rps``:
    petals/server/throughput.py:120-129 times 1024-token forwards without a cache)."""
    cfg = stage_model.cfg
    H = hidden_size or cfg.hidden
    dev = torch.device(device)
    n_tokens = max(1, min(n_tokens, stage_model.kv_capacity - 1))
    mgr = stage_model.new_kv_manager()
    kv = mgr.allocate("__forward_probe__", stage_model.kv_capacity)
    is_first = stage_model.role == "stage0"
    x = torch.randint(0, cfg.vocab, (1, n_tokens), device=dev) if is_first else \
        (torch.randn(1, n_tokens, H, device=dev) * 0.1).to(dtype)
    pos = torch.zeros(1, 1, dtype=torch.long)
    with torch.inference_mode():
        stage_model(x, pos, None, kv)                       # warm-up (plans, graphs, allocator)
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for _ in range(num_iterations):
            stage_model(x, pos, None, kv)                   # position 0 again: the cache is rolled back, not grown
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        elapsed = time.perf_counter() - t0
    mgr.close(kv.session_id)
    return num_iterations * n_tokens / max(elapsed, 1e-9)


def measure_loopback_rpc_rps(hidden_size: int, dtype: torch.dtype = torch.float16, iters: int = 200) -> Optional[float]:
    """Requests/s of the framed-TCP data plane for one ``[1,1,H]`` hidden state, measured against a scratch echo
    server on 127.0.0.1: what a hop between two servers of ONE box really costs, instead of assuming a link speed."""
    try:
        from petals_b200.net.rpc import RpcClient, RpcServer

        srv = RpcServer("127.0.0.1", 0).start()
        srv.add_handler("echo", lambda uid, tensors, md: (tensors, md))
        c = RpcClient()
        x = torch.zeros(1, 1, hidden_size, dtype=dtype)
        for _ in range(10):
            c.call(srv.address, "echo", "", [x], {})
        t0 = time.perf_counter()
        for _ in range(iters):
            c.call(srv.address, "echo", "", [x], {})
        dt = time.perf_counter() - t0
        c.close()
        srv.stop()
        return iters / max(dt, 1e-9)
    except Exception as e:  # noqa: BLE001
        logger.warning("loopback RPC measurement failed: %s", e)
        return None


def get_server_throughput_info(stage_model, device, dtype: torch.dtype = torch.float16, hidden_size: Optional[int] = None,
                               network_bandwidth_mbps: Optional[float] = None, use_cache: bool = True,
                               model_name: str = "", reachable_via_relay: bool = False, relay_penalty: float = 0.2,
                               in_box: Optional[bool] = None, **measure_kw) -> Dict[str, float]:
    """The upstream throughput record (petals/server/throughput.py:37-108):

        inference_rps  single-token steps per second PER BLOCK (decode against a KV cache)
        forward_rps    tokens per second PER BLOCK on long forward passes
        network_rps    hidden states per second the link carries
        throughput     min(forward_rps / ((n + 1) / 2), network_rps [* relay_penalty])  - what the load balancer compares

    (n = blocks of this server; a request enters at a uniformly random block of the span, so it uses (n+1)/2 of
    them on average.)  The per-block numbers are cached per model / device / dtype in a file-locked JSON, so a
    re-placed or restarted server does not measure again.  Network: ``--network_bandwidth_mbps`` if given;
    else, between ranks of one torch.distributed job, a measured p2p ping-pong; else the measured requests/s of
    the loopback RPC itself (servers of one box) - the reference's fixed 100 Mbit/s guess (B10) is only the
    last resort."""
    H = hidden_size or stage_model.cfg.hidden
    n = max(1, stage_model.end - stage_model.start)
    key = _cache_key(model_name, device, dtype, 0, 0).rsplit("|", 1)[0] + "|per_block"
    per_block = None
    if use_cache:
        hit = _cache_load().get(key)
        if isinstance(hit, dict) and hit.get("inference_rps") and hit.get("forward_rps"):
            per_block = hit
    if per_block is None:
        per_block = {}
        try:
            per_block["inference_rps"] = measure_compute_throughput(stage_model, device, dtype, H, **measure_kw) * n
        except Exception as e:  # noqa: BLE001
            logger.warning("compute throughput measurement failed: %s", e)
        try:
            per_block["forward_rps"] = measure_forward_throughput(stage_model, device, dtype, H) * n
        except Exception as e:  # noqa: BLE001
            logger.warning("forward throughput measurement failed: %s", e)
        if use_cache and per_block.get("inference_rps") and per_block.get("forward_rps"):
            _cache_store(key, per_block)
    info: Dict[str, float] = dict(per_block)
    network = None
    if network_bandwidth_mbps is not None:
        network = estimate_network_throughput(H, dtype, network_bandwidth_mbps) * (1.0 - NETWORK_OVERHEAD)
    else:
        mbps = None
        try:
            mbps = measure_p2p_bandwidth_mbps(H, dtype, device)
        except Exception:  # noqa: BLE001
            mbps = None
        if mbps is not None:
            network = estimate_network_throughput(H, dtype, mbps) * (1.0 - NETWORK_OVERHEAD)
        elif in_box is None or in_box:
            network = measure_loopback_rpc_rps(H, dtype)
        if network is None:
            network = estimate_network_throughput(H, dtype, None) * (1.0 - NETWORK_OVERHEAD)
    if reachable_via_relay:
        network *= relay_penalty
    info["network_rps"] = network
    fwd = info.get("forward_rps") or info.get("inference_rps")
    if fwd:
        thr = min(fwd / ((n + 1) / 2), network) if network and network > 0 else fwd / ((n + 1) / 2)
    elif network and network > 0:
        thr = network
    else:
        thr = FALLBACK_RPS
    info["throughput"] = thr
    logger.info("Server throughput: %.2f rps for %d blocks (inference %.1f, forward %.1f per block; network %.1f)", thr, n,
                info.get("inference_rps") or 0.0, info.get("forward_rps") or 0.0, network or 0.0)
    return info


def get_server_throughput(stage_model, device, dtype: torch.dtype = torch.float16, hidden_size: Optional[int] = None,
                          network_bandwidth_mbps: Optional[float] = None, use_cache: bool = True,
                          model_name: str = "", **kw) -> float:
    """The one number the load balancer compares (see :func:`get_server_throughput_info`)."""
    return float(get_server_throughput_info(stage_model, device, dtype, hidden_size, network_bandwidth_mbps, use_cache,
                                            model_name, **kw)["throughput"])

"""Server throughput for load balancing (reference: src/throughput_measurement.py:15-263).

``throughput = min(compute_rps, network_rps * (1 - overhead))`` with the reference's fallbacks
(network only, then 10.0).  What is measured differs:

* compute: the reference times 10 forwards of a random ``[1,1,H]`` input with no KV cache.  Here
  the real decode step is timed — prefill a scratch session to a representative KV length, then
  CUDA-event-time (or perf_counter on CPU) single-token steps against that cache;
* network: the reference assumes 100 Mbit/s.  Inside a box the hop is NVLink: when
  ``torch.distributed`` is initialised a p2p ping-pong of one hidden state is timed; otherwise
  the given ``--network_bandwidth_mbps`` (default 100, as the reference) is used.

A file-locked JSON cache keyed by model/device/dtype/span (upstream petals/server/throughput.py:55-92)
avoids re-measuring on every restart.
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


def get_server_throughput(stage_model, device, dtype: torch.dtype = torch.float16, hidden_size: Optional[int] = None,
                          network_bandwidth_mbps: Optional[float] = None, use_cache: bool = False,
                          model_name: str = "", **measure_kw) -> float:
    H = hidden_size or stage_model.cfg.hidden
    key = _cache_key(model_name, device, dtype, stage_model.start, stage_model.end)
    if use_cache:
        hit = _cache_load().get(key)
        if hit:
            return float(hit)
    compute = None
    try:
        compute = measure_compute_throughput(stage_model, device, dtype, H, **measure_kw)
    except Exception as e:  # noqa: BLE001
        logger.warning("compute throughput measurement failed: %s", e)
    mbps = network_bandwidth_mbps
    if mbps is None:
        try:
            mbps = measure_p2p_bandwidth_mbps(H, dtype, device)
        except Exception:  # noqa: BLE001
            mbps = None
    network = estimate_network_throughput(H, dtype, mbps) * (1.0 - NETWORK_OVERHEAD)
    if compute is not None and compute > 0:
        thr = min(compute, network)
    elif network > 0:
        thr = network
    else:
        thr = FALLBACK_RPS
    logger.info("Server throughput: %.2f rps (compute=%s, network=%.2f)", thr,
                f"{compute:.2f}" if compute else "n/a", network)
    if use_cache:
        _cache_store(key, thr)
    return thr

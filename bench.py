#!/usr/bin/env python
"""Headline benchmark: decode tokens/s + p50 per-token latency, Llama-2-7B fp16, N pipeline stages.

    python bench.py --gpus 1 --steps 64 --warmup 8
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      (the unmodified reference, see baseline/)

One "step" = one new token for every in-flight session.  With N stages the engine keeps N
sessions in flight (one per stage, batch 1 each) so the per-GPU work per step is constant
("weak" scaling); the single-session per-token latency is reported next to the aggregate.

* ``value``      device-timed (CUDA events, max over ranks): K steps replayed back to back with
                 token feedback on the device (no host in the loop).
* ``e2e.value``  the same K steps through the public ``PipelineEngine.decode_round`` API: every
                 step copies the previous tokens host->device from pinned memory and reads the new
                 tokens device->host.
Weights are random-init (no network), prompts synthetic; every step streams all 13 GB of
weights, far more than the 126 MB L2, so no L2 flush is needed between steps.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama-2-7b")
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "bf16"])
    ap.add_argument("--prompt-len", type=int, default=128)
    ap.add_argument("--sessions", type=int, default=0, help="in-flight sessions (default: one per stage)")
    ap.add_argument("--tp", type=int, default=1, help="tensor-parallel width of every stage (gpus = stages * tp)")
    ap.add_argument("--tp-mode", default="fused", choices=["fused", "nccl"])
    ap.add_argument("--group", type=int, default=1,
                    help="sessions per micro-batch travelling through the chain together (serving-throughput mode)")
    ap.add_argument("--weights", default="native", choices=["native", "fp8", "mxfp8"],
                    help="fp8 / mxfp8 = opt-in e4m3 weights with per-row / OCP-MX block scales (NOT the headline configuration)")
    ap.add_argument("--hop", default=None, choices=["p2p", "nccl"],
                    help="decode hop between stages: in-graph peer stores (default) or host-issued NCCL send/recv")
    ap.add_argument("--head-split", type=int, default=None, help="1/0: shard the LM head over the last and first stage")
    ap.add_argument("--no-ttft", action="store_true", help="skip the time-to-first-token measurement")
    return ap.parse_args()


def measured_peaks():
    """Roofline denominators measured by the driver on this pool (fallback: the profiling recipe's figures)."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def even_splits(n_layers: int, n: int):
    return [round(i * n_layers / n) for i in range(1, n)]


def run_b200(args):
    import torch
    import torch.distributed as dist

    from petals_b200.parallel.pipeline import PipelineEngine

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    dtype = torch.float16 if args.dtype == "fp16" else torch.bfloat16
    from petals_b200.models.config import resolve_config

    cfg = resolve_config(args.model)
    stages = world // args.tp
    S = args.sessions or stages
    K, W, L = args.steps, args.warmup, args.prompt_len
    cap = L + 2 * (K + W) + 16
    ttft_lens = [] if args.no_ttft else [128, 2048]
    if ttft_lens:
        cap = max(cap, max(ttft_lens) + 64)
    cap = (cap + 255) // 256 * 256
    eng = PipelineEngine(args.model, even_splits(cfg.n_layers, stages), dtype, dev, kv_capacity=cap,
                         max_sessions=max(S, 1), max_chunk=2048 if stages == 1 else 512, seed=0, weight_format=args.weights, tp=args.tp,
                         tp_mode=args.tp_mode, hop=args.hop,
                         head_split=None if args.head_split is None else bool(args.head_split))
    sess = eng.open_sessions(S)
    g = torch.Generator().manual_seed(1234)
    prompts = [torch.randint(0, cfg.vocab, (L,), generator=g) for _ in range(S)]
    cur = [eng.prefill(i, p) for i, p in zip(sess, prompts)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------------------------------------------------------- device-timed loop
    if eng.is_first:
        eng._tok[:S].copy_(torch.tensor(cur, dtype=torch.int64))
    native = eng.span
    single = world == 1

    def device_steps(n):
        if single:
            native.decode_loop([eng.sessions[i] for i in sess], eng._tok[:S], n)
            eng._tok[:S].copy_(native.ids_in[:S])
        else:
            for _ in range(n):
                eng.decode_round(None, to_host=False, group=args.group)

    device_steps(W)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = native.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    device_steps(K)
    e1.record()
    barrier()
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    launches = native.kernel_launches - launches0
    if world > 1:
        t = torch.tensor([launches], device=dev, dtype=torch.int64)
        dist.all_reduce(t)
        launches = int(t.item())

    # ---------------------------------------------------------------- end-to-end through the public API
    eng.drain()
    barrier()
    if eng.is_first:
        eng._tok_host[:S].copy_(eng._tok[:S])
        cur = [int(x) for x in eng._tok_host[:S]]
    # streaming rounds: every session's token goes device -> host -> device each step, but a session's
    # next step is issued as soon as ITS token is back (no barrier between rounds)
    for _ in range(3):
        r = eng.decode_round_streaming(cur)
        cur = r if r else cur
    barrier()
    h2d0, d2h0 = eng.h2d_bytes, eng.d2h_bytes
    lat = []
    toks0 = []                    # session 0's tokens of the end-to-end phase (checksum: must not depend on N)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e2.record()
    t_prev = time.perf_counter()
    for _ in range(K):
        r = eng.decode_round_streaming(cur)
        cur = r if r else cur
        if r:
            toks0.append(int(r[0]))
        now = time.perf_counter()
        lat.append(now - t_prev)
        t_prev = now
    r = eng.flush_streaming()
    if r:
        toks0.append(int(r[0]))
    e3.record()
    barrier()
    e2e_ms = max_over_ranks(e2.elapsed_time(e3))
    clocks = sampler.stop() if rank == 0 else None

    # ---------------------------------------------------------------- time to first token (public prefill API)
    ttft = {}
    if ttft_lens:
        extra = eng.open_sessions(1)[0]
        gt = torch.Generator().manual_seed(4321)
        for plen in ttft_lens:
            prompt = torch.randint(0, cfg.vocab, (plen,), generator=gt).pin_memory()
            times = []
            for _ in range(4):
                eng.sessions[extra].length = 0
                barrier()
                t0 = time.perf_counter()
                eng.prefill(extra, prompt)          # H2D of the prompt ... D2H of the first token (rank 0)
                times.append((time.perf_counter() - t0) * 1e3)
            ttft[str(plen)] = round(statistics.median(times[1:]), 3)
    if rank == 0:
        import hashlib

        hbm_gbs, peak_src = measured_peaks()
        tok_s = S * K / (dev_ms / 1e3)
        e2e_tok_s = S * K / (e2e_ms / 1e3)
        wbytes = cfg.layer_bytes(2) * cfg.n_layers + cfg.vocab * cfg.hidden * 2
        out = {
            "metric": "decode tokens/sec (aggregate over in-flight sessions), "
                      + ("Llama-2-7B" if args.model == "llama-2-7b" else args.model) + " N-stage pipeline",
            "value": round(tok_s, 2), "unit": "tokens/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(dev_ms / K, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype if args.weights == "native" else f"{args.dtype} activations, {args.weights} (e4m3) weights (opt-in)",
            "data": "synthetic prompts, random-init weights",
            "impl": "b200",
            "config": {"model": args.model, "global_batch": S, "seq_len": L, "parallelism": f"pp{stages}" + (f"tp{args.tp}({args.tp_mode})" if args.tp > 1 else ""),
                       "sessions_in_flight": S, "batch_per_session": 1, "sessions_per_micro_batch": args.group,
                       "l2": "every step streams all weights (13.5 GB >> 126 MB L2); no flush needed",
                       "p50_token_latency_ms": round(statistics.median(lat) * 1e3, 4),
                       "single_session_tokens_per_s": round(1.0 / statistics.median(lat), 2),
                       "hop": eng.hop, "head_split": eng.head_split,
                       "ttft_ms": ttft,
                       "tokens_sha": hashlib.sha256(",".join(str(t) for t in toks0[-K:]).encode()).hexdigest()[:16],
                       f"hbm_roofline_frac_of_{peak_src}": round(
                           (S * wbytes / world) / (dev_ms / K / 1e3) / (hbm_gbs * 1e9), 3)},
            "clocks": clocks,
            "e2e": {"value": round(e2e_tok_s, 2), "unit": "tokens/s",
                    "h2d_bytes_per_step": (eng.h2d_bytes - h2d0) // K, "d2h_bytes_per_step": (eng.d2h_bytes - d2h0) // K,
                    "ms_per_step": round(e2e_ms / K, 4)},
            "gpu_launches": launches,
        }
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        from baseline.reference_arm import run_reference

        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()

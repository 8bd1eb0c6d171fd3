"""Serving throughput over the TCP data plane: one stage server holding a whole model, N concurrent client sessions.

    python scripts/bench_served.py [--model llama-2-7b] [--clients 1,4,16] [--steps 64]

The server runs in a subprocess (`python -m src.main --stage 1` with --splits 0-style span [0, L)); every client is a
process with its own session that sends hidden states (a fixed pseudo-embedding of the previous token, so the loop is
closed and deterministic) and receives tokens.  Requests of different sessions that are queued at the same time are
merged into one batched decode step by the server (src/rpc_handler.py).  Prints aggregate tokens/s per client count,
the mean batch the server formed, and checks that every session's tokens equal the ones it produced alone.
"""
import argparse
import json
import os
import socket
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


SERVER = r"""
import sys, torch, threading
sys.path.insert(0, %(root)r)
from src.main import build_parser, start_stage_server
from src.partition import StageLast, load_stage_model
from petals_b200.models.config import resolve_config
a = build_parser().parse_args(["--model", %(model)r, "--splits", "0", "--stage", "1", "--dht_port", "0", "--rpc_port", %(port)r,
                               "--dtype", "fp16", "--temperature", "0", "--max_batch", %(max_batch)r, "--attn_cache_tokens", "0",
                               "--batch_linger_ms", %(linger)r])
gpu = torch.cuda.is_available()                       # (CPU: only for trying the script out)
a.dtype = torch.float16 if gpu else torch.float32
a.kv_capacity = 512
dev = torch.device("cuda", 0) if gpu else torch.device("cpu")
if gpu:
    torch.cuda.set_device(dev)
cfg = resolve_config(%(model)r)
full = load_stage_model(%(model)r, dev, role="last", start=0, end=cfg.n_layers, dtype=a.dtype, seed=0)
stage = StageLast(full, 0, gpu_device=dev, kv_capacity=512, max_batch=int(%(max_batch)r))
h = start_stage_server(a, dev, stage, True, stage_index=1)
print("READY", flush=True)
h.wait()
"""


def _table(hidden):
    g = torch.Generator().manual_seed(0)
    return (torch.randn(1024, hidden, generator=g) * 0.5).to(torch.float16)     # pseudo-embedding


def session(i, steps, port, hidden, prompt_len, out, barrier=None):
    """One client session (its own process when run concurrently): prefill, then `steps` closed-loop decode steps."""
    sys.path.insert(0, ROOT)
    from petals_b200.net.rpc import RpcClient

    torch.set_num_threads(1)
    addr = ("127.0.0.1", port)
    table = _table(hidden)
    c = RpcClient()
    sid = f"bench-{i}-{time.time_ns()}"
    gi = torch.Generator().manual_seed(100 + i)
    prompt = table[torch.randint(0, 1024, (prompt_len,), generator=gi)].unsqueeze(0)
    md = {"session_id": sid, "seq_len": prompt_len, "cur_len": prompt_len, "is_prefill": True, "max_length": 512,
          "temperature": 0.0}
    _, r = c.call(addr, "StageConnectionHandler.rpc_forward", "k", [prompt], md, timeout=120)
    tok, cur, toks = r["token_id"], prompt_len, []
    if barrier is not None:
        barrier.wait()
    t0 = time.perf_counter()
    for _ in range(steps):
        cur += 1
        md = {"session_id": sid, "seq_len": 1, "cur_len": cur, "is_prefill": False, "max_length": 512, "temperature": 0.0}
        _, r = c.call(addr, "StageConnectionHandler.rpc_forward", "k", [table[tok % 1024].view(1, 1, -1)], md, timeout=120)
        tok = r["token_id"]
        toks.append(tok)
    dt = time.perf_counter() - t0
    if barrier is not None:
        barrier.wait()
    c.call(addr, "StageConnectionHandler.rpc_close_session", "k", [], {"session_id": sid})
    c.close()
    out[i] = (toks, dt)


def main():
    import torch.multiprocessing as mp

    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-2-7b")
    ap.add_argument("--clients", default="1,4,16")
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--prompt", type=int, default=64)
    ap.add_argument("--linger_ms", type=float, default=2.0)
    ap.add_argument("--out", default="gpurun_out/bench_served.json")
    args = ap.parse_args()
    from petals_b200.models.config import resolve_config
    from petals_b200.net.rpc import RpcClient

    cfg = resolve_config(args.model)
    counts = [int(c) for c in args.clients.split(",")]
    port = free_port()
    code = SERVER % {"root": ROOT, "model": args.model, "port": str(port), "max_batch": str(max(counts)),
                     "linger": str(args.linger_ms)}
    srv = subprocess.Popen([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    for line in srv.stdout:
        if line.startswith("READY"):
            break
    threading.Thread(target=lambda: [None for _ in srv.stdout], daemon=True).start()
    addr = ("127.0.0.1", port)
    ctx = mp.get_context("spawn")
    result = {"model": args.model, "steps": args.steps, "batch_linger_ms": args.linger_ms,
              "clients_are": "one process per session", "runs": []}
    try:
        alone = {}
        for i in range(max(counts)):                      # every session alone: the oracle (and the warm-up)
            session(i, args.steps if i < 2 else 8, port, cfg.hidden, args.prompt, alone)
        seen_b = seen_i = 0
        for n in counts:
            c = RpcClient()
            _, info = c.call(addr, "StageConnectionHandler.rpc_info", "k", [], {})
            c.close()
            seen_b, seen_i = info.get("batches", 0), info.get("mean_batch", 0.0) * info.get("batches", 0)
            out, barrier = ctx.Manager().dict(), ctx.Barrier(n + 1)
            ps = [ctx.Process(target=session, args=(i, args.steps, port, cfg.hidden, args.prompt, out, barrier))
                  for i in range(n)]
            for p in ps:
                p.start()
            barrier.wait()                                # every session has done its prefill
            t0 = time.perf_counter()
            barrier.wait()                                # ... and its last step
            dt = time.perf_counter() - t0
            for p in ps:
                p.join()
            c = RpcClient()
            _, info = c.call(addr, "StageConnectionHandler.rpc_info", "k", [], {})
            c.close()
            same = all(out[i][0][:len(alone[i][0])] == alone[i][0][:len(out[i][0])] for i in range(n))
            run = {"clients": n, "tokens_per_s": round(n * args.steps / dt, 1), "ms_per_round": round(dt / args.steps * 1e3, 3),
                   "server_mean_batch": round((info.get("mean_batch", 0.0) * info.get("batches", 0) - seen_i)
                                              / max(info.get("batches", 0) - seen_b, 1), 2),
                   "server_batch_ms_cumulative_mean": info.get("batch_ms"),
                   "server_mean_service_ms_cumulative": info.get("mean_service_ms"),
                   "session_turnaround_ms_cumulative": info.get("mean_turnaround_ms"), "client_cpus": os.cpu_count(), "tokens_equal_to_solo_runs": same}
            print(json.dumps(run), flush=True)
            result["runs"].append(run)
    finally:
        srv.terminate()
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(result, f)


if __name__ == "__main__":
    main()

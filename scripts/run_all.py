#!/usr/bin/env python
"""Single-host launcher: starts every server stage, then the client in the foreground.

Reference: scripts/run_all.py:1-566 (stage1 first, scrape its DHT multiaddr from the log, wait
for "handlers registered", then stage2 (+2 ports), stage3 (+4), client (+6); per-stage
``stageN.log``; SIGINT cleanup).  Differences: any number of stages (from ``--splits``), one GPU
per stage via ``LOCAL_RANK`` when GPUs are present, and defaults that actually run
(``--model gpt2 --splits 6``; the reference's default ``10,20,30`` is empty for 12-layer GPT-2).

    python scripts/run_all.py --model gpt2 --splits 6 --max_new_tokens 16
    python scripts/run_all.py --model llama-2-7b --splits 8,16,24 --dtype fp16
"""
import argparse
import os
import re
import signal
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MADDR_RE = re.compile(r"DHT visible multiaddrs \(use for --dht_initial_peers\): \['([^']+)'")


def wait_for(path, pattern, timeout):
    t0 = time.time()
    rx = re.compile(pattern) if isinstance(pattern, str) else pattern
    while time.time() - t0 < timeout:
        if os.path.exists(path):
            m = rx.search(open(path, errors="replace").read())
            if m:
                return m
        time.sleep(0.2)
    raise TimeoutError(f"{pattern} not seen in {path} after {timeout}s")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="gpt2")
    ap.add_argument("--splits", default="6")
    ap.add_argument("--dtype", default="fp16")
    ap.add_argument("--prompt", default="Hello, how are you?")
    ap.add_argument("--max_new_tokens", type=int, default=32)
    ap.add_argument("--temperature", type=float, default=0.0)
    ap.add_argument("--base_port", type=int, default=8000)
    ap.add_argument("--log_dir", default=".")
    ap.add_argument("--startup_timeout", type=float, default=600.0)
    ap.add_argument("--extra", nargs=argparse.REMAINDER, default=[], help="extra flags passed to every process")
    a = ap.parse_args()
    n_stages = len([s for s in a.splits.split(",") if s.strip()])
    try:
        import torch

        n_gpu = torch.cuda.device_count()
    except Exception:  # noqa: BLE001
        n_gpu = 0
    procs = []

    def cleanup(*_):
        for p in procs:
            if p.poll() is None:
                p.terminate()
        for p in procs:
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                p.kill()

    signal.signal(signal.SIGINT, lambda *_: (cleanup(), sys.exit(130)))
    peers = ""
    try:
        for st in range(1, n_stages + 1):
            log = os.path.join(a.log_dir, f"stage{st}.log")
            if os.path.exists(log):
                os.remove(log)
            env = dict(os.environ, PYTHONUNBUFFERED="1")
            if n_gpu:
                env["LOCAL_RANK"] = str(st % n_gpu)
            cmd = [sys.executable, "-m", "src.main", "--model", a.model, "--splits", a.splits, "--stage", str(st),
                   "--dtype", a.dtype, "--dht_port", str(a.base_port + 2 * (st - 1)),
                   "--rpc_port", str(a.base_port + 2 * (st - 1) + 1), "--dht_initial_peers", peers] + a.extra
            print("starting:", " ".join(cmd), "->", log, flush=True)
            procs.append(subprocess.Popen(cmd, cwd=ROOT, env=env, stdout=open(log, "w"), stderr=subprocess.STDOUT))
            if st == 1:
                peers = wait_for(log, MADDR_RE, a.startup_timeout).group(1)
                print("DHT bootstrap:", peers, flush=True)
            wait_for(log, rf"Stage{st} handlers registered", a.startup_timeout)
        env = dict(os.environ, PYTHONUNBUFFERED="1")
        if n_gpu:
            env["LOCAL_RANK"] = "0"
        cmd = [sys.executable, "-m", "src.main", "--model", a.model, "--splits", a.splits, "--stage", "0",
               "--dtype", a.dtype, "--dht_port", str(a.base_port + 2 * n_stages),
               "--rpc_port", str(a.base_port + 2 * n_stages + 1), "--dht_initial_peers", peers,
               "--prompt", a.prompt, "--max_new_tokens", str(a.max_new_tokens), "--temperature", str(a.temperature)]
        rc = subprocess.call(cmd + a.extra, cwd=ROOT, env=env)
    finally:
        cleanup()
    sys.exit(rc)


if __name__ == "__main__":
    main()

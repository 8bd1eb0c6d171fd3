#!/usr/bin/env python
"""Fault-tolerance demo (reference: scripts/test_fault_tolerance.py:24-93, which prints
instructions and waits for a human to run kill_stage.py).  Non-interactive here: starts two
replicas of every server stage as sub-processes, runs the client, SIGTERMs the replica in use
part-way through, and checks that the continuation equals an undisturbed run token for token.

    python scripts/test_fault_tolerance.py --model tiny-llama-gqa --splits 1,3 --max_new_tokens 24
"""
import argparse
import os
import re
import signal
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="tiny-llama-gqa")
    ap.add_argument("--splits", default="1,3")
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--max_new_tokens", type=int, default=24)
    ap.add_argument("--kill_stage", type=int, default=0, help="stage whose active replica dies (default: last)")
    ap.add_argument("--kill_after_s", type=float, default=0.0, help="0 = after the 5th request of the client run")
    a = ap.parse_args()
    import torch

    from scripts.run_all import MADDR_RE, wait_for
    from src.main import build_parser, parse_splits, run_rank0

    cuts = parse_splits(a.splits)
    n = len(cuts)
    kill_stage = a.kill_stage or n
    procs, logs, peers = {}, {}, ""
    port = 18000
    try:
        for rep in range(2):
            for st in range(1, n + 1):
                log = f"/tmp/ft_stage{st}_{rep}.log"
                cmd = [sys.executable, "-m", "src.main", "--model", a.model, "--splits", a.splits, "--stage", str(st),
                       "--dtype", a.dtype, "--dht_port", str(port), "--rpc_port", str(port + 1),
                       "--dht_initial_peers", peers]
                port += 2
                procs[(st, rep)] = subprocess.Popen(cmd, cwd=ROOT, stdout=open(log, "w"), stderr=subprocess.STDOUT,
                                                    env=dict(os.environ, PYTHONUNBUFFERED="1"))
                logs[(st, rep)] = log
                if not peers:
                    peers = wait_for(log, MADDR_RE, 300).group(1)
                wait_for(log, rf"Stage{st} handlers registered", 300)

        def client(push=False):
            args = build_parser().parse_args(["--model", a.model, "--splits", a.splits, "--stage", "0", "--dtype", a.dtype,
                                              "--dht_initial_peers", peers, "--temperature", "0",
                                              "--max_new_tokens", str(a.max_new_tokens), "--request_timeout", "5"])
            args.dtype = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[a.dtype]
            dev = torch.device("cuda", 0) if torch.cuda.is_available() and a.dtype != "fp32" else torch.device("cpu")
            return run_rank0(args, dev, cuts)

        ref, _ = client()
        print("undisturbed run:", ref)

        def killer():
            # the replica whose log grows is the one the client picked
            sizes = {k: os.path.getsize(v) for k, v in logs.items() if k[0] == kill_stage}
            t0 = time.time()
            while time.time() - t0 < 60:
                time.sleep(a.kill_after_s or 0.05)
                for k in sizes:
                    if procs[k].poll() is None and (a.kill_after_s or os.path.getsize(logs[k]) > sizes[k]):
                        procs[k].send_signal(signal.SIGTERM)
                        print(f"killed stage {k[0]} replica {k[1]} (pid {procs[k].pid})", flush=True)
                        return

        t = threading.Thread(target=killer, daemon=True)
        t.start()
        out, stats = client()
        t.join(timeout=5)
        print("run with a killed stage:", out, "recoveries:", stats["recoveries"])
        ok = out == ref
        print("RESULT:", "token-exact continuation" if ok else "MISMATCH")
        return 0 if ok else 1
    finally:
        for p in procs.values():
            if p.poll() is None:
                p.terminate()


if __name__ == "__main__":
    sys.exit(main())

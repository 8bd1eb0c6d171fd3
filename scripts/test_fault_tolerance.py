#!/usr/bin/env python
"""Fault-tolerance demo (reference: scripts/test_fault_tolerance.py:24-93, which prints
instructions and waits for a human to run kill_stage.py).  Non-interactive here: starts two
replicas of every server stage as sub-processes, runs the client, SIGTERMs the replica in use
part-way through, and checks that the continuation equals an undisturbed run token for token.

    python scripts/test_fault_tolerance.py --model tiny-llama-gqa --splits 1,3 --max_new_tokens 24

On a GPU box the same script is the measurement of BASELINE config 4 (two replica spans, kill a stage):

    python scripts/test_fault_tolerance.py --model llama-2-13b --splits 10,25 --dtype fp16 --gpus 4 \
        --device_handoff --kill_stage 1 --kill_at_token 40 --max_new_tokens 128 --hard --out gpurun_out/failover.json

puts the client on GPU 0 and the servers round-robin on GPUs 1.., hands hidden states over device to device, SIGKILLs the
stage-1 replica that serves the session when token 40 arrives, and reports tokens/s before and after, the length
of the stall, and whether the tokens equal the undisturbed run.
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
    ap.add_argument("--kill_at_token", type=int, default=0, help="kill when this token arrives (overrides kill_after_s)")
    ap.add_argument("--hard", action="store_true", help="SIGKILL instead of SIGTERM")
    ap.add_argument("--gpus", type=int, default=0, help="spread the servers over GPUs 1..gpus-1 (client on GPU 0)")
    ap.add_argument("--device_handoff", action="store_true")
    ap.add_argument("--device_handoff_min_kb", type=int, default=128)
    ap.add_argument("--prompt", default="")
    ap.add_argument("--kv_capacity", type=int, default=0)
    ap.add_argument("--out", default="")
    ap.add_argument("--port_base", type=int, default=18000)
    ap.add_argument("--request_timeout", type=float, default=5.0)
    a = ap.parse_args()
    import torch

    from scripts.run_all import MADDR_RE, wait_for
    from src.main import build_parser, parse_splits, run_rank0

    cuts = parse_splits(a.splits)
    n = len(cuts)
    kill_stage = a.kill_stage or n
    procs, logs, peers = {}, {}, ""
    port = a.port_base
    try:
        for rep in range(2):
            for st in range(1, n + 1):
                log = f"/tmp/ft_stage{st}_{rep}.log"
                cmd = [sys.executable, "-m", "src.main", "--model", a.model, "--splits", a.splits, "--stage", str(st),
                       "--dtype", a.dtype, "--dht_port", str(port), "--rpc_port", str(port + 1),
                       "--dht_initial_peers", peers]
                if a.gpus > 1:
                    cmd += ["--gpu", str(1 + len(procs) % (a.gpus - 1))]
                if a.kv_capacity:
                    cmd += ["--kv_capacity", str(a.kv_capacity)]
                port += 2
                procs[(st, rep)] = subprocess.Popen(cmd, cwd=ROOT, stdout=open(log, "w"), stderr=subprocess.STDOUT,
                                                    env=dict(os.environ, PYTHONUNBUFFERED="1"))
                logs[(st, rep)] = log
                if not peers:
                    peers = wait_for(log, MADDR_RE, 300).group(1)
                wait_for(log, rf"Stage{st} handlers registered", 300)

        sig = signal.SIGKILL if a.hard else signal.SIGTERM
        stamps = []

        def client(push=False, on_token=None):
            stamps.clear()
            args = build_parser().parse_args(["--model", a.model, "--splits", a.splits, "--stage", "0", "--dtype", a.dtype,
                                              "--dht_initial_peers", peers, "--temperature", "0",
                                              "--max_new_tokens", str(a.max_new_tokens), "--request_timeout", str(a.request_timeout)])
            args.dtype = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[a.dtype]
            dev = torch.device("cuda", 0) if torch.cuda.is_available() and a.dtype != "fp32" else torch.device("cpu")
            args.device_handoff = a.device_handoff
            args.device_handoff_min_kb = a.device_handoff_min_kb
            if a.prompt:
                args.prompt = a.prompt
            args.no_repeat_stop = bool(a.kill_at_token)
            if a.kv_capacity:
                args.kv_capacity = a.kv_capacity

            def observe(i, tok):
                stamps.append(time.perf_counter())
                if on_token is not None:
                    on_token(i, tok)

            args.on_token = observe
            return run_rank0(args, dev, cuts)

        def active_replica():
            """The replica of the stage to kill that has served requests (asked over RPC, not guessed from logs)."""
            from petals_b200.net.rpc import RpcClient
            c = RpcClient()
            best, most = None, 0
            for (st, rep), pr in procs.items():
                if st != kill_stage or pr.poll() is not None:
                    continue
                m = re.search(r"--rpc_port (\d+)", " ".join(pr.args))
                try:
                    _, info = c.call(("127.0.0.1", int(m.group(1))), "StageConnectionHandler.rpc_info", "k", [], {}, timeout=2.0)
                except Exception:  # noqa: BLE001
                    continue
                if info.get("sessions", 0) > most or best is None:
                    best, most = (st, rep), info.get("sessions", 0)
            c.close()
            return best

        ref, ref_stats = client()
        ref_stamps = list(stamps)
        print("undisturbed run:", ref)
        if a.kill_at_token:
            killed = {}

            def kill_now(i, tok):
                if i == a.kill_at_token and not killed:
                    k = active_replica()
                    procs[k].send_signal(sig)
                    killed["who"], killed["t"] = k, time.perf_counter()
                    print(f"killed stage {k[0]} replica {k[1]} (pid {procs[k].pid}) at token {i}", flush=True)

            out, stats = client(on_token=kill_now)
            gaps = [b - x for x, b in zip(stamps, stamps[1:])]
            k = a.kill_at_token
            before, after = gaps[1:k], gaps[k + 2:]
            ref_gaps = [b - x for x, b in zip(ref_stamps, ref_stamps[1:])]
            res = {"model": a.model, "splits": a.splits, "dtype": a.dtype, "device_handoff": a.device_handoff,
                   "kill": {"stage": kill_stage, "at_token": k, "signal": "SIGKILL" if a.hard else "SIGTERM"},
                   "tokens": len(out), "token_exact_vs_undisturbed": out == ref, "recoveries": stats["recoveries"],
                   "undisturbed_tok_per_s": round(len(ref_gaps) / max(sum(ref_gaps), 1e-9), 1),
                   "tok_per_s_before_kill": round(len(before) / max(sum(before), 1e-9), 1),
                   "tok_per_s_after_recovery": round(len(after) / max(sum(after), 1e-9), 1),
                   "stall_ms": round(max(gaps[k - 1:k + 2] or [0.0]) * 1e3, 1),
                   "ttft_ms": round(stats["ttft_s"] * 1e3, 1), "decode_p50_ms": stats.get("decode_p50_ms")}
            import json
            print(json.dumps(res), flush=True)
            if a.out:
                os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
                with open(a.out, "w") as f:
                    json.dump(res, f)
            print("RESULT:", "token-exact continuation" if out == ref else "MISMATCH")
            return 0 if out == ref else 1

        def killer():
            # the replica whose log grows is the one the client picked
            sizes = {k: os.path.getsize(v) for k, v in logs.items() if k[0] == kill_stage}
            t0 = time.time()
            while time.time() - t0 < 60:
                time.sleep(a.kill_after_s or 0.05)
                for k in sizes:
                    if procs[k].poll() is None and (a.kill_after_s or os.path.getsize(logs[k]) > sizes[k]):
                        procs[k].send_signal(sig)
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

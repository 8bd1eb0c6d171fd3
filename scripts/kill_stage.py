#!/usr/bin/env python
"""Fault injector: terminate the server process of stage N (reference: scripts/kill_stage.py:16-80).

Finds processes whose command line is ``... -m src.main ... --stage N`` by reading /proc (no
pattern-based ``pkill``), shows them, asks for confirmation unless ``--yes``, and sends SIGTERM
(or SIGKILL with ``--hard``) to exactly those PIDs."""
import argparse
import os
import signal
import sys


def find(stage: int):
    out = []
    for pid in filter(str.isdigit, os.listdir("/proc")):
        try:
            argv = open(f"/proc/{pid}/cmdline", "rb").read().split(b"\0")
        except OSError:
            continue
        argv = [a.decode(errors="replace") for a in argv if a]
        if "src.main" in argv and "--stage" in argv and int(pid) != os.getpid():
            try:
                if int(argv[argv.index("--stage") + 1]) == stage:
                    out.append((int(pid), " ".join(argv)))
            except (ValueError, IndexError):
                pass
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("stage", type=int)
    ap.add_argument("--yes", "-y", action="store_true")
    ap.add_argument("--hard", action="store_true", help="SIGKILL instead of SIGTERM")
    a = ap.parse_args()
    procs = find(a.stage)
    if not procs:
        print(f"no process serving --stage {a.stage}")
        return 1
    for pid, cmd in procs:
        print(f"  pid {pid}: {cmd[:160]}")
    if not a.yes and input(f"kill {len(procs)} process(es)? [y/N] ").strip().lower() != "y":
        return 1
    for pid, _ in procs:
        os.kill(pid, signal.SIGKILL if a.hard else signal.SIGTERM)
        print(f"sent {'SIGKILL' if a.hard else 'SIGTERM'} to {pid}")
    return 0


if __name__ == "__main__":
    sys.exit(main())

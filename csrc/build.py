"""Builds csrc/*.cu into the in-tree shared library loaded by petals_b200.ops.native.

    python csrc/build.py [--force] [--verbose]

Every translation unit is compiled for sm_100a only (``-gencode
arch=compute_100a,code=sm_100a``) with ``-lineinfo`` so ncu's source page maps
back to these files. Objects are cached under build/ and rebuilt when a source
or header is newer. nvcc cross-compiles without a GPU, so this also runs on the
CPU-only development box.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG_DIR = os.path.join(ROOT, "global_capstone_design_distributed-inference-of-llms-over-the-internet_b200")
OUT_LIB = os.path.join(PKG_DIR, "ops", "libb200petals.so")
OBJ_DIR = os.path.join(ROOT, "build", "obj")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _sources() -> list[str]:
    return sorted(os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".cu"))


def _headers() -> list[str]:
    hs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".cuh", ".h"))]
    inc = os.path.join(ROOT, "include")
    hs += [os.path.join(inc, f) for f in os.listdir(inc) if f.endswith(".h")]
    return hs


def _newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, trace: bool = False) -> str:
    """trace=True builds the variant with the device-side event trace compiled in (csrc/trace.cuh) as
    ops/libb200petals_trace.so; select it with B200_LIB (scripts/trace_decode.py does)."""
    obj_dir = OBJ_DIR + ("_trace" if trace else "")
    out_lib = OUT_LIB.replace(".so", "_trace.so") if trace else OUT_LIB
    extra = ["-DB200_TRACE"] if trace else []
    os.makedirs(obj_dir, exist_ok=True)
    nvcc = _nvcc()
    headers = _headers()
    jobs = []
    objs = []
    for src in _sources():
        obj = os.path.join(obj_dir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if force or _newer(obj, [src] + headers):
            cmd = [nvcc, *NVCC_FLAGS, *extra, "-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        return cmd, r

    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for cmd, r in ex.map(run, jobs):
                if verbose or r.returncode != 0:
                    sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
                if r.returncode != 0:
                    raise RuntimeError(f"nvcc failed for {cmd[-3]}")
    if jobs or force or _newer(out_lib, objs):
        cmd = [nvcc, "-shared", "-o", out_lib, *objs]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return out_lib


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--trace", action="store_true", help="build the event-trace variant (libb200petals_trace.so)")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose, trace=a.trace))

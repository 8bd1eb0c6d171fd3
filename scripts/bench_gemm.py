"""Micro-benchmark of the tcgen05 stream-K GEMM on decode / prefill shapes.

Rotates through enough distinct weight copies (> L2) that every launch streams
its weight from HBM; times with CUDA events around a batch of launches.
Prints achieved GB/s (weights bytes / time) and TFLOP/s against
MEASURED_PEAKS.json.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import gemm, packing  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, nargs="+", default=[1, 16])
    ap.add_argument("--dtype", default="fp16")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--pdl", type=int, default=0)
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    dt = torch.float16 if args.dtype == "fp16" else torch.bfloat16
    peaks = {"hbm_gbs": 6551.0, "bf16_tflops": 1662.9}
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peaks.update(json.load(open(p)))
    shapes = [("qkv", 12288, 4096, False), ("o_proj", 4096, 4096, False), ("gate_up", 11008, 4096, True),
              ("down", 4096, 11008, False), ("lm_head", 32000, 4096, False)]
    rows = []
    for T in args.tokens:
        for name, N, K, dual in shapes:
            wbytes = N * K * 2 * (2 if dual else 1)
            copies = max(2, int(600e6 // wbytes) + 1)
            ws = []
            for _ in range(copies):
                if dual:
                    ws.append(packing.pack_dual((torch.randn(N, K, device="cuda") * 0.02).to(dt),
                                                (torch.randn(N, K, device="cuda") * 0.02).to(dt)))
                else:
                    ws.append(packing.pack_weight((torch.randn(N, K, device="cuda") * 0.02).to(dt)))
            x = torch.randn(T, K, device="cuda").to(dt)
            out = torch.empty(T, N, device="cuda", dtype=dt)
            wsp = gemm.GemmWorkspace.create("cuda", max(T, 16))
            for w in ws[:3]:
                gemm.linear(x, w, out=out, workspace=wsp, pdl=bool(args.pdl))
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            # one CUDA graph = one pass over all weight copies (no host launch cost in the timing)
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side):
                    for w in ws:
                        gemm.linear(x, w, out=out, workspace=wsp, pdl=bool(args.pdl))
            torch.cuda.current_stream().wait_stream(side)
            g.replay()
            torch.cuda.synchronize()
            n = 0
            e0.record()
            for _ in range(args.iters):
                g.replay()
                n += len(ws)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / n
            gbs = wbytes / us / 1e3
            tf = 2.0 * T * N * K * (2 if dual else 1) / us / 1e6
            rows.append(dict(shape=name, T=T, N=N, K=K, dual=dual, us=round(us, 2), gbs=round(gbs, 1),
                             frac_hbm=round(gbs / peaks["hbm_gbs"], 3), tflops=round(tf, 1),
                             frac_tc=round(tf / peaks["bf16_tflops"], 3)))
            print(rows[-1], flush=True)
            del ws
            torch.cuda.empty_cache()
    if args.json:
        json.dump(rows, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()

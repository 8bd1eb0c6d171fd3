"""Per-kernel timeline of one decode step from the device-side event trace (csrc/trace.cuh).

    python scripts/trace_decode.py [--layers 8] [--model llama-2-7b] [--out gpurun_out/trace_decode.json]

For every GEMM / attention launch of one CUDA-graph replay: when its first and last CTA entered, when the
dependency wait ended, when the first accumulator was ready and when the CTA left - all in microseconds from
the first event, plus the gaps between consecutive kernels.
"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
# the stamps are compiled in only in the trace variant of the library (python csrc/build.py --trace)
os.environ.setdefault("B200_LIB", os.path.join(
    ROOT, "global_capstone_design_distributed-inference-of-llms-over-the-internet_b200", "ops", "libb200petals_trace.so"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-2-7b")
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--prompt", type=int, default=128)
    ap.add_argument("--out", default="gpurun_out/trace_decode.json")
    a = ap.parse_args()
    from petals_b200.models.config import resolve_config
    from petals_b200.ops import native
    from petals_b200.runtime.kv_cache import KVCacheManager
    from petals_b200.runtime.native_span import NativeSpan
    from petals_b200.weights.loader import load_span

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cfg = resolve_config(a.model)
    dt = torch.float16
    w = load_span(a.model, cfg, 0, a.layers, need_embed=True, need_head=True, dtype=dt, device=dev, seed=0)
    span = NativeSpan(cfg, w, dev, dt, kv_capacity=512, max_batch=1, max_chunk=256)
    kv = KVCacheManager(a.layers, cfg.n_kv_heads, cfg.head_dim, dt, dev).allocate("s", 512)
    ids = torch.randint(0, cfg.vocab, (a.prompt,), device=dev)
    _, tok = span.prefill(kv, ids=ids)
    for _ in range(8):
        _, tok = span.decode([kv], ids=tok.clone())
    torch.cuda.synchronize()
    cap = 1 << 17
    buf = torch.zeros(1 + 2 * cap, dtype=torch.int64, device=dev)
    L = native.lib()
    L.b200_trace_set.argtypes = [ctypes.c_void_p, ctypes.c_int32]
    native.check(L.b200_trace_set(buf.data_ptr(), cap), "trace on")
    _, tok = span.decode([kv], ids=tok.clone())
    torch.cuda.synchronize()
    native.check(L.b200_trace_set(None, 0), "trace off")
    h = buf.cpu().tolist()
    n = min(h[0], cap)
    ev = sorted((h[1 + 2 * i], h[2 + 2 * i]) for i in range(n))
    t0 = ev[0][0]
    chain_ev = [(t, tag) for t, tag in ev if (tag >> 56) == 3]
    if chain_ev:
        # persistent chain kernel: per op and CTA (first / last): begin, staged (inputs consumed), accumulator ready, end
        rows = {}
        for t, tag in chain_ev:
            oi, ph, which = (tag >> 32) & 0xffffff, (tag >> 16) & 0xff, tag & 3
            rows.setdefault((oi, which), {})[ph] = (t - t0) / 1e3
        print(f"persistent chain kernel: {len(chain_ev)} events; step = {(ev[-1][0] - t0) / 1e3:.1f} us")
        print(f"{'op':>4} | CTA 0 workers: {'begin':>8} {'staged':>8} {'acc':>8} {'end':>8} | loader {'begin':>8} {'issued':>8} | mma {'begin':>8} {'commit':>8}")
        nops = max(k[0] for k in rows) + 1
        for oi in list(range(min(nops, 14))) + list(range(max(14, nops - 3), nops)):
            a, b = rows.get((oi, 0), {}), rows.get((oi, 1), {})
            f = lambda d, k: f"{d[k]:8.1f}" if k in d else "       -"   # noqa: E731
            print(f"{oi:>4} |                " + " ".join(f(a, k) for k in range(4)) + " |        " + " ".join(f(a, k) for k in (4, 5))
                  + " |     " + " ".join(f(a, k) for k in (6, 7)))
        ev = [(t, tag) for t, tag in ev if (tag >> 56) != 3]
    names = {}

    def kname(tag):
        kid, aux = tag >> 56, (tag >> 32) & 0xffffff
        if kid == 2:
            return "attn"
        nt, kb = aux & 0xfff, aux >> 12
        key = (nt, kb)
        if key not in names:
            names[key] = f"gemm[{nt}x{kb}]"
        return names[key]

    # the i-th event of a (kernel type, phase, cta) belongs to the i-th launch of that type (launches of one type
    # never overlap each other); launches are then ordered by the time their first CTA entered
    per = {}
    for t, tag in ev:
        which = tag & 3
        ph = ["enter", "dep", "acc", "exit"][(tag >> 16) & 0xff]
        if ph == "exit":                      # every CTA stamps its exit: keep them all for the true end of the launch
            per.setdefault((kname(tag), "exit_all", "*"), []).append((t - t0) / 1e3)
        if which == 2:
            continue
        k = (kname(tag), ph, "L" if which == 1 else "0")
        per.setdefault(k, []).append((t - t0) / 1e3)
    launches = []
    for kn in sorted({k[0] for k in per}):
        cnt = len(per.get((kn, "enter", "0"), []))
        for i in range(cnt):
            d = {"kernel": kn}
            for ph in ("enter", "dep", "acc", "exit"):
                for c in ("0", "L"):
                    lst = per.get((kn, ph, c), [])
                    # "acc" is stamped once per part; keep the first one after this launch's enter
                    if ph == "acc":
                        ent = per[(kn, "enter", c)][i] if i < len(per.get((kn, "enter", c), [])) else None
                        nxt = per[(kn, "enter", c)][i + 1] if ent is not None and i + 1 < len(per[(kn, "enter", c)]) else 1e18
                        cand = [x for x in lst if ent is not None and ent <= x < nxt]
                        d[ph + c] = cand[0] if cand else float("nan")
                    else:
                        d[ph + c] = lst[i] if i < len(lst) else float("nan")
            # true end: the last exit of any CTA between this launch's entry and the next launch of the same type
            ent = per[(kn, "enter", "0")][i]
            nxt = per[(kn, "enter", "0")][i + 1] if i + 1 < cnt else 1e18
            ex = [x for x in per.get((kn, "exit_all", "*"), []) if ent <= x < nxt]
            d["exit_max"] = max(ex) if ex else float("nan")
            d["exit_med"] = sorted(ex)[len(ex) // 2] if ex else float("nan")
            d["n_exit"] = len(ex)
            launches.append(d)
    launches.sort(key=lambda d: d["enter0"])
    print(f"{n} events, {len(launches)} launches; step = {(ev[-1][0] - t0) / 1e3:.1f} us for {a.layers} layers + head")
    print(f"{'#':>3} {'kernel':14} {'enter0':>8} {'dep0':>8} {'acc0':>8} {'exit0':>8} | {'enterL':>8} {'depL':>8} {'accL':>8} {'exitL':>8} | prev_end->dep  dep->end")
    summary = []
    prev_end = 0.0
    for i, d in enumerate(launches):
        end_ = d["exit_max"]
        dep = max(x for x in (d["dep0"], d["depL"]) if x == x) if (d["dep0"] == d["dep0"] or d["depL"] == d["depL"]) else float("nan")
        d["end"], d["dep"] = end_, dep
        summary.append(d)
        if 5 <= i < 16 or i >= len(launches) - 3:
            print(f"{i:>3} {d['kernel']:14} " + " ".join(f"{d[k]:8.1f}" for k in ("enter0", "dep0", "acc0", "exit0")) + " | "
                  + " ".join(f"{d[k]:8.1f}" for k in ("enterL", "depL", "accL", "exitL"))
                  + f" | {dep - prev_end:8.1f} {end_ - dep:8.1f}  end(all {d['n_exit']} CTAs) {end_:8.1f} med {d['exit_med']:8.1f}")
        prev_end = end_
    # mean over layers per kernel type
    print("mean per kernel type (us): wait after the previous kernel's end until the dependency is visible, then dep -> end")
    for kn in sorted({d["kernel"] for d in summary}):
        sel = [(j, d) for j, d in enumerate(summary) if d["kernel"] == kn and j > 0]
        if not sel:
            continue
        gaps = [d["dep"] - summary[j - 1]["end"] for j, d in sel]
        runs = [d["end"] - d["dep"] for j, d in sel]
        pre = [d["dep"] - d["enter0"] for j, d in sel]
        print(f"  {kn:14} n={len(sel):3d}  prev_end->dep {sum(gaps) / len(gaps):6.2f}   dep->end {sum(runs) / len(runs):6.2f}   resident before dep {sum(pre) / len(pre):6.2f}")
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    with open(a.out, "w") as f:
        json.dump({"model": a.model, "layers": a.layers, "launches": summary}, f)


if __name__ == "__main__":
    main()

"""Are two CTAs of the decode GEMM ever resident on one SM at the same time?  (trace build: every CTA stamps its SM id
at entry and exit)"""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("B200_LIB", os.path.join(ROOT, "global_capstone_design_distributed-inference-of-llms-over-the-internet_b200", "ops", "libb200petals_trace.so"))
from petals_b200.ops import gemm as G, native, packing

dev = torch.device("cuda")
w = torch.randn(12288, 4096, device=dev, dtype=torch.float16) * 0.02
pw = packing.pack_weight(w)
x = torch.randn(1, 4096, device=dev, dtype=torch.float16)
ws = G.GemmWorkspace.create(dev, 16)
for _ in range(3):
    G.linear(x, pw, workspace=ws)
torch.cuda.synchronize()
cap = 4096
buf = torch.zeros(1 + 2 * cap, dtype=torch.int64, device=dev)
L = native.lib()
L.b200_trace_set(ctypes.c_void_p(buf.data_ptr()), cap)
G.linear(x, pw, workspace=ws)
torch.cuda.synchronize()
L.b200_trace_set(None, 0)
h = buf.cpu().tolist()
n = min(h[0], cap)
ev = [(h[1 + 2 * i], h[2 + 2 * i]) for i in range(n)]
t0 = min(t for t, _ in ev)
per_sm = {}
for t, tag in ev:
    ph = (tag >> 16) & 0xff
    sm = (tag >> 2) & 0xfff
    if ph in (0, 3):
        per_sm.setdefault(sm, []).append(((t - t0) / 1e3, ph))
overlap = 0
for sm, lst in sorted(per_sm.items()):
    lst.sort()
    live = 0
    mx = 0
    for _, ph in lst:
        live += 1 if ph == 0 else -1
        mx = max(mx, live)
    overlap += mx >= 2
print(f"{n} events on {len(per_sm)} SMs; SMs that ever held 2 CTAs at once: {overlap}")
for sm in sorted(per_sm)[:4]:
    print(sm, [(round(t, 1), "in" if ph == 0 else "out") for t, ph in sorted(per_sm[sm])])

"""GEMM-only decode chain: qkv -> o_proj(+res) -> gate/up(SwiGLU) -> down(+res), L layers
with distinct weights captured in one CUDA graph.  Shows the per-layer cost of the four
weight-streaming kernels back to back (launch gaps included) against the HBM roofline."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import gemm, packing

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=16)
ap.add_argument("--tokens", type=int, default=1)
ap.add_argument("--pdl", type=int, default=1)
ap.add_argument("--norm", type=int, default=1)
ap.add_argument("--hidden", type=int, default=4096)
ap.add_argument("--inter", type=int, default=11008)
ap.add_argument("--kv", type=int, default=4096)
a = ap.parse_args()
dt = torch.float16
H, I, T = a.hidden, a.inter, a.tokens
mk = lambda n, k: (torch.randn(n, k, device="cuda") * 0.02).to(dt)
layers = []
for _ in range(a.layers):
    layers.append(dict(qkv=packing.pack_weight(mk(H + 2 * a.kv, H)), o=packing.pack_weight(mk(H, H)),
                       gu=packing.pack_dual(mk(I, H), mk(I, H)), down=packing.pack_weight(mk(H, I)),
                       ln1=torch.ones(H, device="cuda", dtype=dt), ln2=torch.ones(H, device="cuda", dtype=dt)))
wbytes = sum(l["qkv"].nbytes + l["o"].nbytes + l["gu"].nbytes + l["down"].nbytes for l in layers)
h = torch.randn(T, H, device="cuda").to(dt) * 0.1
h2 = torch.empty_like(h)
qkv = torch.empty(T, H + 2 * a.kv, device="cuda", dtype=dt)
act = torch.empty(T, I, device="cuda", dtype=dt)
ws = gemm.GemmWorkspace.create("cuda", max(T, 16))
pdl = bool(a.pdl)
nm = gemm.NORM_RMS if (a.norm and T <= 32) else gemm.NORM_NONE

def step():
    for l in layers:
        gemm.linear(h, l["qkv"], out=qkv, norm_mode=nm, norm_w=l["ln1"], workspace=ws, pdl=pdl)
        gemm.linear(qkv[:, :H], l["o"], out=h2, residual=h, workspace=ws, pdl=pdl)
        gemm.linear(h2, l["gu"], out=act, norm_mode=nm, norm_w=l["ln2"], workspace=ws, pdl=pdl)
        gemm.linear(act, l["down"], out=h, residual=h2, workspace=ws, pdl=pdl)

step(); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph(); side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    with torch.cuda.graph(g, stream=side):
        step()
torch.cuda.current_stream().wait_stream(side)
for _ in range(3): g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(20): g.replay()
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / 20
print(json.dumps(dict(layers=a.layers, tokens=T, pdl=a.pdl, norm=a.norm, us_per_layer=round(us / a.layers, 2),
                      gbs=round(wbytes / us / 1e3, 1), frac_hbm_measured=round(wbytes / us / 1e3 / 6551.0, 3))))

"""What does the stream-K fix-up cost?  Equal weight bytes (~77 MB), T=1, 16 k-blocks per CTA on 296 CTAs:
A: 296 tiles x K=1024 (every CTA owns a whole tile: no partials), B: 37 tiles x K=8192 (8 partials per tile),
C: 96 tiles x K=3072 (3 partials per tile, grid 288).  CUDA-graphed chains over > L2 distinct weights."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import gemm, packing
dt = torch.float16
for pdl in (0, 1):
    for name, N, K in (("A no-split", 37888, 1024), ("B 8-way", 4736, 8192), ("C 3-way", 12288, 3072)):
        ws = [packing.pack_weight((torch.randn(N, K, device="cuda") * 0.02).to(dt)) for _ in range(10)]
        x = torch.randn(1, K, device="cuda").to(dt); out = torch.empty(1, N, device="cuda", dtype=dt)
        wsp = gemm.GemmWorkspace.create("cuda", 16)
        for w in ws[:2]: gemm.linear(x, w, out=out, workspace=wsp, pdl=bool(pdl))
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph(); side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for _ in range(3):
                    for w in ws: gemm.linear(x, w, out=out, workspace=wsp, pdl=bool(pdl))
        torch.cuda.current_stream().wait_stream(side)
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(5): g.replay()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (5 * 30)
        print(f"pdl={pdl} {name:12s} N={N} K={K}: {us:6.2f} us/launch, {N*K*2/us/1e3:7.1f} GB/s", flush=True)

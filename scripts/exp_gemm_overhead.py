import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import gemm, packing
dt = torch.float16
def run(name, N, K, grid, T=1, dual=False, pdl=False, debug=0):
    wbytes = N*K*2*(2 if dual else 1)
    copies = max(2, int(600e6//wbytes)+1)
    mk = lambda: (torch.randn(N, K, device="cuda")*0.02).to(dt)
    ws = [packing.pack_dual(mk(), mk()) if dual else packing.pack_weight(mk()) for _ in range(copies)]
    x = torch.randn(T, K, device="cuda").to(dt); out = torch.empty(T, N, device="cuda", dtype=dt)
    wsp = gemm.GemmWorkspace.create("cuda", 16)
    gemm.linear(x, ws[0], out=out, workspace=wsp, grid=grid, debug=debug); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph(); side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for w in ws: gemm.linear(x, w, out=out, workspace=wsp, grid=grid, pdl=pdl, debug=debug)
    torch.cuda.current_stream().wait_stream(side)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(20): g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1)*1e3/(20*len(ws))
    print(f"{name:8s} N={N} K={K} grid={grid} T={T} pdl={int(pdl)} dbg={debug}: {us:.2f} us  {wbytes/us/1e3:.0f} GB/s", flush=True)
for dbg in (0, 3):
    run("o_proj", 4096, 4096, 32, debug=dbg)
run("o_proj", 4096, 4096, 128)
run("o_proj", 4096, 4096, 148)
run("qkv", 12288, 4096, 148)
run("gate_up", 11008, 4096, 148, dual=True)
run("down", 4096, 11008, 148)
run("lm_head", 32000, 4096, 148)
run("lm_head", 32000, 4096, 148, debug=3)

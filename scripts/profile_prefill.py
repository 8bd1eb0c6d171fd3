"""Prefill workload for ncu: the four Llama-2-7B GEMMs and the attention kernel at T=2048."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import attention as A, gemm, packing
dt = torch.float16; dev = "cuda"; T = 2048; H = 4096; I = 11008
mk = lambda n, k: (torch.randn(n, k, device=dev) * 0.02).to(dt)
x = torch.randn(T, H, device=dev).to(dt); xi = torch.randn(T, I, device=dev).to(dt)
ws = {"qkv": packing.pack_weight(mk(3 * H, H)), "o": packing.pack_weight(mk(H, H)),
      "gu": packing.pack_dual(mk(I, H), mk(I, H)), "down": packing.pack_weight(mk(H, I))}
res = torch.randn(T, H, device=dev).to(dt)
for _ in range(2):
    q = gemm.linear(x, ws["qkv"]); o = gemm.linear(x, ws["o"], residual=res)
    g = gemm.linear(x, ws["gu"]); d = gemm.linear(xi, ws["down"], residual=res)
    kc = torch.randn(32, T, 128, device=dev).to(dt); vc = torch.randn(32, T, 128, device=dev).to(dt)
    out = torch.empty(T, H, device=dev, dtype=dt)
    A.attn_prefill(q, out, kc, vc, 32, 32, 128, T, 0)
torch.cuda.synchronize(); print("ok")
if os.environ.get("TIMING"):
    def tm(f, n=20):
        for _ in range(3): f()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(n): f()
        e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n * 1e3
    print({"qkv_us": tm(lambda: gemm.linear(x, ws["qkv"])), "o_res_us": tm(lambda: gemm.linear(x, ws["o"], residual=res)),
           "o_nores_us": tm(lambda: gemm.linear(x, ws["o"])), "gate_up_us": tm(lambda: gemm.linear(x, ws["gu"])),
           "down_res_us": tm(lambda: gemm.linear(xi, ws["down"], residual=res)),
           "down_nores_us": tm(lambda: gemm.linear(xi, ws["down"])),
           "attn_us": tm(lambda: A.attn_prefill(q, out, kc, vc, 32, 32, 128, T, 0))})

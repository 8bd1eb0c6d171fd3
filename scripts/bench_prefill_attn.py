"""Prefill attention: tcgen05 kernel vs the CUDA-core kernel (Llama-2-7B shape: 32 heads, d=128)."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import attention as A
dev = "cuda"; dt = torch.float16
nh, nkv, D = 32, 32, 128
rows = []
for T in (256, 1024, 2048, 4096):
    s_max = T
    qkv = torch.randn(T, (nh + 2 * nkv) * D, device=dev).to(dt)
    kc = torch.randn(nkv, s_max, D, device=dev).to(dt); vc = torch.randn(nkv, s_max, D, device=dev).to(dt)
    out = torch.empty(T, nh * D, device=dev, dtype=dt)
    for tc in (True, False):
        for _ in range(3): A.attn_prefill(qkv, out, kc, vc, nh, nkv, D, s_max, 0, tensor_cores=tc)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        n = 10
        e0.record()
        for _ in range(n): A.attn_prefill(qkv, out, kc, vc, nh, nkv, D, s_max, 0, tensor_cores=tc)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        flops = 4.0 * nh * D * T * T / 2
        rows.append(dict(T=T, kernel="tcgen05" if tc else "cuda-core", ms=round(ms, 3), tflops=round(flops / ms / 1e9, 1)))
        print(rows[-1], flush=True)
if len(sys.argv) > 1: json.dump(rows, open(sys.argv[1], "w"), indent=1)

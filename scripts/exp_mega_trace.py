import os, sys, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.models.config import ModelConfig
from petals_b200.runtime.kv_cache import KVCacheManager
from petals_b200.runtime.native_span import NativeSpan
from petals_b200.weights.loader import load_span
from petals_b200.ops import native
layers = int(os.environ.get("LAYERS", "8"))
cfg = ModelConfig("llama", layers, 4096, 32, 32, 128, 11008, 32000, name="slice")
dev = torch.device("cuda", 0)
w = load_span("slice", cfg, 0, layers, need_embed=True, need_head=True, dtype=torch.float16, device=dev)
span = NativeSpan(cfg, w, dev, torch.float16, kv_capacity=2048, max_batch=1, max_chunk=256, use_graphs=False)
kv = KVCacheManager(layers, 32, 128, torch.float16, dev).allocate("p", 2048)
ids = torch.randint(0, 32000, (int(os.environ.get("PROMPT", "128")),), device=dev)
_, tok = span.prefill(kv, ids=ids)
for _ in range(3): _, tok = span.decode([kv], ids=tok.clone())
n_ops = layers * 5 + 1
trace = torch.zeros(2 * n_ops * 2, dtype=torch.int64, device=dev)
L = native.lib(); L.b200_decode_mega_set_trace.argtypes = [ctypes.c_void_p]; L.b200_decode_mega_set_trace(trace.data_ptr())
_, tok = span.decode([kv], ids=tok.clone()); torch.cuda.synchronize()
L.b200_decode_mega_set_trace(None)
t = trace.cpu().view(2, n_ops, 2).double() / 1e3
names = ["qkv", "attn", "o", "gate_up", "down"]
t0 = t[0, 0, 0]
print("op        cta0: start   work  gap_to_next | ctaL: work")
for i in range(min(n_ops, 16)):
    nm = names[i % 5] if i < n_ops - 1 else "lm_head"
    gap = (t[0, i + 1, 0] - t[0, i, 1]) if i + 1 < n_ops else 0
    print(f"{i:3d} {nm:8s} {t[0,i,0]-t0:8.1f} {t[0,i,1]-t[0,i,0]:6.1f} {gap:6.1f}       | {t[1,i,1]-t[1,i,0]:6.1f}")
tot = t[0, n_ops - 1, 1] - t0
work = (t[0, :, 1] - t[0, :, 0])
gaps = (t[0, 1:, 0] - t[0, :-1, 1])
print(f"total {tot:.1f} us; sum work {work.sum():.1f}; sum barrier gaps {gaps.sum():.1f}; per-type work:")
for k, nm in enumerate(names):
    sel = work[k:n_ops - 1:5]
    print(f"  {nm}: mean {sel.mean():.1f} us")
print(f"  lm_head: {work[-1]:.1f} us")

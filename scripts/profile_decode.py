"""Small decode workload for ncu: a 4-block Llama-2-7B-shaped span, no CUDA graph, a few steps.
(Per-kernel behaviour equals the 32-block model's; the short run keeps ncu's replays cheap.)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.models.config import ModelConfig  # noqa: E402
from petals_b200.runtime.kv_cache import KVCacheManager  # noqa: E402
from petals_b200.runtime.native_span import NativeSpan  # noqa: E402
from petals_b200.weights.loader import load_span  # noqa: E402

layers = int(os.environ.get("LAYERS", "4"))
steps = int(os.environ.get("STEPS", "3"))
batch = int(os.environ.get("BATCH", "1"))          # sessions per decode step
weights = os.environ.get("WEIGHTS", "native")      # native | fp8 | mxfp8
cfg = ModelConfig("llama", layers, 4096, 32, 32, 128, 11008, 32000, name="llama-2-7b-slice")
dev = torch.device("cuda", 0)
w = load_span("llama-2-7b-slice", cfg, 0, layers, need_embed=True, need_head=True, dtype=torch.float16, device=dev)
span = NativeSpan(cfg, w, dev, torch.float16, kv_capacity=1024, max_batch=batch, max_chunk=256, use_graphs=False,
                  weight_format=weights)
mgr = KVCacheManager(layers, 32, 128, torch.float16, dev)
kvs = [mgr.allocate(f"p{i}", 1024) for i in range(batch)]
toks = []
for kv in kvs:
    ids = torch.randint(0, 32000, (int(os.environ.get("PROMPT", "128" if batch > 1 else "512")),), device=dev)
    _, tok = span.prefill(kv, ids=ids)
    toks.append(tok.clone())
tok = torch.cat(toks)
torch.cuda.synchronize()
for _ in range(steps):
    _, tok = span.decode(kvs, ids=tok.clone())
torch.cuda.synchronize()
print("launches", span.kernel_launches, "last token", int(tok[0]))

"""TTFT: full-model prefill time on one GPU for several prompt lengths (CUDA events)."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.parallel.pipeline import PipelineEngine
model = sys.argv[1] if len(sys.argv) > 1 else "llama-2-7b"
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
eng = PipelineEngine(model, [], torch.float16, dev, kv_capacity=4352, max_sessions=1, max_chunk=int(os.environ.get("CHUNK", "2048")))
rows = []
for L in (128, 512, 2048, 4096):
    ids = torch.randint(0, eng.cfg.vocab, (L,))
    for it in range(3):
        eng.close_sessions(); s = eng.open_sessions(1)[0]
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); eng.prefill(s, ids); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    flops = 2.0 * L * (eng.cfg.layer_param_count() * eng.cfg.n_layers) + 4.0 * eng.cfg.n_heads * eng.cfg.head_dim * L * L / 2 * eng.cfg.n_layers
    rows.append(dict(model=model, prompt=L, ttft_ms=round(ms, 2), tflops=round(flops / ms / 1e9, 1)))
    print(rows[-1], flush=True)
if len(sys.argv) > 2: json.dump(rows, open(sys.argv[2], "w"), indent=1)

import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.models.config import PRESETS
from petals_b200.parallel.tensor_parallel import make_tp_span
from petals_b200.runtime.kv_cache import KVCacheManager
from petals_b200.runtime.native_span import NativeSpan
from petals_b200.weights.loader import load_span
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
cfg = PRESETS["tiny-llama-gqa"]; dtype = torch.bfloat16
ids = torch.randint(0, cfg.vocab, (40,), generator=torch.Generator().manual_seed(0)).to(dev)
res = {}
for mode, graphs, pdl in [("single", False, False), ("fused", False, False), ("fused", False, True), ("fused", True, False), ("nccl", False, False)]:
    w = load_span("tiny-llama-gqa", cfg, 0, cfg.n_layers, need_embed=True, need_head=True, dtype=dtype, device=dev, seed=9)
    if mode == "single":
        span = NativeSpan(cfg, w, dev, dtype, kv_capacity=256, max_batch=2, max_chunk=64, use_graphs=False, pdl=False); lc = cfg
    else:
        span = make_tp_span(cfg, w, dev, dtype, dist.group.WORLD, mode=mode, kv_capacity=256, max_batch=2, max_chunk=64, use_graphs=graphs, pdl=pdl); lc = span.cfg
    kv = KVCacheManager(cfg.n_layers, lc.n_kv_heads, lc.head_dim, dtype, dev).allocate("s", 256)
    logits, tok = span.prefill(kv, ids=ids); outs = [logits.float().clone()]
    toks = [int(tok[0])]
    for _ in range(4):
        logits, tok = span.decode([kv], ids=tok.clone()); outs.append(logits.float().clone()); toks.append(int(tok[0]))
    torch.cuda.synchronize(); dist.barrier()
    key = f"{mode}/g{int(graphs)}/p{int(pdl)}"
    res[key] = torch.cat(outs)
    if rank == 0 and mode != "single":
        d = (res[key] - res["single/g0/p0"]).abs().amax(-1)
        print(key, "per-step max err:", [round(x, 4) for x in d.tolist()], "toks", toks, flush=True)
    elif rank == 0:
        print(key, "toks", toks, flush=True)
dist.destroy_process_group()

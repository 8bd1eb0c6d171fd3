"""2/4/8-way tensor-parallel span: fused GEMM+all-reduce (peer memory) vs the NCCL baseline vs one GPU."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(rank, world, port, q, preset="tiny-llama-gqa"):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from petals_b200.models.config import PRESETS
    from petals_b200.parallel.tensor_parallel import make_tp_span
    from petals_b200.runtime.kv_cache import KVCacheManager
    from petals_b200.runtime.native_span import NativeSpan
    from petals_b200.weights.loader import load_span

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    cfg = PRESETS[preset]
    dtype = torch.bfloat16
    res = {}
    ids = torch.randint(0, cfg.vocab, (40,), generator=torch.Generator().manual_seed(0)).to(dev)
    for mode in ("single", "fused", "fused-nodefer", "nccl"):
        w = load_span(preset, cfg, 0, cfg.n_layers, need_embed=True, need_head=True, dtype=dtype, device=dev, seed=9)
        if mode == "single":
            span = NativeSpan(cfg, w, dev, dtype, kv_capacity=256, max_batch=2, max_chunk=64, use_graphs=False)
            lc = cfg
        else:
            span = make_tp_span(cfg, w, dev, dtype, dist.group.WORLD, mode=mode.split("-")[0], kv_capacity=256,
                                max_batch=2, max_chunk=64, use_graphs=(mode == "fused"))
            span.tp_defer = mode == "fused"
            lc = span.cfg
        kv = KVCacheManager(cfg.n_layers, lc.n_kv_heads, lc.head_dim, dtype, dev).allocate("s", 256)
        logits, tok = span.prefill(kv, ids=ids)
        outs = [logits.float().clone()]
        forced = torch.randint(0, cfg.vocab, (5,), generator=torch.Generator().manual_seed(1)).to(dev)
        for i in range(5):   # teacher forcing: a near-tie may flip the greedy token between modes
            logits, tok = span.decode([kv], ids=forced[i:i + 1])
            outs.append(logits.float().clone())
        torch.cuda.synchronize()
        res[mode] = torch.cat(outs).cpu()
        if mode == "fused":
            res["nvls_used"] = torch.tensor(float(getattr(span, "nvls", None) is not None))
        dist.barrier()
    # every rank must hold identical logits in the fused mode
    mine = res["fused"].to(dev)
    other = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(other, mine)
    same = all(torch.equal(o, mine) for o in other)
    if rank == 0:
        q.put((res, same))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,preset", [(2, "tiny-llama-gqa"), (4, "tiny-llama-tp8"), (8, "tiny-llama-tp8")])
def test_tp_fused_allreduce_matches_single_gpu_and_nccl(world, preset):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_run, args=(r, world, port, q, preset)) for r in range(world)]
    for p in procs:
        p.start()
    res, same = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert same, "fused all-reduce must leave bit-identical results on every rank"
    assert res.pop("nvls_used").item() == 1.0, "the 40-token prefill chunk should have been reduced through the switch (NVLS)"
    scale = res["single"].abs().max().item()
    assert (res["fused"] - res["single"]).abs().max().item() < 0.05 * scale
    assert (res["nccl"] - res["single"]).abs().max().item() < 0.05 * scale
    assert (res["fused"] - res["nccl"]).abs().max().item() < 0.02 * scale
    assert torch.equal(res["fused"], res["fused-nodefer"]), "deferred and in-producer reductions must agree bit for bit"

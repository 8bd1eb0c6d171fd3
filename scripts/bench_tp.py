"""TP-2/4/8 decode micro-benchmark: fused GEMM+all-reduce (peer memory) vs GEMM + NCCL all-reduce.
Llama-2-7B-shaped row-parallel linears (o_proj 4096x4096, down 4096x11008) at T=1, CUDA-graphed,
device-timed (max over ranks).  torchrun --nproc-per-node R scripts/bench_tp.py"""
import json, os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import gemm, packing
from petals_b200.parallel.tensor_parallel import TPWorkspace, linear_allreduce
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
dt = torch.float16
tp = TPWorkspace(dist.group.WORLD, dev, 4096)
ws = gemm.GemmWorkspace.create(dev, 16)
rows = []
for name, N, Kfull in (("o_proj", 4096, 4096), ("down", 4096, 11008)):
    K = Kfull // world
    K = K // 64 * 64
    copies = max(2, int(300e6 // (N * K * 2)) + 1)
    wts = [packing.pack_weight((torch.randn(N, K, device=dev) * 0.02).to(dt)) for _ in range(copies)]
    x = torch.randn(1, K, device=dev).to(dt); res = torch.randn(1, N, device=dev).to(dt)
    out = torch.empty(1, N, device=dev, dtype=dt)
    def fused():
        for w in wts: linear_allreduce(x, w, tp, out=out, residual=res, workspace=ws)
    gu = packing.pack_weight((torch.randn(256, N, device=dev) * 0.02).to(dt))   # a small consumer GEMM
    out2 = torch.empty(1, 256, device=dev, dtype=dt); hbuf = torch.empty(1, N, device=dev, dtype=dt)
    from petals_b200.parallel.tensor_parallel import linear_gather
    def deferred():
        for w in wts:
            linear_allreduce(x, w, tp, out=out, workspace=ws, defer=True)
            linear_gather(hbuf, gu, tp, residual=res, out=out2, workspace=ws)
    def fused_plus_consumer():
        for w in wts:
            linear_allreduce(x, w, tp, out=hbuf, residual=res, workspace=ws)
            gemm.linear(hbuf, gu, out=out2, workspace=ws)
    def nccl_plus_consumer():
        for w in wts:
            gemm.linear(x, w, out=hbuf, workspace=ws); dist.all_reduce(hbuf); hbuf.add_(res)
            gemm.linear(hbuf, gu, out=out2, workspace=ws)
    def nccl():
        for w in wts:
            gemm.linear(x, w, out=out, workspace=ws); dist.all_reduce(out); out.add_(res)
    for label, fn in (("fused", fused), ("nccl", nccl), ("push+gather-gemm", deferred),
                      ("fused+consumer-gemm", fused_plus_consumer), ("nccl+consumer-gemm", nccl_plus_consumer)):
        fn(); torch.cuda.synchronize(); dist.barrier()
        g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s): fn()
        torch.cuda.current_stream().wait_stream(s)
        g.replay(); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(20): g.replay()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        us = float(t) * 1e3 / (20 * copies)
        rows.append(dict(op=name, mode=label, tp=world, N=N, K_shard=K, us=round(us, 2),
                         weight_gbs_per_gpu=round(N * K * 2 / us / 1e3, 1)))
        if rank == 0: print(rows[-1], flush=True)
        dist.barrier()
if rank == 0 and len(sys.argv) > 1: json.dump(rows, open(sys.argv[1], "w"), indent=1)
torch.cuda.synchronize(); dist.barrier()
sys.stdout.flush()
os._exit(0)   # symmetric-memory teardown can block in destroy_process_group; nothing left to flush

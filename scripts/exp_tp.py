import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import gemm, packing
from petals_b200.parallel.tensor_parallel import TPWorkspace, linear_allreduce
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
dt = torch.bfloat16
tp = TPWorkspace(dist.group.WORLD, dev, 4096)
ws = gemm.GemmWorkspace.create(dev, 16)
for (T, N, K) in [(1, 512, 256), (1, 4096, 2048), (3, 512, 1024), (16, 4096, 5504)]:
    torch.manual_seed(5)
    xf = (torch.randn(T, K * world, device=dev) * 0.5).to(dt)
    wf = (torch.randn(N, K * world, device=dev) / (K * world) ** 0.5).to(dt)
    res = torch.randn(T, N, device=dev).to(dt)
    x = xf[:, rank * K:(rank + 1) * K].contiguous()
    w = packing.pack_weight(wf[:, rank * K:(rank + 1) * K].contiguous())
    ref = (xf.float() @ wf.float().t()).to(dt).float() + res.float()
    for it in range(4):
        out = torch.full((T, N), float("nan"), device=dev, dtype=dt)
        linear_allreduce(x, w, tp, out=out, residual=res, workspace=ws)
        torch.cuda.synchronize()
        err = (out.float() - ref).abs().max().item()
        print(f"rank{rank} T={T} N={N} K={K} it={it} err={err:.4f} seq={int(tp.local[0])} nan={int(out.float().isnan().sum())}", flush=True)
    dist.barrier()
dist.destroy_process_group()

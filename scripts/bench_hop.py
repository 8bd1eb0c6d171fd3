"""Latency of one pipeline hop (csrc/p2p.cu) between two GPUs: a ping-pong of hidden states captured in CUDA graphs.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/bench_hop.py [--hidden 4096]

Rank 0: send -> recv, rank 1: recv -> send, `--rounds` times per graph replay; one hop = half a round trip and contains
the send kernel (LL words stored into the peer's ring over NVLink), the flight, and the polling receive kernel that
unpacks them - everything a stage's graph adds per slot for its input.  For comparison the same ping-pong over NCCL
send/recv (host-issued, not capturable together with the compute in the same way).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, default=4096)
    ap.add_argument("--rounds", type=int, default=50)
    ap.add_argument("--tokens", type=int, default=1)
    ap.add_argument("--out", default="gpurun_out/bench_hop.json")
    a = ap.parse_args()
    rank = int(os.environ["RANK"])
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    from petals_b200.parallel.hop import HopLinks

    hl = HopLinks(dev, dist.group.WORLD, a.hidden, depth=8, t_max=max(a.tokens, 1))
    x = torch.randn(a.tokens, a.hidden, device=dev).to(torch.float16)
    y = torch.zeros_like(x)
    other = 1 - rank

    def pingpong():
        for _ in range(a.rounds):
            if rank == 0:
                hl.send_hidden(x, other)
                hl.recv_hidden(y)
            else:
                hl.recv_hidden(y)
                hl.send_hidden(y, other)

    def timed(fn, reps=20):
        fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (reps * a.rounds * 2)      # us per hop

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        pingpong()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s, capture_error_mode="thread_local"):
        pingpong()
    p2p_us = timed(g.replay)
    ok = bool(torch.equal(y, x)) if rank == 0 else True

    def nccl_pingpong():
        for _ in range(a.rounds):
            if rank == 0:
                dist.send(x, other)
                dist.recv(y, other)
            else:
                dist.recv(y, other)
                dist.send(y, other)

    nccl_us = timed(nccl_pingpong, reps=5)
    t = torch.tensor([p2p_us, nccl_us], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        res = {"hidden": a.hidden, "tokens": a.tokens, "bytes": a.tokens * a.hidden * 2,
               "hop_us_in_graph_peer_store": round(float(t[0]), 2), "hop_us_nccl_send_recv": round(float(t[1]), 2),
               "payload_intact": ok, "rounds_per_graph": a.rounds}
        print(json.dumps(res), flush=True)
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(res, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

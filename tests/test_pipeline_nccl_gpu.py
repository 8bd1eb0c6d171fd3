"""2-GPU NCCL chain (native kernels) must reproduce the single-GPU tokens exactly."""
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


def _prompts(vocab):
    g = torch.Generator().manual_seed(0)
    return [torch.randint(0, vocab, (n,), generator=g) for n in (40, 9)]


def _worker(rank, world, port, model, splits, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from petals_b200.parallel.pipeline import PipelineEngine

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    eng = PipelineEngine(model, splits, torch.bfloat16, dev, kv_capacity=256, max_sessions=4, seed=5)
    res = eng.generate(_prompts(eng.cfg.vocab), 10)
    if rank == 0:
        q.put(res[0])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_stage_nccl_chain_token_exact():
    sys.path.insert(0, ROOT)
    from petals_b200.parallel.pipeline import PipelineEngine

    model = "tiny-llama-gqa"
    torch.cuda.set_device(0)
    eng = PipelineEngine(model, [], torch.bfloat16, torch.device("cuda", 0), kv_capacity=256, max_sessions=4, seed=5)
    ref = eng.generate(_prompts(eng.cfg.vocab), 10)[0]
    del eng
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, model, [2], q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out == ref

"""PipelineEngine host logic with gloo on CPU (world size 2 and 3): the chain must reproduce
the single-process result token for token."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, model, splits, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from petals_b200.parallel.pipeline import PipelineEngine

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    eng = PipelineEngine(model, splits, torch.float32, torch.device("cpu"), kv_capacity=64, max_sessions=4, seed=5)
    g = torch.Generator().manual_seed(0)
    prompts = [torch.randint(0, eng.cfg.vocab, (n,), generator=g) for n in (5, 9, 3)]
    res = eng.generate(prompts, 6)
    # micro-batched prefill (3-token chunks flowing through the chain) must give the same first token
    eng.close_sessions()
    sid = eng.open_sessions(1)[0]
    first = eng.prefill(sid, prompts[1], micro_batch=3)
    if rank == 0:
        assert first == res[0][1][0], "micro-batched prefill changed the result"
    # micro-batches of two sessions travelling together must not change any token
    eng.close_sessions()
    ids3 = eng.open_sessions(len(prompts))
    cur = [eng.prefill(i, p) for i, p in zip(ids3, prompts)]
    grouped = [[c] for c in cur] if rank == 0 else None
    for _ in range(5):
        cur = eng.decode_round(cur, group=2)
        if rank == 0:
            for o, tkn in zip(grouped, cur):
                o.append(tkn)
    if rank == 0:
        assert grouped == res[0], "grouped micro-batches changed the tokens"
    res2 = eng.generate(prompts, 6, streaming=True)
    if rank == 0:
        assert res2[0] == res[0], "streaming rounds must produce the same tokens"
        q.put(res[0])
    dist.barrier()
    dist.destroy_process_group()


def _single(model):
    sys.path.insert(0, ROOT)
    from petals_b200.parallel.pipeline import PipelineEngine

    eng = PipelineEngine(model, [], torch.float32, torch.device("cpu"), kv_capacity=64, max_sessions=4, seed=5)
    g = torch.Generator().manual_seed(0)
    prompts = [torch.randint(0, eng.cfg.vocab, (n,), generator=g) for n in (5, 9, 3)]
    return eng.generate(prompts, 6)[0]


@pytest.mark.parametrize("model,splits", [("tiny-llama-gqa", [2]), ("tiny-gpt2", [1, 3])])
def test_chain_equals_single_process(model, splits):
    ref = _single(model)
    world = len(splits) + 1
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, model, splits, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out == ref

"""Multi-GPU pipeline chains (native kernels) must reproduce the single-GPU tokens exactly, whichever way the
hidden state travels: host-issued NCCL send/recv, or the peer-store hops captured into the stages' graphs, with
the LM head on the last stage or sharded over the two ends of the chain."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

MODES = [("nccl", False), ("p2p", False), ("p2p", True)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _prompts(vocab):
    g = torch.Generator().manual_seed(0)
    return [torch.randint(0, vocab, (n,), generator=g) for n in (40, 9, 17, 33)]


def _device_loop(eng, prompts, n_steps, group):
    """Prefill, then n_steps rounds with the tokens fed back on the device; returns [steps+1][sessions] on rank 0."""
    eng.close_sessions()
    ids = eng.open_sessions(len(prompts))
    first = [eng.prefill(i, p) for i, p in zip(ids, prompts)]
    if eng.is_first:
        eng._tok[:len(ids)].copy_(torch.tensor(first, dtype=torch.int64))
    rows = [first]
    for _ in range(n_steps):
        if eng.world == 1:
            rows.append(eng.decode_round(rows[-1]))
        else:
            eng.decode_round(None, to_host=False, group=group)
            eng.drain()
            if eng.is_first:
                rows.append([int(x) for x in eng._tok[:len(ids)].cpu()])
    return rows


def _sampled(eng, prompts):
    """Sampled generation (device-side sampler on the last stage, per-session Philox streams) and an EOS stop."""
    from petals_b200.runtime.sampler import SamplingParams

    sp = SamplingParams(0.9, 0.92, 50, 1.3)
    a = eng.generate(prompts[:2], 12, sampling=sp, seed=42)
    b = eng.generate(prompts[:2], 12, sampling=sp, seed=42)
    g = eng.generate(prompts[:2], 24, eos_token_id=None)
    eos = None
    if eng.is_first:
        assert a[0] == b[0], "same seed, same tokens"
        eos = g[0][0][5]                       # a token the greedy run is known to produce
    import torch.distributed as dist
    if dist.is_initialized():
        t = torch.tensor([eos if eos is not None else 0], device=eng.device)
        dist.broadcast(t, 0)
        eos = int(t.item())
    e = eng.generate(prompts[:2], 24, eos_token_id=eos, eos_check_every=2)
    if eng.is_first:
        return a[0], e[0][0], eos
    return None


def _worker(rank, world, port, model, splits, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from petals_b200.parallel.pipeline import PipelineEngine

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    out = {}
    for hop, split in MODES:
        eng = PipelineEngine(model, splits, torch.bfloat16, dev, kv_capacity=256, max_sessions=4, seed=5, hop=hop,
                             head_split=split)
        assert eng.hop == hop and eng.head_split == split
        prompts = _prompts(eng.cfg.vocab)
        res = eng.generate(prompts[:2], 10)
        res_s = eng.generate(prompts[:2], 10, streaming=True)
        loop1 = _device_loop(eng, prompts, 6, 1)
        loop2 = _device_loop(eng, prompts, 6, 2)
        samp = None
        if not split:      # sampling needs the whole vocabulary on the last stage
            samp = _sampled(eng, prompts)
        if rank == 0:
            out[(hop, split)] = (res[0], res_s[0], loop1, loop2, samp)
        dist.barrier()
        del eng
    if rank == 0:
        q.put(out)
    dist.barrier()
    dist.destroy_process_group()


def _reference(model):
    sys.path.insert(0, ROOT)
    from petals_b200.parallel.pipeline import PipelineEngine

    torch.cuda.set_device(0)
    eng = PipelineEngine(model, [], torch.bfloat16, torch.device("cuda", 0), kv_capacity=256, max_sessions=4, seed=5)
    prompts = _prompts(eng.cfg.vocab)
    ref = eng.generate(prompts[:2], 10)[0]
    loop = _device_loop(eng, prompts, 6, 1)
    samp = _sampled(eng, prompts)
    del eng
    return ref, loop, samp


def _chain(world, model, splits):
    ref, ref_loop, ref_samp = _reference(model)
    assert ref_samp[1][-1] == ref_samp[2] or ref_samp[2] not in ref_samp[1], "an EOS in the output ends it"
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, model, splits, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for mode, (res, res_s, loop1, loop2, samp) in out.items():
        if samp is not None:
            assert samp[0] == ref_samp[0], f"{mode}: sampled tokens differ from one GPU (same seed)"
            assert samp[1] == ref_samp[1], f"{mode}: EOS-stopped generation differs from one GPU"
        assert res == ref, f"{mode}: generate() differs from one GPU"
        assert res_s == ref, f"{mode}: streaming generate() differs from one GPU"
        assert loop1 == ref_loop, f"{mode}: device loop differs from one GPU"
        assert loop2 == ref_loop, f"{mode}: device loop with micro-batches of 2 differs from one GPU"


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_stage_chain_token_exact():
    _chain(2, "tiny-llama-gqa", [2])


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs 4 GPUs")
def test_four_stage_chain_token_exact():
    _chain(4, "tiny-llama-gqa", [1, 2, 3])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_stage_chain_gpt2_token_exact():
    _chain(2, "tiny-gpt2", [2])

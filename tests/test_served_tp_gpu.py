"""A served stage that is wider than one GPU: `--stage 1 --tp 2` (leader + follower process, fused GEMM+all-reduce
inside the span) must give the client exactly the tokens of a one-GPU server, over the TCP data plane, also with
several concurrent sessions batched into one step."""
import os
import socket
import sys
import threading

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


def _server_args(model, splits, stage, dht_port, **kw):
    sys.path.insert(0, ROOT)
    from src.main import build_parser

    a = build_parser().parse_args(["--model", model, "--splits", splits, "--stage", str(stage), "--dht_port", str(dht_port),
                                   "--rpc_port", "0", "--temperature", "0", "--dtype", "bf16"])
    a.dtype = torch.bfloat16
    a.mean_block_selection_delay = 0.0
    a.no_throughput_cache = True
    a.seed = 21
    a.kv_capacity = 256
    for k, v in kw.items():
        setattr(a, k, v)
    return a


def _tp_server(rank, world, master_port, dht_port, model, splits, ready):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from src.main import parse_splits, run_stage_server_fixed

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    args = _server_args(model, splits, 1, dht_port, tp=world, max_batch=4)
    if rank == 0:
        h = run_stage_server_fixed(args, dev, parse_splits(splits), block=False)
        ready.set()
        h.wait()
    else:
        run_stage_server_fixed(args, dev, parse_splits(splits))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_tensor_parallel_served_stage_matches_one_gpu():
    sys.path.insert(0, ROOT)
    from src.main import parse_splits, run_rank0, run_stage_server_fixed

    model, splits = "tiny-llama-gqa", "1"
    cuts = parse_splits(splits)
    dev = torch.device("cuda", 0)
    prompts = ["the quick brown fox", "a b c d e f g h i j k", "hello"]

    def client(peers, prompt):
        a = _server_args(model, splits, 0, 0, dht_initial_peers=peers, max_new_tokens=10, prompt=prompt, request_timeout=60.0)
        return run_rank0(a, dev, cuts)[0]

    # one-GPU server: the oracle
    one = run_stage_server_fixed(_server_args(model, splits, 1, 0), dev, cuts, block=False)
    try:
        ref = [client(one.dht.get_visible_maddrs()[0], p) for p in prompts]
    finally:
        one.stop()

    ctx = mp.get_context("spawn")
    ready = ctx.Event()
    master_port, dht_port = _free_port(), _free_port()
    procs = [ctx.Process(target=_tp_server, args=(r, 2, master_port, dht_port, model, splits, ready), daemon=True) for r in range(2)]
    for p in procs:
        p.start()
    try:
        assert ready.wait(300), "the tensor-parallel server did not come up"
        peers = f"/ip4/127.0.0.1/tcp/{dht_port}"
        out = [client(peers, p) for p in prompts]
        assert out == ref, "a tensor-parallel span must serve the tokens of a one-GPU span"
        # concurrent sessions: batched decode steps through the mirrored decode_batch
        res = [None] * len(prompts)

        def run(i):
            res[i] = client(peers, prompts[i])

        ts = [threading.Thread(target=run, args=(i,)) for i in range(len(prompts))]
        for t in ts:
            t.start()
        for t in ts:
            t.join(300)
        assert res == ref
    finally:
        for p in procs:
            p.terminate()
        for p in procs:
            p.join(10)

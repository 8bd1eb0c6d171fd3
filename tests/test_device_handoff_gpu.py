"""Served stages in one box hand hidden states over device to device (CUDA IPC arena + peer copy), with the TCP
messages carrying only a descriptor: same tokens as the byte path, in relay and in push mode, and a killed server is
still recovered from the (device-resident) journal."""
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


def _args(model, splits, stage, dht_port, **kw):
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


def _server(stage, gpu, dht_port, peers, model, splits, ready, die_flag=None):
    sys.path.insert(0, ROOT)
    from src.main import parse_splits, run_stage_server_fixed

    torch.cuda.set_device(gpu)
    dev = torch.device("cuda", gpu)
    a = _args(model, splits, stage, dht_port, dht_initial_peers=peers)
    h = run_stage_server_fixed(a, dev, parse_splits(splits), block=False)
    if die_flag is not None:        # scripted crash at the 5th request; only the first replica to get there dies

        def first_to_die():
            try:
                os.close(os.open(die_flag, os.O_CREAT | os.O_EXCL))
                return True
            except FileExistsError:
                return False

        h.handler.die_guard = first_to_die
        h.handler.die_after = 5
    ready.set()
    h.wait()


def test_arena_roundtrip_and_stale_slot_detection():
    sys.path.insert(0, ROOT)
    from petals_b200.net.devshare import DeviceExporter, DeviceImporter, StaleDeviceTensor

    dev = torch.device("cuda", 0)
    ex, im = DeviceExporter(dev, arena_bytes=1 << 16), DeviceImporter(dev)
    x = torch.randn(3, 5, 64, device=dev).to(torch.bfloat16)
    d = ex.export(x)
    assert im.accepts(d)
    y = im.pull(d, local=ex.arena)
    assert torch.equal(x, y) and y.dtype == x.dtype
    first = d
    for _ in range(40):                                   # wrap the 64 KiB ring over the first slot
        ex.export(torch.randn(1, 1024, device=dev))
    with pytest.raises(StaleDeviceTensor):
        im.pull(first, local=ex.arena)


@pytest.mark.parametrize("push", [False, True])
def test_served_chain_device_handoff_token_exact(push):
    sys.path.insert(0, ROOT)
    from petals_b200.net.rpc import RpcClient
    from src.main import parse_splits, run_rank0

    model, splits = "tiny-llama-gqa", "1,2"
    cuts = parse_splits(splits)
    n_gpu = torch.cuda.device_count()
    dev = torch.device("cuda", 0)
    ctx = mp.get_context("spawn")
    p1, p2 = _free_port(), _free_port()
    peers = f"/ip4/127.0.0.1/tcp/{p1}"
    r1, r2 = ctx.Event(), ctx.Event()
    procs = [ctx.Process(target=_server, args=(1, 1 % n_gpu, p1, "", model, splits, r1), daemon=True)]
    procs[0].start()
    try:
        assert r1.wait(120)
        procs.append(ctx.Process(target=_server, args=(2, 2 % n_gpu, p2, peers, model, splits, r2), daemon=True))
        procs[1].start()
        assert r2.wait(120)

        def client(handoff):
            a = _args(model, splits, 0, 0, dht_initial_peers=peers, max_new_tokens=12, prompt="the quick brown fox jumps",
                      request_timeout=60.0, device_handoff=handoff, push=push, no_repeat_stop=True,
                      device_handoff_min_kb=0)      # every tensor, single tokens included (default: prompts only)
            return run_rank0(a, dev, cuts)

        ref, _ = client(False)
        out, stats = client(True)
        assert out == ref, "device hand-off must not change a token"
        # the servers really pulled their inputs from a peer arena instead of reading them from the socket
        from petals_b200.net.dht import DHT, parse_maddr
        from src.dht_utils import get_stage_entries

        c = RpcClient()
        d = DHT(start=True, initial_peers=[peers], client_mode=True)
        try:
            for stage in (1, 2):
                entries = get_stage_entries(d, stage)
                assert entries, f"stage {stage} is announced"
                host, port, _ = parse_maddr(entries[0]["p2p_maddrs"][0])
                _, info = c.call((host, port), "StageConnectionHandler.rpc_info", f"mini_petals:stage{stage}", [], {})
                assert info["box"], "a CUDA server names its box"
                assert info["handoff_in"] >= 12, info          # inputs pulled from a peer arena, not read off the socket
                if stage == 1:
                    assert info["handoff_out"] >= 12, info
        finally:
            c.close()
            d.shutdown()
    finally:
        for p in procs:
            p.terminate()
        for p in procs:
            p.join(10)


def test_killed_replica_is_recovered_from_the_device_journal(tmp_path):
    sys.path.insert(0, ROOT)
    from src.main import parse_splits, run_rank0

    model, splits = "tiny-llama-gqa", "1,2"
    cuts = parse_splits(splits)
    n_gpu = torch.cuda.device_count()
    dev = torch.device("cuda", 0)
    ctx = mp.get_context("spawn")
    ports = [_free_port() for _ in range(3)]
    peers = f"/ip4/127.0.0.1/tcp/{ports[0]}"
    flag = str(tmp_path / "died")
    procs, events = [], []
    try:
        # stage 2 is the bootstrap (it survives); two replicas of stage 1, one of which will crash mid-generation
        for i, (stage, die) in enumerate([(2, None), (1, flag), (1, flag)]):
            ev = ctx.Event()
            p = ctx.Process(target=_server, args=(stage, (i + 1) % n_gpu, ports[i], "" if i == 0 else peers, model, splits,
                                                  ev, die), daemon=True)
            p.start()
            procs.append(p)
            events.append(ev)
            assert ev.wait(120)

        def client(handoff):
            a = _args(model, splits, 0, 0, dht_initial_peers=peers, max_new_tokens=14, prompt="one two three four five",
                      request_timeout=20.0, device_handoff=handoff, no_repeat_stop=True, device_handoff_min_kb=0)
            return run_rank0(a, dev, cuts)

        out, stats = client(True)
        assert os.path.exists(flag), "one replica crashed"
        assert stats["recoveries"] >= 1
        ref, _ = client(False)          # the byte path through the surviving replica: the oracle
        assert out == ref, "recovery over the device path must be token-exact"
    finally:
        for p in procs:
            p.terminate()
        for p in procs:
            p.join(10)


def test_served_stage_with_mx_fp8_weights():
    """`--weight_format mxfp8` (upstream: --quant_type): the served span holds block-scaled e4m3 weights - about half
    the bytes - and serves a deterministic greedy continuation over the TCP path."""
    sys.path.insert(0, ROOT)
    from src.main import parse_splits, run_rank0, run_stage_server_fixed

    model, splits = "tiny-llama-gqa", "1"
    cuts = parse_splits(splits)
    dev = torch.device("cuda", 0)
    handles, sizes, outs = [], {}, {}
    try:
        for wf in ("native", "mxfp8"):
            h = run_stage_server_fixed(_args(model, splits, 1, 0, weight_format=wf), dev, cuts, block=False)
            handles.append(h)
            span = h.handler.stage_model.span
            assert span.weight_format == wf
            sizes[wf] = sum(P[k].nbytes for P in span.layers for k in ("qkv", "o", "mlp_in", "down"))
            peers = h.dht.get_visible_maddrs()[0]
            runs = [run_rank0(_args(model, splits, 0, 0, dht_initial_peers=peers, max_new_tokens=10, prompt="a b c d e f",
                                    request_timeout=60.0, no_repeat_stop=True), dev, cuts)[0] for _ in range(2)]
            assert runs[0] == runs[1] and len(runs[0]) == 10
            outs[wf] = runs[0]
            h.stop()
            handles.remove(h)
        assert sizes["mxfp8"] < 0.6 * sizes["native"]
    finally:
        for h in handles:
            h.stop()

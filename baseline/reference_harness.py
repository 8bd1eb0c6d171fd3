"""Reference arm: the reference's own stage classes, decoder block, handler and sampler, unmodified
from baseline/_ref, driven over a loopback TCP transport.

What is the reference's and what is not
---------------------------------------
* reference code on the timed path (imported from baseline/_ref, not edited):
  ``src.llama_partition.Stage0 / StageSegment / StageLast``, ``petals.llama.block.
  OptimizedLlamaDecoderLayer`` (incl. its CUDA-graphed RoPE/RMSNorm), ``src.rpc_handler.
  StageConnectionHandler._run_forward`` (session KV dict, position ids, ``.item()`` activation guard,
  ``.cpu()`` + serialisation of the result) and ``_sample_token``.
* NOT the reference: its transport.  ``src.main`` / ``src.rpc_transport`` import hivemind 1.1.11
  (+ the Go p2pd daemon) at module top and cannot be installed without network, so the hop is a
  plain TCP socket on 127.0.0.1 that reproduces the reference's data path per hop — GPU -> ``.cpu()``
  -> bytes -> socket -> tensor -> ``.to(device)``, star topology through the client, one token per
  round trip.  Loopback TCP is FASTER than libp2p, so this flatters the reference.
* compatibility shims only: transformers 5.x removed four attributes the reference's block reads
  (``num_heads``, ``num_key_value_heads``, ``hidden_size``, ``rotary_emb``) and ``pretraining_tp``;
  they are set on the objects.  Weights are random-init of the same architecture
  (no checkpoints offline), ``--keep_layers_on_gpu`` = span length (best case: resident weights;
  the as-shipped default re-streams every layer over PCIe per token).

N GPUs: rank 0 is the client (embedding + first 32/N blocks), ranks 1..N-1 are server stages
(the last one samples); N = 1 runs the client half and one last-stage server thread on the same GPU
(the reference cannot run without at least one server stage).

Sessions: like the other arm, ``--sessions`` (default: one per GPU) independent batch-1 client sessions run
concurrently, each with its own session id, its own sockets to every stage and its own client-side KV; a
stage server accepts all of them and serialises ``_run_forward`` calls the way the reference's single event
loop does, so while session A is on stage 2 session B can be on stage 1.  The client half (``Stage0``) is
one module on rank 0's GPU shared by the session threads under a lock (the reference would start one client
process per session, each with its own copy of those blocks on that GPU).
"""
from __future__ import annotations

import io
import json
import logging
import os
import pickle
import socket
import statistics
import struct
import sys
import threading
import time
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


# ----------------------------------------------------------------------------- hivemind stand-ins
class _Tensor:   # runtime_pb2.Tensor stand-in: raw bytes + dtype + shape (hivemind's uncompressed codec)
    def __init__(self, buffer, dtype, size):
        self.buffer, self.dtype, self.size = buffer, dtype, size


def _install_stubs():
    import torch

    class PeerID(str):
        @classmethod
        def from_base58(cls, s):
            return cls(s)

        def to_base58(self):
            return str(self)

    def serialize_torch_tensor(t, *a, **k):
        t = t.detach().contiguous()
        if t.dtype == torch.bfloat16:       # hivemind up-casts bf16 on the wire
            t = t.to(torch.float32)
        return _Tensor(t.numpy().tobytes(), str(t.dtype), tuple(t.shape))

    def deserialize_torch_tensor(m):
        import numpy as np

        dt = getattr(torch, m.dtype.replace("torch.", ""))
        arr = np.frombuffer(m.buffer, dtype=np.uint8).copy()
        return torch.from_numpy(arr).view(dt).view(m.size)

    class MSGPackSerializer:
        @staticmethod
        def dumps(o):
            import msgpack

            return msgpack.packb(o, use_bin_type=True)

        @staticmethod
        def loads(b):
            import msgpack

            return msgpack.unpackb(b, raw=False)

    class ExpertResponse:
        def __init__(self, tensors=None, metadata=b""):
            self.tensors, self.metadata = list(tensors or []), metadata

    class ExpertRequest(ExpertResponse):
        def __init__(self, uid="", tensors=None, metadata=b""):
            super().__init__(tensors, metadata)
            self.uid = uid

    class ConnectionHandler:
        def __init__(self, *a, **k):
            pass

    def mk(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class _D:
        def __init__(self, *a, **k):
            pass

    mk("hivemind", DHT=_D, PeerID=PeerID, get_logger=logging.getLogger, get_dht_time=time.time, P2PContext=_D,
       deserialize_torch_tensor=deserialize_torch_tensor, serialize_torch_tensor=serialize_torch_tensor,
       MSGPackSerializer=MSGPackSerializer)
    mk("hivemind.p2p", P2P=_D, P2PContext=_D, PeerID=PeerID)
    mk("hivemind.utils", get_logger=logging.getLogger)
    mk("hivemind.utils.logging", get_logger=logging.getLogger)
    mk("hivemind.compression", serialize_torch_tensor=serialize_torch_tensor,
       deserialize_torch_tensor=deserialize_torch_tensor)
    mk("hivemind.compression.serialization", serialize_torch_tensor=serialize_torch_tensor,
       deserialize_torch_tensor=deserialize_torch_tensor, deserialize_tensor_stream=None)
    mk("hivemind.moe")
    mk("hivemind.moe.server")
    mk("hivemind.moe.server.connection_handler", ConnectionHandler=ConnectionHandler)
    mk("hivemind.proto", runtime_pb2=types.SimpleNamespace(ExpertRequest=ExpertRequest, ExpertResponse=ExpertResponse,
                                                           Tensor=_Tensor))
    mk("async_timeout", timeout=_D)
    return serialize_torch_tensor, deserialize_torch_tensor, MSGPackSerializer


# ----------------------------------------------------------------------------- loopback transport
def _send(sock, obj):
    data = pickle.dumps(obj, protocol=4)
    sock.sendall(struct.pack("<Q", len(data)) + data)


def _recv(sock):
    hdr = b""
    while len(hdr) < 8:
        c = sock.recv(8 - len(hdr))
        if not c:
            raise ConnectionError("closed")
        hdr += c
    n = struct.unpack("<Q", hdr)[0]
    buf = bytearray()
    while len(buf) < n:
        c = sock.recv(min(1 << 20, n - len(buf)))
        if not c:
            raise ConnectionError("closed")
        buf += c
    return pickle.loads(bytes(buf))


def _serve(handler, port_holder, ready, stop, deserialize, msgpack_ser, device=None, done=None):
    if device is not None:
        import torch

        torch.cuda.set_device(device)     # the current CUDA device is per host thread
    srv = socket.socket()
    srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    srv.bind(("127.0.0.1", port_holder[0]))
    srv.listen(64)
    port_holder[0] = srv.getsockname()[1]
    ready.set()
    srv.settimeout(0.5)
    fwd_lock = threading.Lock()           # one _run_forward at a time, like the reference's event loop
    state = {"open": 0, "seen": 0}
    state_lock = threading.Lock()

    def connection(conn):
        if device is not None:
            import torch

            torch.cuda.set_device(device)
        try:
            while True:
                tensor_msg, metadata = _recv(conn)
                hidden = deserialize(tensor_msg)
                with fwd_lock:
                    resp = handler._run_forward(hidden, msgpack_ser.loads(metadata))      # <- reference code
                _send(conn, (resp.tensors[0], resp.metadata))
        except (ConnectionError, OSError):
            conn.close()
        with state_lock:
            state["open"] -= 1
            if state["open"] == 0 and done is not None:
                done.set()                # every client hung up: this stage's work is over

    while not stop.is_set():
        try:
            conn, _ = srv.accept()
        except socket.timeout:
            continue
        conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        with state_lock:
            state["open"] += 1
            state["seen"] += 1
        threading.Thread(target=connection, args=(conn,), daemon=True).start()
    srv.close()


# ----------------------------------------------------------------------------- model pieces
def _hf_span(cfg_dict, n_layers, device, dtype, with_embed, with_head):
    import torch
    import transformers

    c = transformers.LlamaConfig(**{k: v for k, v in cfg_dict.items() if k not in ("model_type", "architectures")})
    c.num_hidden_layers = n_layers
    c.pretraining_tp = 1
    c._attn_implementation = "eager"
    with torch.device(device):
        m = transformers.LlamaForCausalLM(c).to(dtype).eval()
    return m, c


def _shim(stage, cfg):
    from transformers.models.llama.modeling_llama import LlamaRotaryEmbedding

    for layer in stage.layers:
        a = layer.self_attn
        a.num_heads, a.num_key_value_heads, a.hidden_size = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.hidden_size
        a.rotary_emb = LlamaRotaryEmbedding(cfg).to(next(layer.parameters()).device)


def run(args) -> None:
    import torch
    import torch.distributed as dist

    if not os.path.isdir(os.path.join(REF, "src")):
        import subprocess

        subprocess.run([sys.executable, os.path.join(HERE, "install_reference.py")], check=False, capture_output=True)
    if not os.path.isdir(os.path.join(REF, "src")):
        raise RuntimeError("baseline/_ref is empty: run `python baseline/install_reference.py` first")
    serialize, deserialize, mser = _install_stubs()
    sys.path.insert(0, REF)
    import src.llama_partition as lp                      # reference, unmodified
    from src.rpc_handler import StageConnectionHandler    # reference, unmodified

    sys.path.insert(0, os.path.dirname(HERE))
    from bench import ClockSampler, even_splits
    from petals_b200.models.config import resolve_config

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    dtype = torch.float16 if args.dtype == "fp16" else torch.bfloat16
    cfg = resolve_config(args.model)
    hf_dict = cfg.to_hf_dict()
    L = cfg.n_layers
    # stage layout: N=1 -> client [0, L/2) + one last stage [L/2, L) on the same GPU
    n_stages = max(world, 2)
    bounds = [0] + even_splits(L, n_stages) + [L]
    K, W, P = args.steps, args.warmup, args.prompt_len
    torch.manual_seed(0)

    def make_server(si, device):
        start, end = bounds[si], bounds[si + 1]
        last = si == n_stages - 1
        full, c = _hf_span(hf_dict, end - start, device, dtype, False, last)
        if last:
            stage = lp.StageLast(full, start, gpu_device=device, keep_layers_on_gpu=end - start)
        else:
            stage = lp.StageSegment(full, start, end, gpu_device=device, keep_layers_on_gpu=end - start)
        _shim(stage, c)
        h = StageConnectionHandler(None, stage, device, request_timeout=30.0, final_stage=last)
        return h

    stop = threading.Event()
    done = threading.Event()
    ports = [[0] for _ in range(n_stages)]

    def watchdog():
        # never leave the driver waiting on a wedged reference run
        if not stop.wait(float(os.environ.get("REFERENCE_ARM_TIMEOUT", "420"))):
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "reference harness timed out"}), flush=True)
            os._exit(0)

    threading.Thread(target=watchdog, daemon=True).start()
    threads = []
    if world == 1:
        h = make_server(1, dev)
        ready = threading.Event()
        t = threading.Thread(target=_serve, args=(h, ports[1], ready, stop, deserialize, mser, dev, done), daemon=True)
        t.start(); ready.wait(30); threads.append(t)
        port_list = [0, ports[1][0]]
    else:
        port_list = [0] * n_stages
        if rank > 0:
            h = make_server(rank, dev)
            ready = threading.Event()
            t = threading.Thread(target=_serve, args=(h, ports[rank], ready, stop, deserialize, mser, dev, done), daemon=True)
            t.start(); ready.wait(30); threads.append(t)
        pt = torch.tensor([ports[rank][0] if rank > 0 else 0], device=dev, dtype=torch.int64)
        allp = [torch.zeros_like(pt) for _ in range(world)]
        dist.all_gather(allp, pt)
        port_list = [int(x.item()) for x in allp]

    result = None
    if rank == 0:
        import hashlib

        full0, c0 = _hf_span(hf_dict, bounds[1], dev, dtype, True, False)
        s0 = lp.Stage0(full0, bounds[1]).to(dev)
        _shim(s0, c0)
        S = args.sessions or world
        s0_lock = threading.Lock()
        g = torch.Generator().manual_seed(1234)
        max_len = max(P + 2 * (K + W) + 16, 2048 + 64)

        class Session:
            def __init__(self, i, prompt_len):
                self.sid = f"bench-session-{i}"
                self.socks = []
                for si in range(1, n_stages):
                    sk = socket.create_connection(("127.0.0.1", port_list[si]))
                    sk.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                    self.socks.append(sk)
                self.prompt = torch.randint(0, cfg.vocab, (1, prompt_len), generator=g).pin_memory()
                self.generated = []
                self.tok_host = torch.zeros(1, 1, dtype=torch.long).pin_memory()
                self.lat = []

            def hops(self, hidden, md):
                cur = hidden.cpu()                                             # reference: rpc_transport.py:724
                for sock in self.socks:
                    _send(sock, (serialize(cur), mser.dumps(md)))
                    tensor_msg, rmd = _recv(sock)
                    out = deserialize(tensor_msg)
                    rmd = mser.loads(rmd)
                    if "token_id" in rmd:
                        return int(rmd["token_id"])
                    cur = out
                raise RuntimeError("no token returned")

            def prefill(self):
                Pn = self.prompt.shape[1]
                with s0_lock:
                    ids = self.prompt.to(dev, non_blocking=True)
                    pos = torch.arange(Pn, device=dev).unsqueeze(0)
                    hidden, self.past0 = s0(ids, pos, None, None, use_cache=True)
                md = dict(session_id=self.sid, seq_len=Pn, cur_len=Pn, is_prefill=True, max_length=max_len,
                          temperature=0.0, top_p=0.92, top_k=50)
                self.nxt = self.hops(hidden, md)
                self.generated.append(self.nxt)
                self.cur_len = Pn + 1

            def step(self):
                with s0_lock:
                    self.tok_host[0, 0] = self.nxt
                    new = self.tok_host.to(dev, non_blocking=True)               # H2D of the step's input
                    past_len = self.past0[0][0].shape[-2]
                    pos = torch.tensor([[past_len]], device=dev, dtype=torch.long)
                    hidden, self.past0 = s0(new, pos, None, self.past0, use_cache=True)
                md = dict(session_id=self.sid, seq_len=1, cur_len=self.cur_len, is_prefill=False, max_length=max_len,
                          temperature=0.0, top_p=0.92, top_k=50, generated_tokens=self.generated[-50:])
                self.nxt = self.hops(hidden, md)                                 # D2H of hidden + token back
                self.generated.append(self.nxt)
                self.cur_len += 1

            def close(self):
                for sk in self.socks:
                    sk.close()

        def run_steps(sessions, n, record):
            def work(se):
                torch.cuda.set_device(dev)
                with torch.inference_mode():
                    t_prev = time.perf_counter()
                    for _ in range(n):
                        se.step()
                        now = time.perf_counter()
                        if record:
                            se.lat.append(now - t_prev)
                        t_prev = now
            if len(sessions) == 1:
                work(sessions[0])
                return
            ts = [threading.Thread(target=work, args=(se,)) for se in sessions]
            for t in ts:
                t.start()
            for t in ts:
                t.join()

        with torch.inference_mode():
            sessions = [Session(i, P) for i in range(S)]
            for se in sessions:
                se.prefill()
            run_steps(sessions, W, False)
            torch.cuda.synchronize()
            sampler = ClockSampler(local)
            sampler.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run_steps(sessions, K, True)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            clocks = sampler.stop()
            # time to first token: a fresh session's prompt through Stage0 and every server stage, token back on the host
            ttft = {}
            for plen in (128, 2048):
                try:
                    times = []
                    for rep in range(3):
                        se = Session(1000 + plen * 10 + rep, plen)
                        torch.cuda.synchronize()
                        t0 = time.perf_counter()
                        se.prefill()
                        times.append((time.perf_counter() - t0) * 1e3)
                        se.close()
                    ttft[str(plen)] = round(statistics.median(times[1:]), 3)
                except Exception as e:  # noqa: BLE001  (the decode line must survive a failing prompt length)
                    ttft[str(plen)] = f"failed: {type(e).__name__}: {e}"[:120]
        H = cfg.hidden
        esz = 2
        lat0 = sessions[0].lat
        sha = hashlib.sha256(",".join(str(t) for t in sessions[0].generated[-K:]).encode()).hexdigest()[:16]
        result = {
            "metric": "decode tokens/sec (aggregate over in-flight sessions), "
                      + ("Llama-2-7B" if args.model == "llama-2-7b" else args.model) + " N-stage pipeline",
            "value": round(S * K / (ms / 1e3), 2), "unit": "tokens/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic prompts, random-init weights", "impl": "reference",
            "config": {"model": args.model, "global_batch": S, "seq_len": P, "parallelism": f"pp{n_stages}",
                       "sessions_in_flight": S, "batch_per_session": 1,
                       "keep_layers_on_gpu": "span (best case, resident weights)",
                       "transport": "loopback TCP stand-in for hivemind/libp2p (not installable offline)",
                       "p50_token_latency_ms": round(statistics.median(lat0) * 1e3, 4),
                       "single_session_tokens_per_s": round(1.0 / statistics.median(lat0), 2),
                       "ttft_ms": ttft, "tokens_sha": sha,
                       "note": "reference Stage*/OptimizedLlamaDecoderLayer/_run_forward/_sample_token run unmodified "
                               "from baseline/_ref; N=1 runs client half + one last-stage server on the same GPU"},
            "clocks": clocks,
            "e2e": {"value": round(S * K / (ms / 1e3), 2), "unit": "tokens/s",
                    "h2d_bytes_per_step": S * (8 + (n_stages - 1) * H * esz),
                    "d2h_bytes_per_step": S * ((n_stages - 1) * H * esz + 8), "ms_per_step": round(ms / K, 4),
                    "note": "the reference has no device-only loop: value == e2e"},
            "gpu_launches": 0,
        }
        for se in sessions:
            se.close()
        print(json.dumps(result), flush=True)
    elif world > 1:
        done.wait()          # serve until the client closes the connection (no collective while serving:
                             # a blocking NCCL barrier here would starve the server thread)
    stop.set()
    sys.stdout.flush()
    os._exit(0)              # skip process-group teardown: ranks finish at different times by design

"""Framed-TCP RPC carrying tensors: the cross-host data plane (fallback to NCCL in a box).

The reference talks hivemind ``P2P.call_protobuf_handler`` / ``iterate_protobuf_handler``
with protobuf ``ExpertRequest{uid, tensors, metadata(msgpack)}`` messages over libp2p
(src/rpc_transport.py:524-562; src/rpc_handler.py:405-464).  Neither hivemind nor its Go
daemon exists here, so the same contract rides on plain TCP:

    frame  := magic u32 | header_len u32 | payload_len u64 | header (msgpack) | payload (raw)
    header := {method, uid, metadata{...}, tensors[{dtype, shape, nbytes}], error?}

Tensor bytes travel in their own dtype (bf16 stays 2 bytes; hivemind up-casts it to fp32).
Handlers are registered under hivemind-style names (``"StageConnectionHandler.rpc_forward"``).
The server accepts connections on an asyncio loop in a background thread and runs the
handlers on a single worker thread, so GPU work never blocks heartbeats.
Large payloads need no separate streaming RPC on TCP; ``rpc_forward_stream`` is the same
handler under a second name for wire-level parity.

Streams: a method registered with ``add_stream_handler`` owns its connection for as long as it
runs - the handler gets a ``stream`` object with ``recv(timeout)`` / ``send(tensors, metadata)``
and exchanges any number of frames with the client before it returns (its return value is the
closing frame).  This carries the per-session ``rpc_inference`` protocol of upstream Petals
(petals/server/handler.py:132-195); the client side is ``RpcClient.open_stream``.
"""
from __future__ import annotations

import asyncio
import concurrent.futures as cf
import socket
import struct
import threading
from typing import Any, Callable, Dict, List, Optional, Tuple

import msgpack
import numpy as np
import torch

MAGIC = 0xB200_0001
_HDR = struct.Struct("<IIQ")
MAX_FRAME = 4 << 30

_DT = {"float16": torch.float16, "bfloat16": torch.bfloat16, "float32": torch.float32, "int64": torch.int64,
       "int32": torch.int32, "uint8": torch.uint8, "float64": torch.float64}
_DT_NAME = {v: k for k, v in _DT.items()}


class RpcError(RuntimeError):
    """Remote handler raised, or the frame was malformed."""


def serialize_torch_tensor(t: torch.Tensor) -> Tuple[dict, bytes]:
    t = t.detach().contiguous().cpu()
    raw = t.view(torch.uint8).numpy().tobytes() if t.numel() else b""
    return {"dtype": _DT_NAME[t.dtype], "shape": list(t.shape), "nbytes": len(raw)}, raw


def deserialize_torch_tensor(desc: dict, raw: bytes) -> torch.Tensor:
    dt = _DT[desc["dtype"]]
    if not raw:
        return torch.empty(desc["shape"], dtype=dt)
    arr = np.frombuffer(raw, dtype=np.uint8).copy()
    return torch.from_numpy(arr).view(dt).view(desc["shape"])


def pack_frame(method: str, uid: str, metadata: dict, tensors: List[torch.Tensor], error: Optional[str] = None) -> bytes:
    descs, blobs = [], []
    for t in tensors:
        d, raw = serialize_torch_tensor(t)
        descs.append(d)
        blobs.append(raw)
    header = {"method": method, "uid": uid, "metadata": metadata, "tensors": descs}
    if error is not None:
        header["error"] = error
    hb = msgpack.packb(header, use_bin_type=True)
    payload = b"".join(blobs)
    return _HDR.pack(MAGIC, len(hb), len(payload)) + hb + payload


def unpack_frame(hb: bytes, payload: bytes) -> Tuple[dict, List[torch.Tensor]]:
    header = msgpack.unpackb(hb, raw=False, strict_map_key=False)
    tensors, off = [], 0
    for d in header.get("tensors", []):
        n = int(d["nbytes"])
        tensors.append(deserialize_torch_tensor(d, payload[off:off + n]))
        off += n
    return header, tensors


Handler = Callable[[str, List[torch.Tensor], dict], Tuple[List[torch.Tensor], dict]]


class RpcServer:
    def __init__(self, host: str = "127.0.0.1", port: int = 0):
        self.host, self.port = host, port
        self.handlers: Dict[str, Handler] = {}
        self._loop: Optional[asyncio.AbstractEventLoop] = None
        self._thread: Optional[threading.Thread] = None
        self._server = None
        # Handlers mostly WAIT (for the stage's task pool, for the next hop): many may be in flight at once, the
        # compute itself is serialised - and batched - by the stage's own task pool thread.
        self._pool = cf.ThreadPoolExecutor(max_workers=64, thread_name_prefix="rpc-handler")
        self._stream_pool = cf.ThreadPoolExecutor(max_workers=256, thread_name_prefix="rpc-stream")
        self.stream_handlers: Dict[str, Callable] = {}
        self._ready = threading.Event()
        self.drop_next = 0          # fault injection: close the connection instead of answering
        self.requests_served = 0

    def add_handler(self, name: str, fn: Handler) -> None:
        self.handlers[name] = fn

    def add_stream_handler(self, name: str, fn: Callable) -> None:
        """``fn(uid, tensors, metadata, stream) -> (tensors, metadata)``; see the module docstring."""
        self.stream_handlers[name] = fn

    def start(self) -> "RpcServer":
        self._thread = threading.Thread(target=self._run, name="rpc-server", daemon=True)
        self._thread.start()
        if not self._ready.wait(10):
            raise RuntimeError("RPC server failed to start")
        return self

    def _run(self) -> None:
        self._loop = asyncio.new_event_loop()
        asyncio.set_event_loop(self._loop)

        async def boot():
            bind = "0.0.0.0" if self.host in ("0.0.0.0", "") else self.host
            self._server = await asyncio.start_server(self._serve_conn, bind, self.port)
            self.port = self._server.sockets[0].getsockname()[1]
            self._ready.set()

        self._loop.run_until_complete(boot())
        try:
            self._loop.run_forever()
        finally:
            self._loop.close()

    async def _serve_conn(self, reader: asyncio.StreamReader, writer: asyncio.StreamWriter) -> None:
        try:
            while True:
                raw = await reader.readexactly(_HDR.size)
                magic, hl, pl = _HDR.unpack(raw)
                if magic != MAGIC or hl + pl > MAX_FRAME:
                    break
                hb = await reader.readexactly(hl)
                payload = await reader.readexactly(pl) if pl else b""
                if self.drop_next > 0:
                    self.drop_next -= 1
                    break
                header, tensors = unpack_frame(hb, payload)
                sfn = self.stream_handlers.get(header.get("method", ""))
                if sfn is not None:
                    stream = _ServerStream(self._loop, reader, writer, header["method"], header.get("uid", ""))
                    try:
                        res_t, res_md = await self._loop.run_in_executor(
                            self._stream_pool, sfn, header.get("uid", ""), tensors, header.get("metadata") or {}, stream)
                        out = pack_frame(header["method"], header.get("uid", ""), res_md, res_t)
                    except Exception as e:  # noqa: BLE001
                        out = pack_frame(header["method"], header.get("uid", ""), {}, [], error=f"{type(e).__name__}: {e}")
                    self.requests_served += 1
                    writer.write(out)
                    await writer.drain()
                    if stream.broken:
                        break
                    continue
                fn = self.handlers.get(header.get("method", ""))
                if fn is None:
                    out = pack_frame(header.get("method", ""), header.get("uid", ""), {}, [],
                                     error=f"no handler {header.get('method')!r}")
                else:
                    try:
                        res_t, res_md = await self._loop.run_in_executor(
                            self._pool, fn, header.get("uid", ""), tensors, header.get("metadata") or {})
                        out = pack_frame(header["method"], header.get("uid", ""), res_md, res_t)
                    except Exception as e:  # noqa: BLE001 - forwarded to the caller
                        out = pack_frame(header["method"], header.get("uid", ""), {}, [],
                                         error=f"{type(e).__name__}: {e}")
                self.requests_served += 1
                writer.write(out)
                await writer.drain()
        except (asyncio.IncompleteReadError, ConnectionError, OSError):
            pass
        finally:
            try:
                writer.close()
            except Exception:  # noqa: BLE001
                pass

    def stop(self) -> None:
        if self._loop is None:
            return

        loop = self._loop

        async def _shutdown():
            if self._server is not None:
                self._server.close()
                try:
                    await asyncio.wait_for(self._server.wait_closed(), 1.0)
                except Exception:  # noqa: BLE001
                    pass
            tasks = [t for t in asyncio.all_tasks() if t is not asyncio.current_task()]
            for t in tasks:
                t.cancel()
            await asyncio.gather(*tasks, return_exceptions=True)
            loop.stop()

        asyncio.run_coroutine_threadsafe(_shutdown(), loop)
        if self._thread is not None:
            self._thread.join(timeout=5)
        self._pool.shutdown(wait=False, cancel_futures=True)
        self._stream_pool.shutdown(wait=False, cancel_futures=True)
        self._loop = None

    @property
    def address(self) -> Tuple[str, int]:
        return ("127.0.0.1" if self.host in ("0.0.0.0", "") else self.host), self.port


class _ServerStream:
    """The handler thread's view of its connection: blocking frame I/O executed on the server's event loop."""

    def __init__(self, loop, reader, writer, method: str, uid: str):
        self._loop, self._reader, self._writer, self.method, self.uid = loop, reader, writer, method, uid
        self.broken = False

    async def _read(self):
        raw = await self._reader.readexactly(_HDR.size)
        magic, hl, pl = _HDR.unpack(raw)
        if magic != MAGIC or hl + pl > MAX_FRAME:
            raise ConnectionError("bad frame")
        hb = await self._reader.readexactly(hl)
        payload = await self._reader.readexactly(pl) if pl else b""
        return unpack_frame(hb, payload)

    async def _write(self, frame: bytes):
        self._writer.write(frame)
        await self._writer.drain()

    def recv(self, timeout: Optional[float] = None) -> Tuple[List[torch.Tensor], dict]:
        fut = asyncio.run_coroutine_threadsafe(self._read(), self._loop)
        try:
            header, tensors = fut.result(timeout)
        except cf.TimeoutError:
            fut.cancel()
            self.broken = True          # a frame may be half read: the connection cannot be reused
            raise TimeoutError("no message on the stream") from None
        except (asyncio.IncompleteReadError, ConnectionError, OSError, asyncio.CancelledError) as e:
            self.broken = True
            raise ConnectionError(f"stream closed: {e}") from None
        return tensors, header.get("metadata") or {}

    def send(self, tensors: List[torch.Tensor], metadata: dict) -> None:
        frame = pack_frame(self.method, self.uid, metadata, tensors)
        try:
            asyncio.run_coroutine_threadsafe(self._write(frame), self._loop).result(30)
        except Exception as e:  # noqa: BLE001
            self.broken = True
            raise ConnectionError(f"stream closed: {e}") from None


class ClientStream:
    """Client end of a stream method: its own connection, frames in both directions until ``close``."""

    def __init__(self, sock: socket.socket, method: str, uid: str):
        self._s, self.method, self.uid = sock, method, uid
        self.closed = False

    def send(self, tensors: List[torch.Tensor], metadata: dict) -> None:
        self._s.sendall(pack_frame(self.method, self.uid, metadata, tensors))

    def recv(self, timeout: Optional[float] = 30.0) -> Tuple[List[torch.Tensor], dict]:
        self._s.settimeout(timeout)
        try:
            raw = _recv_exact(self._s, _HDR.size)
            magic, hl, pl = _HDR.unpack(raw)
            if magic != MAGIC:
                raise ConnectionError("bad frame magic")
            hb = _recv_exact(self._s, hl)
            payload = _recv_exact(self._s, pl) if pl else b""
        except socket.timeout as e:
            raise TimeoutError(f"stream {self.method}: no answer in {timeout}s") from e
        except OSError as e:
            raise ConnectionError(f"stream {self.method} failed: {e}") from e
        header, out = unpack_frame(hb, payload)
        if header.get("error"):
            raise RpcError(header["error"])
        md = header.get("metadata") or {}
        if md.get("error"):
            raise RpcError(md["error"])
        return out, md

    def call(self, tensors: List[torch.Tensor], metadata: dict, timeout: float = 30.0):
        self.send(tensors, metadata)
        return self.recv(timeout)

    def close(self, timeout: float = 5.0) -> Optional[dict]:
        if self.closed:
            return None
        self.closed = True
        md = None
        try:
            self.send([], {"close": True})
            _, md = self.recv(timeout)
        except Exception:  # noqa: BLE001 - best effort
            pass
        try:
            self._s.close()
        except OSError:
            pass
        return md


class RpcClient:
    """Blocking client with one pooled connection per server address."""

    def open_stream(self, addr: Tuple[str, int], method: str, uid: str, tensors: List[torch.Tensor], metadata: dict,
                    timeout: float = 30.0) -> ClientStream:
        """Starts a stream method on its own connection and returns after the server's opening answer."""
        try:
            s = socket.create_connection(addr, timeout=timeout)
        except OSError as e:
            raise ConnectionError(f"RPC {method} to {addr} failed: {e}") from e
        s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        st = ClientStream(s, method, uid)
        try:
            st.send(tensors, metadata)
            _, st.opened = st.recv(timeout)
        except Exception:
            s.close()
            raise
        return st

    def __init__(self):
        self._socks: Dict[Tuple[str, int], socket.socket] = {}
        self._lock = threading.Lock()

    def _sock(self, addr: Tuple[str, int], timeout: float) -> socket.socket:
        s = self._socks.get(addr)
        if s is None:
            s = socket.create_connection(addr, timeout=timeout)
            s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            self._socks[addr] = s
        s.settimeout(timeout)
        return s

    def _drop(self, addr: Tuple[str, int]) -> None:
        s = self._socks.pop(addr, None)
        if s is not None:
            try:
                s.close()
            except OSError:
                pass

    def call(self, addr: Tuple[str, int], method: str, uid: str, tensors: List[torch.Tensor], metadata: dict,
             timeout: float = 30.0) -> Tuple[List[torch.Tensor], dict]:
        frame = pack_frame(method, uid, metadata, tensors)
        with self._lock:
            try:
                s = self._sock(addr, timeout)
                s.sendall(frame)
                raw = _recv_exact(s, _HDR.size)
                magic, hl, pl = _HDR.unpack(raw)
                if magic != MAGIC:
                    raise ConnectionError("bad frame magic")
                hb = _recv_exact(s, hl)
                payload = _recv_exact(s, pl) if pl else b""
            except socket.timeout as e:
                self._drop(addr)
                raise TimeoutError(f"RPC {method} to {addr} timed out after {timeout}s") from e
            except (OSError, ConnectionError) as e:
                self._drop(addr)
                raise ConnectionError(f"RPC {method} to {addr} failed: {e}") from e
        header, out = unpack_frame(hb, payload)
        if header.get("error"):
            raise RpcError(header["error"])
        return out, header.get("metadata") or {}

    def close(self) -> None:
        with self._lock:
            for a in list(self._socks):
                self._drop(a)


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    buf = bytearray(n)
    view = memoryview(buf)
    got = 0
    while got < n:
        r = sock.recv_into(view[got:], n - got)
        if r == 0:
            raise ConnectionError("peer closed the connection")
        got += r
    return bytes(buf)

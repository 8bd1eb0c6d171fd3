"""Framed-TCP RPC carrying tensors: the cross-host data plane (fallback to NCCL in a box).

The reference talks hivemind ``P2P.call_protobuf_handler`` / ``iterate_protobuf_handler``
with protobuf ``ExpertRequest{uid, tensors, metadata(msgpack)}`` messages over libp2p
(src/rpc_transport.py:524-562; src/rpc_handler.py:405-464).  Neither hivemind nor its Go
daemon exists here, so the same contract rides on plain TCP:

    frame  := magic u32 | header_len u32 | payload_len u64 | header (msgpack) | payload (raw)
    header := {method, uid, metadata{...}, tensors[{dtype, shape, nbytes}], error?}

Tensor bytes travel in their own dtype (bf16 stays 2 bytes; hivemind up-casts it to fp32).
Handlers are registered under hivemind-style names (``"StageConnectionHandler.rpc_forward"``).
The server gives every connection its own thread (blocking sockets); handlers run on that
thread and hand the GPU work to the stage's task pool, so it never blocks heartbeats.
Large payloads need no separate streaming RPC on TCP; ``rpc_forward_stream`` is the same
handler under a second name for wire-level parity.

Streams: a method registered with ``add_stream_handler`` owns its connection for as long as it
runs - the handler gets a ``stream`` object with ``recv(timeout)`` / ``send(tensors, metadata)``
and exchanges any number of frames with the client before it returns (its return value is the
closing frame).  This carries the per-session ``rpc_inference`` protocol of upstream Petals
(petals/server/handler.py:132-195); the client side is ``RpcClient.open_stream``.
"""
from __future__ import annotations

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
    """Thread-per-connection server over blocking sockets.

    A connection's thread reads a frame, runs the handler itself and writes the answer: no event loop and no executor
    hop between the socket and the handler, so a request costs two thread wake-ups less than the first (asyncio)
    version did - with many sessions stepping against one stage that was most of the server's time per request.
    Handlers mostly WAIT (for the stage's task pool, for the next hop); the compute itself is serialised - and
    batched - by the stage's own task pool thread."""

    def __init__(self, host: str = "127.0.0.1", port: int = 0):
        self.host, self.port = host, port
        self.handlers: Dict[str, Handler] = {}
        self.stream_handlers: Dict[str, Callable] = {}
        self._listener: Optional[socket.socket] = None
        self._thread: Optional[threading.Thread] = None
        self._conns: set = set()
        self._conns_lock = threading.Lock()
        self._stopping = False
        self.drop_next = 0          # fault injection: close the connection instead of answering
        self.requests_served = 0

    def add_handler(self, name: str, fn: Handler) -> None:
        self.handlers[name] = fn

    def add_stream_handler(self, name: str, fn: Callable) -> None:
        """``fn(uid, tensors, metadata, stream) -> (tensors, metadata)``; see the module docstring."""
        self.stream_handlers[name] = fn

    def start(self) -> "RpcServer":
        bind = "0.0.0.0" if self.host in ("0.0.0.0", "") else self.host
        ls = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        ls.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        ls.bind((bind, self.port))
        ls.listen(256)
        self.port = ls.getsockname()[1]
        self._listener = ls
        self._thread = threading.Thread(target=self._accept_loop, name="rpc-server", daemon=True)
        self._thread.start()
        return self

    def _accept_loop(self) -> None:
        ls = self._listener
        while not self._stopping:
            try:
                conn, _ = ls.accept()
            except OSError:
                return
            conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            with self._conns_lock:
                if self._stopping:
                    conn.close()
                    return
                self._conns.add(conn)
            threading.Thread(target=self._serve_conn, args=(conn,), name="rpc-conn", daemon=True).start()

    def _serve_conn(self, conn: socket.socket) -> None:
        try:
            while not self._stopping:
                raw = _recv_exact(conn, _HDR.size)
                magic, hl, pl = _HDR.unpack(raw)
                if magic != MAGIC or hl + pl > MAX_FRAME:
                    break
                hb = _recv_exact(conn, hl)
                payload = _recv_exact(conn, pl) if pl else b""
                if self.drop_next > 0:
                    self.drop_next -= 1
                    break
                header, tensors = unpack_frame(hb, payload)
                method, uid = header.get("method", ""), header.get("uid", "")
                sfn = self.stream_handlers.get(method)
                stream = None
                try:
                    if sfn is not None:
                        stream = _ServerStream(conn, method, uid)
                        res_t, res_md = sfn(uid, tensors, header.get("metadata") or {}, stream)
                    else:
                        fn = self.handlers.get(method)
                        if fn is None:
                            raise LookupError(f"no handler {method!r}")
                        res_t, res_md = fn(uid, tensors, header.get("metadata") or {})
                    out = pack_frame(method, uid, res_md, res_t)
                except LookupError as e:
                    out = pack_frame(method, uid, {}, [], error=str(e.args[0]) if e.args else "no handler")
                except Exception as e:  # noqa: BLE001 - forwarded to the caller
                    out = pack_frame(method, uid, {}, [], error=f"{type(e).__name__}: {e}")
                self.requests_served += 1
                if self._stopping:
                    break
                conn.settimeout(None)
                conn.sendall(out)
                if stream is not None and stream.broken:
                    break
        except (ConnectionError, OSError):
            pass
        finally:
            with self._conns_lock:
                self._conns.discard(conn)
            try:
                conn.close()
            except OSError:
                pass

    def stop(self) -> None:
        if self._listener is None:
            return
        self._stopping = True
        ls, self._listener = self._listener, None
        try:
            ls.shutdown(socket.SHUT_RDWR)       # wakes the accept() up
        except OSError:
            pass
        ls.close()
        with self._conns_lock:
            conns = list(self._conns)
        for c in conns:                         # wakes the connection threads up; in-flight answers are not sent
            try:
                c.shutdown(socket.SHUT_RDWR)
            except OSError:
                pass
        if self._thread is not None:
            self._thread.join(timeout=5)

    @property
    def address(self) -> Tuple[str, int]:
        return ("127.0.0.1" if self.host in ("0.0.0.0", "") else self.host), self.port


class _ServerStream:
    """A stream handler's view of its connection: blocking frame I/O on the connection's own thread."""

    def __init__(self, conn: socket.socket, method: str, uid: str):
        self._conn, self.method, self.uid = conn, method, uid
        self.broken = False

    def recv(self, timeout: Optional[float] = None) -> Tuple[List[torch.Tensor], dict]:
        try:
            self._conn.settimeout(timeout)
            raw = _recv_exact(self._conn, _HDR.size)
            magic, hl, pl = _HDR.unpack(raw)
            if magic != MAGIC or hl + pl > MAX_FRAME:
                raise ConnectionError("bad frame")
            hb = _recv_exact(self._conn, hl)
            payload = _recv_exact(self._conn, pl) if pl else b""
        except socket.timeout:
            self.broken = True          # a frame may be half read: the connection cannot be reused
            raise TimeoutError("no message on the stream") from None
        except (ConnectionError, OSError) as e:
            self.broken = True
            raise ConnectionError(f"stream closed: {e}") from None
        header, tensors = unpack_frame(hb, payload)
        return tensors, header.get("metadata") or {}

    def send(self, tensors: List[torch.Tensor], metadata: dict) -> None:
        frame = pack_frame(self.method, self.uid, metadata, tensors)
        try:
            self._conn.settimeout(30)
            self._conn.sendall(frame)
        except (socket.timeout, OSError) as e:
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

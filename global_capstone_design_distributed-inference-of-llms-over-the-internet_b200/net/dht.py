"""Registry with hivemind-DHT semantics (TTL, sub-keys, ``latest``) over localhost TCP.

The reference's control plane is a hivemind Kademlia DHT (src/main.py:304-312, 442-447;
src/dht_utils.py:71,122-127,170-174).  hivemind (and its Go ``p2pd`` daemon) cannot be
installed here, and inside one box a Kademlia swarm buys nothing, so this module
re-provides the part of the API the framework consumes:

    dht = DHT(start=True, initial_peers=[...], host_maddrs=["/ip4/0.0.0.0/tcp/8000"])
    dht.store(key, value, expiration_time, subkey=None) -> bool
    dht.get(key, latest=True) -> ValueWithExpiration | None      (.value, .expiration_time)
    dht.get_visible_maddrs(), dht.peer_id, dht.shutdown(), get_dht_time()

Records with sub-keys merge into a dict ``subkey -> ValueWithExpiration`` whose entries
expire individually (what replica announcement relies on).  Every non-client node serves its
table over TCP.  Nodes learn of each other transitively: each request carries the sender's
listen address and each answer the answering node's peer list, so a node bootstrapped from
one peer soon knows every node that peer knows (and they know it).  Stores are replicated to
every known peer and gets are merged over all of them (newest expiration wins); a peer that
stops answering is skipped for a few seconds instead of being dialled on every call.  The
registry therefore outlives its bootstrap node: servers keep re-announcing to each other and a
client that has talked to the swarm once keeps finding them (hivemind's Kademlia gives the
reference the same property, src/main.py:304-312).  Swarms here are a handful of nodes, so
"k closest" replication is simply "all known peers".
Multiaddrs look like ``/ip4/127.0.0.1/tcp/8000/p2p/<peer_id>``.
"""
from __future__ import annotations

import os
import socket
import socketserver
import threading
import time
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple

import msgpack

_B58 = "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz"


def get_dht_time() -> float:
    return time.time()


class PeerID:
    """Opaque, hashable, base58-printable peer identity."""

    def __init__(self, raw: bytes):
        self._raw = bytes(raw)

    @classmethod
    def generate(cls) -> "PeerID":
        return cls(os.urandom(20))

    def to_base58(self) -> str:
        n = int.from_bytes(b"\x01" + self._raw, "big")
        out = ""
        while n:
            n, r = divmod(n, 58)
            out = _B58[r] + out
        return "Qm" + out

    @classmethod
    def from_base58(cls, s: str) -> "PeerID":
        if s.startswith("Qm"):
            s = s[2:]
        n = 0
        for ch in s:
            n = n * 58 + _B58.index(ch)
        raw = n.to_bytes((n.bit_length() + 7) // 8, "big")
        return cls(raw[1:])

    def __str__(self) -> str:
        return self.to_base58()

    __repr__ = __str__

    def __hash__(self) -> int:
        return hash(self._raw)

    def __eq__(self, other) -> bool:
        return isinstance(other, PeerID) and other._raw == self._raw


@dataclass
class ValueWithExpiration:
    value: Any
    expiration_time: float


def parse_maddr(maddr: str) -> Tuple[str, int, Optional[str]]:
    """'/ip4/1.2.3.4/tcp/8000[/p2p/<id>]' -> (host, port, peer_id)."""
    parts = [p for p in maddr.strip().split("/") if p]
    host, port, pid = None, None, None
    for i in range(0, len(parts) - 1, 2):
        k, v = parts[i], parts[i + 1]
        if k in ("ip4", "ip6", "dns", "dns4"):
            host = v
        elif k == "tcp":
            port = int(v)
        elif k == "p2p":
            pid = v
    if host is None or port is None:
        raise ValueError(f"cannot parse multiaddr {maddr!r}")
    return host, port, pid


class _Table:
    def __init__(self):
        self.lock = threading.Lock()
        self.plain: Dict[str, ValueWithExpiration] = {}
        self.sub: Dict[str, Dict[str, ValueWithExpiration]] = {}

    def store(self, key: str, value, exp: float, subkey: Optional[str]) -> bool:
        now = get_dht_time()
        if exp <= now:
            return False
        with self.lock:
            if subkey is None:
                old = self.plain.get(key)
                if old is not None and old.expiration_time > exp:
                    return False
                self.plain[key] = ValueWithExpiration(value, exp)
            else:
                d = self.sub.setdefault(key, {})
                old = d.get(subkey)
                if old is not None and old.expiration_time > exp:
                    return False
                d[subkey] = ValueWithExpiration(value, exp)
        return True

    def get(self, key: str):
        """-> ('plain', value, exp) | ('sub', {subkey: (value, exp)}) | None, expired entries dropped."""
        now = get_dht_time()
        with self.lock:
            d = self.sub.get(key)
            if d:
                for k in [k for k, v in d.items() if v.expiration_time <= now]:
                    del d[k]
                if d:
                    return ("sub", {k: (v.value, v.expiration_time) for k, v in d.items()})
                del self.sub[key]
            v = self.plain.get(key)
            if v is not None:
                if v.expiration_time <= now:
                    del self.plain[key]
                    return None
                return ("plain", v.value, v.expiration_time)
        return None


def _send_msg(sock: socket.socket, obj) -> None:
    data = msgpack.packb(obj, use_bin_type=True)
    sock.sendall(len(data).to_bytes(4, "big") + data)


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(n - len(buf))
        if not chunk:
            raise ConnectionError("peer closed the connection")
        buf += chunk
    return bytes(buf)


def _recv_msg(sock: socket.socket):
    n = int.from_bytes(_recv_exact(sock, 4), "big")
    if n > 64 << 20:
        raise ValueError("oversized DHT message")
    return msgpack.unpackb(_recv_exact(sock, n), raw=False, strict_map_key=False)


class _Handler(socketserver.BaseRequestHandler):
    def handle(self):
        table: _Table = self.server.table  # type: ignore[attr-defined]
        node = getattr(self.server, "node", None)
        try:
            while True:
                req = _recv_msg(self.request)
                op = req.get("op")
                if node is not None and req.get("from"):
                    node._learn([req["from"]])          # the caller is a serving node: remember it
                peers = node.known_peers() if node is not None else []
                if op == "store":
                    ok = table.store(req["key"], req["value"], req["exp"], req.get("subkey"))
                    _send_msg(self.request, {"ok": ok, "peers": peers})
                elif op == "get":
                    _send_msg(self.request, {"r": table.get(req["key"]), "peers": peers})
                elif op == "ping":
                    _send_msg(self.request, {"peer_id": self.server.peer_id, "peers": peers})  # type: ignore[attr-defined]
                else:
                    _send_msg(self.request, {"error": f"unknown op {op!r}"})
        except (ConnectionError, OSError, ValueError):
            return


class _Server(socketserver.ThreadingTCPServer):
    allow_reuse_address = True
    daemon_threads = True


class DHT:
    def __init__(self, start: bool = True, initial_peers: Optional[List[str]] = None,
                 host_maddrs: Optional[List[str]] = None, announce_maddrs: Optional[List[str]] = None,
                 client_mode: bool = False, **_ignored):
        self.peer_id = PeerID.generate()
        self.initial_peers = [p for p in (initial_peers or []) if p]
        self._table = _Table()
        self._server: Optional[_Server] = None
        self._thread: Optional[threading.Thread] = None
        self._conns: Dict[Tuple[str, int], socket.socket] = {}
        self._conn_lock = threading.Lock()
        self._host, self._port = "127.0.0.1", 0
        self._announce = announce_maddrs or []
        if host_maddrs:
            h, p, _ = parse_maddr(host_maddrs[0])
            self._host, self._port = h, p
        self._peers: List[Tuple[str, int]] = []
        self._peer_lock = threading.Lock()
        self._dead_until: Dict[Tuple[str, int], float] = {}
        self._fails: Dict[Tuple[str, int], int] = {}
        for m in self.initial_peers:
            h, p, _ = parse_maddr(m)
            self._peers.append((h, p))
        self.client_mode = client_mode
        if start:
            self.run_in_background()
            self.refresh_peers()

    # ---- lifecycle
    def run_in_background(self) -> None:
        if self.client_mode or self._server is not None:
            return
        bind_host = "0.0.0.0" if self._host in ("0.0.0.0", "") else self._host
        self._server = _Server((bind_host, self._port), _Handler)
        self._server.table = self._table          # type: ignore[attr-defined]
        self._server.peer_id = str(self.peer_id)  # type: ignore[attr-defined]
        self._server.node = self                  # type: ignore[attr-defined]
        self._port = self._server.server_address[1]
        self._thread = threading.Thread(target=self._server.serve_forever, name="dht-server", daemon=True)
        self._thread.start()

    def shutdown(self) -> None:
        with self._peer_lock:                      # idempotent, also when two threads stop the node at once
            server, self._server = self._server, None
        if server is not None:
            server.shutdown()
            server.server_close()
        with self._conn_lock:
            for s in self._conns.values():
                try:
                    s.close()
                except OSError:
                    pass
            self._conns.clear()

    def get_visible_maddrs(self) -> List[str]:
        if self._announce:
            return [f"{m}/p2p/{self.peer_id}" if "/p2p/" not in m else m for m in self._announce]
        host = "127.0.0.1" if self._host in ("0.0.0.0", "") else self._host
        return [f"/ip4/{host}/tcp/{self._port}/p2p/{self.peer_id}"]

    # ---- peer bookkeeping
    DEAD_BACKOFF = 5.0
    FORGET_AFTER = 8

    def _listen_addr(self) -> Optional[List]:
        """Address other nodes can dial this node's table at (None for client-mode nodes)."""
        if self._server is None:
            return None
        if self._announce:
            h, p, _ = parse_maddr(self._announce[0])
            return [h, p]
        host = "127.0.0.1" if self._host in ("0.0.0.0", "") else self._host
        return [host, self._port]

    def known_peers(self) -> List[List]:
        with self._peer_lock:
            out = [[h, p] for h, p in self._peers]
        me = self._listen_addr()
        if me is not None:
            out.append(me)
        return out

    def _learn(self, peers) -> None:
        me = self._listen_addr()
        with self._peer_lock:
            for hp in peers or []:
                try:
                    a = (str(hp[0]), int(hp[1]))
                except (TypeError, ValueError, IndexError):
                    continue
                if me is not None and a == (me[0], me[1]):
                    continue
                if a not in self._peers:
                    self._peers.append(a)

    def _live_peers(self) -> List[Tuple[str, int]]:
        now = time.monotonic()
        with self._peer_lock:
            return [a for a in self._peers if self._dead_until.get(a, 0.0) <= now]

    def refresh_peers(self, rounds: int = 2) -> int:
        """Introduce this node to every peer it knows of and collect theirs (transitive closure in `rounds`)."""
        for _ in range(rounds):
            before = len(self._peers)
            for a in self._live_peers():
                self._call(a, {"op": "ping"}, timeout=1.5)
            if len(self._peers) == before:
                break
        return len(self._peers)

    # ---- remote plumbing
    def _call(self, addr: Tuple[str, int], req: dict, timeout: float = 3.0):
        me = self._listen_addr()
        if me is not None:
            req = dict(req)
            req["from"] = me
        r = self._call_raw(addr, req, timeout)
        if r is None:
            n = self._fails.get(addr, 0) + 1
            self._fails[addr] = n
            self._dead_until[addr] = time.monotonic() + self.DEAD_BACKOFF
            if n >= self.FORGET_AFTER:
                # gone for good (or re-placed under another port): stop gossiping it
                with self._peer_lock:
                    if addr in self._peers and len(self._peers) > 1:
                        self._peers.remove(addr)
                self._fails.pop(addr, None)
        else:
            self._fails.pop(addr, None)
            self._dead_until.pop(addr, None)
            self._learn(r.get("peers"))
        return r

    def _call_raw(self, addr: Tuple[str, int], req: dict, timeout: float = 3.0):
        with self._conn_lock:
            s = self._conns.get(addr)
            try:
                if s is None:
                    s = socket.create_connection(addr, timeout=timeout)
                    s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                    self._conns[addr] = s
                s.settimeout(timeout)
                _send_msg(s, req)
                return _recv_msg(s)
            except (OSError, ConnectionError, ValueError):
                if s is not None:
                    try:
                        s.close()
                    except OSError:
                        pass
                self._conns.pop(addr, None)
                return None

    # ---- API
    def store(self, key: str, value, expiration_time: float, subkey: Optional[str] = None,
              return_future: bool = False, **_ignored) -> bool:
        ok = self._table.store(key, value, float(expiration_time), subkey)
        for addr in self._live_peers():
            r = self._call(addr, {"op": "store", "key": key, "value": value, "exp": float(expiration_time),
                                  "subkey": subkey})
            ok = ok or bool(r and r.get("ok"))
        return ok

    def get(self, key: str, latest: bool = True, **_ignored) -> Optional[ValueWithExpiration]:
        results = [self._table.get(key)]
        for addr in self._live_peers():
            r = self._call(addr, {"op": "get", "key": key})
            if r is not None:
                results.append(r.get("r"))
        merged_sub: Dict[str, ValueWithExpiration] = {}
        best_plain: Optional[ValueWithExpiration] = None
        for r in results:
            if not r:
                continue
            r = list(r)
            if r[0] == "sub":
                for sk, (v, exp) in r[1].items():
                    cur = merged_sub.get(sk)
                    if cur is None or exp > cur.expiration_time:
                        merged_sub[sk] = ValueWithExpiration(v, exp)
            elif r[0] == "plain":
                if best_plain is None or r[2] > best_plain.expiration_time:
                    best_plain = ValueWithExpiration(r[1], r[2])
        if merged_sub:
            return ValueWithExpiration(merged_sub, max(v.expiration_time for v in merged_sub.values()))
        return best_plain

    def ping_peers(self) -> int:
        return sum(1 for a in list(self._peers) if self._call(a, {"op": "ping"}) is not None)

"""Reachability probing (upstream petals/server/reachability.py:22-164).

Upstream asks a public API and other peers (``rpc_check``) whether a joining server can be dialled
from the Internet and falls back to relays.  Offline and inside one box the only meaningful part
is the direct probe: can the announced address be connected to, and does the peer behind it answer
a ping?  ``check_direct_reachability`` does that; servers expose it to others as ``rpc_check``.
"""
from __future__ import annotations

import socket
import time
from typing import Optional, Tuple

from .dht import parse_maddr


def check_direct_reachability(maddr: str, timeout: float = 2.0) -> Tuple[bool, Optional[float]]:
    """(reachable, connect RTT in seconds) for a ``/ip4/../tcp/..`` multiaddr."""
    host, port, _ = parse_maddr(maddr)
    t0 = time.perf_counter()
    try:
        with socket.create_connection((host, port), timeout=timeout):
            return True, time.perf_counter() - t0
    except OSError:
        return False, None


def validate_reachability(maddrs, timeout: float = 2.0) -> None:
    """Raises if none of the announced addresses accepts connections (a server that nobody can
    reach only pollutes the registry)."""
    if not any(check_direct_reachability(m, timeout)[0] for m in maddrs):
        raise RuntimeError(f"none of the announced addresses is reachable: {list(maddrs)}")

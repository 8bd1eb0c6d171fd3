"""DHT key schema and registry helpers (reference: src/dht_utils.py:20-281).

Record families, all refreshed by server heartbeats and expiring by TTL:

    mini_petals:stage{N}                 subkey=peer_id  -> {peer_id, timestamp, stage, p2p_maddrs, ...}
    petals:server:{model}:{peer_id}                      -> {start_block, end_block, throughput, state, ...}
    petals:module:{model}:block_{i}      subkey=peer_id  -> {block_idx, start_block, end_block, throughput,
                                                             state, p2p_maddrs, final_stage}

``{model}`` is the raw ``--model`` string and ``peer_id`` is the RPC endpoint's identity.
"""
from __future__ import annotations

import logging
from typing import Any, Dict, List, Optional, Sequence

from petals_b200.net.dht import DHT, get_dht_time

from .load_balancing import RemoteModuleInfo, ServerInfo, ServerState

logger = logging.getLogger(__name__)

MODULE_KEY_PREFIX = "petals:module:"
SERVER_KEY_PREFIX = "petals:server:"
STAGE_KEY_PREFIX = "mini_petals:stage"
DEFAULT_TTL = 90.0


def get_module_key(block_idx: int, model_name: str = "default") -> str:
    return f"{MODULE_KEY_PREFIX}{model_name}:block_{block_idx}"


def get_server_key(peer_id, model_name: str = "default") -> str:
    return f"{SERVER_KEY_PREFIX}{model_name}:{peer_id}"


def get_stage_key(stage: int) -> str:
    return f"{STAGE_KEY_PREFIX}{stage}"


def _exp(expiration_time: Optional[float]) -> float:
    return get_dht_time() + DEFAULT_TTL if expiration_time is None else expiration_time


def register_stage_on_dht(dht: DHT, stage: int, peer_id, p2p_maddrs: Sequence[str],
                          expiration_time: Optional[float] = None, **extra) -> bool:
    """``mini_petals:stage{N}`` record for fixed-split routing (src/main.py:517-527)."""
    rec = {"peer_id": str(peer_id), "timestamp": get_dht_time(), "stage": int(stage),
           "p2p_maddrs": list(p2p_maddrs), **extra}
    return dht.store(get_stage_key(stage), rec, _exp(expiration_time), subkey=str(peer_id))


def register_server_on_dht(dht: DHT, peer_id, start_block: int, end_block: int, throughput: float,
                           model_name: str = "default", server_address: Optional[str] = None,
                           p2p_maddrs: Optional[List[str]] = None, final_stage: bool = False,
                           state: ServerState = ServerState.ONLINE, expiration_time: Optional[float] = None) -> bool:
    rec = {"peer_id": str(peer_id), "timestamp": get_dht_time(), "start_block": int(start_block),
           "end_block": int(end_block), "throughput": float(throughput), "state": state.value,
           "server_address": server_address, "p2p_maddrs": list(p2p_maddrs or []), "final_stage": bool(final_stage)}
    try:
        ok = dht.store(get_server_key(peer_id, model_name), rec, _exp(expiration_time))
        logger.info("Registered server %s on DHT: blocks [%d:%d], throughput=%.2f rps, final_stage=%s",
                    peer_id, start_block, end_block, throughput, final_stage)
        return ok
    except Exception as e:  # noqa: BLE001
        logger.error("Failed to register server on DHT: %s", e)
        return False


def register_blocks_on_dht(dht: DHT, peer_id, block_indices: Sequence[int], model_name: str = "default",
                           p2p_maddrs: Optional[List[str]] = None, start_block: Optional[int] = None,
                           end_block: Optional[int] = None, throughput: Optional[float] = None,
                           final_stage: bool = False, state: ServerState = ServerState.ONLINE,
                           expiration_time: Optional[float] = None, **extra) -> bool:
    """One sub-keyed record per served block, so replicas coexist under the same block key."""
    exp = _exp(expiration_time)
    ok = True
    for b in block_indices:
        rec = {"peer_id": str(peer_id), "timestamp": get_dht_time(), "block_idx": int(b),
               "start_block": None if start_block is None else int(start_block),
               "end_block": None if end_block is None else int(end_block),
               "throughput": None if throughput is None else float(throughput), "state": state.value,
               "p2p_maddrs": list(p2p_maddrs or []), "final_stage": bool(final_stage), **extra}
        try:
            ok = dht.store(get_module_key(b, model_name), rec, exp, subkey=str(peer_id)) and ok
        except Exception as e:  # noqa: BLE001
            logger.error("Failed to register block %d: %s", b, e)
            ok = False
    return ok


MODELS_KEY = "_petals.models"


def declare_model(dht: DHT, model_name: str, peer_id, info: Dict[str, Any],
                  expiration_time: Optional[float] = None) -> bool:
    """The swarm-wide list of served models (upstream ``_petals.models``, petals/server/server.py:738-744): every
    server adds ``{model -> {peer -> info}}`` so that a monitor can discover what a swarm serves without knowing
    model names in advance.  ``info``: architecture facts a client needs before it asks for blocks."""
    rec = {"peer_id": str(peer_id), "model": model_name, "timestamp": get_dht_time(), **info}
    return dht.store(MODELS_KEY, rec, _exp(expiration_time), subkey=f"{model_name}|{peer_id}")


def get_models(dht: DHT) -> Dict[str, List[Dict[str, Any]]]:
    """{model name: [live server records]} from the ``_petals.models`` key."""
    out: Dict[str, List[Dict[str, Any]]] = {}
    rec = dht.get(MODELS_KEY, latest=True)
    if rec is None or not isinstance(rec.value, dict):
        return out
    for x in rec.value.values():
        v = getattr(x, "value", None)
        if isinstance(v, dict) and v.get("model"):
            out.setdefault(str(v["model"]), []).append(v)
    return out


def _live_entries(record) -> Dict[str, Dict[str, Any]]:
    """{subkey: value dict} of a sub-keyed record (tolerates plain records)."""
    if record is None:
        return {}
    v = record.value
    if isinstance(v, dict) and v and all(hasattr(x, "value") for x in v.values()):
        return {k: x.value for k, x in v.items() if isinstance(x.value, dict)}
    if isinstance(v, dict) and "peer_id" in v:
        return {str(v["peer_id"]): v}
    return {}


def get_block_entries(dht: DHT, block_idx: int, model_name: str = "default") -> List[Dict[str, Any]]:
    return list(_live_entries(dht.get(get_module_key(block_idx, model_name), latest=True)).values())


def get_stage_entries(dht: DHT, stage: int) -> List[Dict[str, Any]]:
    return list(_live_entries(dht.get(get_stage_key(stage), latest=True)).values())


def get_remote_module_infos(dht: DHT, model_name: str = "default", total_blocks: int = 32) -> List[RemoteModuleInfo]:
    """Scan ``block_0 .. block_{N-1}``; one RemoteModuleInfo per (block, serving peer)."""
    infos: List[RemoteModuleInfo] = []
    for b in range(total_blocks):
        entries = get_block_entries(dht, b, model_name)
        if not entries:
            infos.append(RemoteModuleInfo(uid=f"block_{b}", server_info=None))
            continue
        for e in entries:
            try:
                state = ServerState(e.get("state", "online"))
            except ValueError:
                state = ServerState.ONLINE
            srv = ServerInfo(peer_id=e.get("peer_id"), state=state, throughput=float(e.get("throughput") or 0.0),
                             start_block=int(e["start_block"]) if e.get("start_block") is not None else b,
                             end_block=int(e["end_block"]) if e.get("end_block") is not None else b + 1,
                             server_address=(e.get("p2p_maddrs") or [None])[0])
            infos.append(RemoteModuleInfo(uid=f"block_{b}", server_info=srv))
    return infos


def update_server_throughput_on_dht(dht: DHT, peer_id, throughput: float, model_name: str = "default",
                                    expiration_time: Optional[float] = None) -> bool:
    rec = dht.get(get_server_key(peer_id, model_name), latest=True)
    if rec is None or not isinstance(rec.value, dict):
        return False
    new = dict(rec.value)
    new["throughput"] = float(throughput)
    new["timestamp"] = get_dht_time()
    return dht.store(get_server_key(peer_id, model_name), new, _exp(expiration_time))


def announce_offline(dht: DHT, peer_id, block_indices: Sequence[int], model_name: str = "default", **kw) -> bool:
    """Graceful shutdown announcement (upstream server.py:650).

    A DHT record only yields to one with a LATER expiration, so the OFFLINE marker carries a full
    TTL; readers skip OFFLINE entries, and the marker lapses on its own afterwards."""
    return register_blocks_on_dht(dht, peer_id, block_indices, model_name, state=ServerState.OFFLINE,
                                  expiration_time=get_dht_time() + DEFAULT_TTL, **kw)

#!/usr/bin/env python
"""What a swarm serves, from nothing but a registry address (the role of upstream's health monitor).

    python scripts/swarm_status.py /ip4/10.0.0.1/tcp/8000 [--json]

Reads the ``_petals.models`` key (every server declares its model there), then per model the stage / block records,
asks every server for ``rpc_info`` and prints: who holds which blocks, throughput, sessions, cache tokens left, queue
length, mean batch, whether the model is covered end to end.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def collect(peers: str) -> dict:
    from petals_b200.net.dht import DHT, parse_maddr
    from petals_b200.net.rpc import RpcClient
    from src.dht_utils import get_models

    dht = DHT(start=True, initial_peers=[peers], client_mode=True)
    client = RpcClient()
    out = {}
    try:
        for model, servers in sorted(get_models(dht).items()):
            rows, n_layers = [], None
            for s in sorted(servers, key=lambda r: (r.get("blocks") or [0])[0] or 0):
                n_layers = n_layers or s.get("n_layers")
                row = {"peer": str(s.get("peer_id"))[:12], "blocks": s.get("blocks"), "stage": s.get("stage"),
                       "final": s.get("final_stage"), "throughput_rps": s.get("throughput"), "dtype": s.get("dtype")}
                try:
                    host, port, _ = parse_maddr((s.get("p2p_maddrs") or [""])[0])
                    _, info = client.call((host, port), "StageConnectionHandler.rpc_info", "status", [], {}, timeout=3.0)
                    row.update(alive=bool(info.get("healthy", True)), sessions=info.get("sessions"), cache_tokens_left=info.get("cache_tokens_left"),
                               queued=info.get("queued"), mean_batch=round(info.get("mean_batch", 0.0), 2),
                               requests=info.get("requests"), uptime_s=round(info.get("uptime_s", 0.0), 1))
                except Exception as e:  # noqa: BLE001
                    row.update(alive=False, error=str(e)[:80])
                rows.append(row)
            covered = set()
            for r in rows:
                if r.get("alive") and r.get("blocks"):
                    covered.update(range(int(r["blocks"][0]), int(r["blocks"][1])))
            first = min(covered) if covered else None
            out[model] = {"n_layers": n_layers, "servers": rows,
                          "served_blocks": [first, max(covered) + 1] if covered else None,
                          "gaps": sorted(set(range(first, n_layers)) - covered) if covered and n_layers else None}
    finally:
        client.close()
        dht.shutdown()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("peers", help="/ip4/<ip>/tcp/<dht_port> of any node of the swarm")
    ap.add_argument("--json", action="store_true")
    a = ap.parse_args()
    st = collect(a.peers)
    if a.json:
        print(json.dumps(st, indent=1))
        return 0
    if not st:
        print("no model is declared on this registry")
        return 1
    for model, m in st.items():
        gaps = m["gaps"]
        print(f"{model}: {m['n_layers']} blocks, served {m['served_blocks']}, "
              f"{'no gaps' if gaps == [] else 'MISSING blocks ' + str(gaps)}")
        for r in m["servers"]:
            state = (f"sessions {r.get('sessions')} queue {r.get('queued')} cache_left {r.get('cache_tokens_left')} "
                     f"mean_batch {r.get('mean_batch')} up {r.get('uptime_s')}s") if r["alive"] else f"DEAD ({r.get('error')})"
            print(f"  {r['peer']}  blocks {r['blocks']}  stage {r['stage']}  final {r['final']}  "
                  f"{r['throughput_rps'] or 0:.1f} rps  {r['dtype']}  {state}")
    return 0


if __name__ == "__main__":
    sys.exit(main())

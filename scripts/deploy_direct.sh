#!/usr/bin/env bash
# Start one stage server on this host in the background (reference: scripts/deploy_direct.sh:1-101).
# usage: deploy_direct.sh <stage> <model> <splits> [dht_initial_peers] [public_ip] [dtype]
# Ports default to 8000+2*(stage-1) (DHT) and 8001+2*(stage-1) (RPC) like the reference's 8002+2(stage-1) scheme.
# No venv / pip step: the image already carries torch; hivemind is not needed (TCP control plane).
set -euo pipefail
STAGE=${1:?stage}; MODEL=${2:?model}; SPLITS=${3:?splits}; PEERS=${4:-}; PUBLIC_IP=${5:-}; DTYPE=${6:-fp16}
cd "$(dirname "$0")/.."
python csrc/build.py >/dev/null
DHT_PORT=$((8000 + 2 * (STAGE - 1))); RPC_PORT=$((DHT_PORT + 1))
PIDFILE="stage${STAGE}.pid"
if [[ -f "$PIDFILE" ]] && kill -0 "$(cat "$PIDFILE")" 2>/dev/null; then
  echo "stopping previous stage ${STAGE} server (pid $(cat "$PIDFILE"))"; kill "$(cat "$PIDFILE")"; sleep 1
fi
nohup python -m src.main --model "$MODEL" --splits "$SPLITS" --stage "$STAGE" --dtype "$DTYPE" \
  --dht_port "$DHT_PORT" --rpc_port "$RPC_PORT" --dht_initial_peers "$PEERS" \
  ${PUBLIC_IP:+--public_ip "$PUBLIC_IP"} > "stage${STAGE}.log" 2>&1 &
echo $! > "$PIDFILE"
echo "stage ${STAGE} started (pid $!), log: stage${STAGE}.log"

#!/usr/bin/env bash
# Load-balancing demo on one host (reference: scripts/elice_test_load_balancing.sh:1-130:
# 4 LB servers with --num_blocks 8 --total_blocks 32 on Llama-3.1-8B + a client).
# usage: test_load_balancing.sh [model] [num_servers] [num_blocks] [client_blocks]
set -euo pipefail
MODEL=${1:-llama-3.1-8b}; N=${2:-4}; NB=${3:-8}; C=${4:-0}
cd "$(dirname "$0")/.."
PEERS=""; PIDS=()
trap 'kill "${PIDS[@]}" 2>/dev/null || true' EXIT
GPUS=$(python -c "import torch;print(torch.cuda.device_count())")
for i in $(seq 1 "$N"); do
  LOG="lb_server_${i}.log"
  LOCAL_RANK=$(( GPUS > 0 ? (i - 1) % GPUS : 0 )) python -m src.main --model "$MODEL" --splits "$C" --stage 1 \
    --use_load_balancing --num_blocks "$NB" --dtype bf16 --dht_port $((8100 + 2 * i)) --rpc_port $((8101 + 2 * i)) \
    --dht_initial_peers "$PEERS" > "$LOG" 2>&1 &
  PIDS+=($!)
  until grep -q "Load Balancing server ready" "$LOG"; do sleep 1; kill -0 "${PIDS[-1]}" || { cat "$LOG"; exit 1; }; done
  grep -E "Selected blocks|Load Balancing server ready" "$LOG"
  if [[ -z "$PEERS" ]]; then PEERS=$(grep -o "/ip4/[^']*" "$LOG" | head -1); fi
done
python -m src.main --model "$MODEL" --splits "$C" --stage 0 --use_load_balancing --dtype bf16 \
  --dht_initial_peers "$PEERS" --max_new_tokens 32 --temperature 0

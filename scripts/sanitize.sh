#!/usr/bin/env bash
# compute-sanitizer pass over the kernel tests (one tool per GPU session, see B200_PROFILING.md):
#   scripts/sanitize.sh memcheck|racecheck|synccheck [pytest args]
# Small shapes only: the sanitizer slows kernels by 10-100x.
set -euo pipefail
TOOL=${1:-memcheck}; shift || true
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
compute-sanitizer --tool "$TOOL" --error-exitcode 99 --launch-timeout 600 \
  python -m pytest tests/test_gemm_gpu.py tests/test_attention_gpu.py -x -q -k "${SANITIZE_K:-1-256-128 or swiglu_dual or fused_norm or attn_decode and 8-8-128 and 63 or prefill and 33-0}" "$@" \
  2>&1 | tee "gpurun_out/sanitizer_${TOOL}.log" | tail -15

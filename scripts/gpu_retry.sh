#!/usr/bin/env bash
# Runs a gpurun call, retrying while the pod answers "busy" (exit code 3, nothing charged).
#   scripts/gpu_retry.sh [gpurun options] -- '<command>'
for i in $(seq 1 40); do
    /usr/local/graft/bin/gpurun "$@"
    rc=$?
    if [ "$rc" != "3" ]; then exit $rc; fi
    sleep 45
done
exit 3

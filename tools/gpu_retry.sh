#!/bin/bash
# Retries one gpurun call until the pod has a free slot (exit 3 / "transient" answers are not charged).
#   tools/gpu_retry.sh <timeout-seconds> '<command>' [extra gpurun flags...]
# NOTE: gpurun snapshots /root/repo when the call is ACCEPTED: do not leave the tree half-edited while this waits.
cd "$(dirname "$0")/.."
t=$1; cmd=$2; shift 2
for i in $(seq 1 40); do
  out=$(gpurun --timeout "$t" "$@" -- "$cmd" 2>&1)
  echo "$out" | tail -12
  if echo "$out" | grep -q "status=transient"; then sleep 60; continue; fi
  exit 0
done
exit 1

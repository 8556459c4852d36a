#!/bin/bash
# usage: scripts/gpu_retry.sh OUTFILE TIMEOUT 'command' [--gpus N]   — retries gpurun while the pod answers busy / transient
out=$1; to=$2; cmd=$3; shift 3
mkdir -p "$(dirname "$out")"
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" --timeout "$to" -- "$cmd" > "$out" 2>&1
  rc=$?
  if [ $rc -eq 3 ] || grep -qi "transient" "$out"; then sleep 60; continue; fi
  break
done
echo "gpu_retry rc=$rc" >> "$out"

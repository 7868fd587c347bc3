#!/bin/bash
# usage: tools/dev/gpu_retry.sh <timeout> <script> <logfile> : retries while the pod is busy
for i in $(seq 1 40); do
  gpurun --timeout "$1" -- bash "$2" > "$3" 2>&1
  if ! grep -q "status=transient" "$3"; then exit 0; fi
  sleep 60
done

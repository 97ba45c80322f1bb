#!/bin/bash
# gpurun with retries while the pod answers "transient" (exit 3 / status=transient): usage gpurun_retry.sh <timeout> '<command>'
for i in $(seq 1 12); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$1" -- "$2" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient\|status=busy" || [ $rc -eq 3 ]; then echo "[retry $i] busy"; sleep 150; continue; fi
  echo "$out" | tail -${3:-60}; exit $rc
done
echo "gave up"; exit 3

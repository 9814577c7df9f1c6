#!/bin/bash
# retry a gpurun call while the pod answers "busy" (nothing is charged for those)
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1)
  echo "$out" | tail -12
  if echo "$out" | grep -q "status=transient"; then echo "[retry $i] busy, sleeping 120 s"; sleep 120; continue; fi
  break
done

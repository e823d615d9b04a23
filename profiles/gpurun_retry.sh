#!/bin/bash
# scratch helper: retry a gpurun call while the pod answers "transient" (no slot; nothing charged).
# usage: profiles/gpurun_retry.sh <timeout-seconds> '<command>' [--gpus N]
T=$1; CMD=$2; shift 2
for i in $(seq 1 40); do
  OUT=$(/usr/local/graft/bin/gpurun "$@" --timeout "$T" -- "$CMD" 2>&1)
  if echo "$OUT" | grep -q "status=transient"; then sleep 120; continue; fi
  echo "$OUT" | tail -60
  exit 0
done
echo "gpurun_retry: gave up after 40 tries"

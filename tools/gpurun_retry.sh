#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <gpurun args...>   -- repeats a gpurun call while the pod answers "transient" (nothing charged)
log=$1; shift
for i in $(seq 1 20); do
  gpurun "$@" > "$log" 2>&1
  grep -q "status=transient" "$log" || break
  sleep 60
done
tail -3 "$log"

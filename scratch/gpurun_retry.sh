#!/bin/bash
# usage: scratch/gpurun_retry.sh OUTFILE [gpurun args...]   -- retries while the pod answers "busy" (exit 3)
out=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$out" 2>&1; rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc after $i attempt(s)" >> "$out"; exit $rc; fi
  sleep 120
done
echo "gave up" >> "$out"; exit 3

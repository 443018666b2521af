#!/bin/bash
# usage: gpurun_retry.sh <out-file> <gpurun args...>   — retries while the pod answers busy (exit 3)
out=$1; shift
for i in $(seq 1 40); do
  gpurun "$@" > "$out" 2>&1; rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc" >> "$out"; exit $rc; fi
  sleep 120
done
echo "gpurun: gave up" >> "$out"

#!/bin/bash
# developer aid: run selected tests each in its own process with a short timeout, report which ones hang
mkdir -p gpurun_out
for t in $(python -m pytest tests/test_gpu_parity.py -m gpu --collect-only -q -k "$1" 2>/dev/null | grep "::"); do
  s=$(date +%s)
  timeout 40 python -m pytest "$t" -m gpu -q -x -p no:cacheprovider --tb=line > gpurun_out/hunt_one.log 2>&1
  rc=$?
  e=$(date +%s)
  echo "$rc $((e - s))s $t"
done

#!/bin/bash
# developer aid: repeat the flaky train tests in both launch modes
for mode in cluster grid; do
  for i in 1 2 3; do
    BPE_B200_TRAIN_MODE=$mode timeout 60 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider --tb=line -k "ts:1 or freshly or corpus.en-4000 or owt:3" 2>&1 | tail -1
    echo "mode=$mode run=$i rc=$?"
  done
done

#!/bin/bash
# Developer tool: ncu --set full of the one-pass emit kernel (warm encode) on a 256 MB corpus; run under gpurun.  $1 = report name
python tools/run_encode_once.py 256000000 3 > gpurun_out/plain.log 2>&1 && tail -2 gpurun_out/plain.log && \
ncu --set full --clock-control none --import-source on -k regex:fused_emit -s 1 -c 1 -o gpurun_out/$1 python tools/run_encode_once.py 256000000 3 > gpurun_out/ncu.log 2>&1; tail -1 gpurun_out/ncu.log

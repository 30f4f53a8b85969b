#!/bin/bash
# Developer tool: the round's evidence in one gpurun call.  $1 = tag (file prefix under gpurun_out/).
#   GPU test suite, smoke, the default bench line, the launch list of the bench command, one `ncu --set full` capture
#   of the byte kernels of the product path (count, then encode) at the bench size.
t=${1:-r02}
python -m pytest tests -m gpu -x -q > gpurun_out/${t}_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${t}_gpu_tests.log
python __graft_entry__.py smoke > gpurun_out/${t}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/${t}_bench.json 2> gpurun_out/${t}_bench.log; echo "bench rc=$?"; cat gpurun_out/${t}_bench.json
python bench.py --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${t}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/${t}_launches.log 2>&1; echo "launch list rc=$?"
python tools/run_encode_once.py 2100000000 2 0 > gpurun_out/${t}_once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'fused_count|scan_kernel|emit_fused' -o gpurun_out/${t}_full \
    python tools/run_encode_once.py 2100000000 2 0 > gpurun_out/${t}_full.log 2>&1; echo "ncu full rc=$?"; tail -2 gpurun_out/${t}_full.log

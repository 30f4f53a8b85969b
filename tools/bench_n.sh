#!/bin/bash
# Developer tool: the bench line at N GPUs of one box, as the driver launches it.  $1 = N, $2 = tag
N=$1; t=${2:-r02}
BPE_B200_TRAIN_WATCHDOG=20 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/${t}_bench_n$N.json 2> gpurun_out/${t}_bench_n$N.log
echo "rc=$?"; cat gpurun_out/${t}_bench_n$N.json | cut -c1-1500; grep -E "^\[bench\]" gpurun_out/${t}_bench_n$N.log | tail -12

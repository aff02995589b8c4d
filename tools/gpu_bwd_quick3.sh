#!/bin/bash
# BASELINE config 2 shape: one 20 000 x 2048 bag per step
mkdir -p gpurun_out
timeout 30 python tools/bench_train.py --bags 1 --patches 20000 --steps 2 --warmup 1 > gpurun_out/train_bench_20k.json 2> gpurun_out/train_bench_20k.err
echo "exit $?" >> gpurun_out/train_bench_20k.err
cat gpurun_out/train_bench_20k.json; tail -3 gpurun_out/train_bench_20k.err

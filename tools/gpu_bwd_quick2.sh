#!/bin/bash
# second short on-device pass: the whole backward GPU suite, then one timed training step line
mkdir -p gpurun_out
timeout 50 python -m pytest tests/test_zz_backward_gpu.py -x -q > gpurun_out/bwd_suite.log 2>&1
echo "exit $?" >> gpurun_out/bwd_suite.log
tail -4 gpurun_out/bwd_suite.log
timeout 25 python tools/bench_train.py --steps 2 --warmup 1 > gpurun_out/train_bench.json 2> gpurun_out/train_bench.err
echo "exit $?" >> gpurun_out/train_bench.err
cat gpurun_out/train_bench.json; tail -3 gpurun_out/train_bench.err

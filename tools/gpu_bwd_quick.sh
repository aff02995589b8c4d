#!/bin/bash
# quickest possible on-device check of the backward: one small bag, every parameter gradient vs the float64 oracle
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_zz_backward_gpu.py -x -q -k "parameter_gradients and 512" > gpurun_out/bwd_quick.log 2>&1
echo "exit $?" >> gpurun_out/bwd_quick.log
tail -5 gpurun_out/bwd_quick.log

#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 90 python tools/timeline_pinv.py > gpurun_out/pinv_check.log 2>&1; tail -4 gpurun_out/pinv_check.log
timeout 300 python tools/batched_pinv.py > gpurun_out/batched_pinv.log 2>&1; cat gpurun_out/batched_pinv.log | tail -8

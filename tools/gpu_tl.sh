#!/bin/bash
set -x
mkdir -p gpurun_out
TM_EXTRA_NVCC_FLAGS=-DTM_PINV_TIMELINE python transmil-deepgraft_b200/build.py --force > gpurun_out/build_tl.log 2>&1
timeout 120 python tools/timeline_pinv.py > gpurun_out/timeline_push.log 2>&1; head -13 gpurun_out/timeline_push.log; tail -3 gpurun_out/timeline_push.log

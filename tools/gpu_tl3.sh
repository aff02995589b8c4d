#!/bin/bash
set -x
mkdir -p gpurun_out
TM_EXTRA_NVCC_FLAGS=-DTM_TIMELINE python transmil-deepgraft_b200/build.py --force > gpurun_out/build_tl.log 2>&1
timeout 120 python tools/timeline_nystrom_out.py > gpurun_out/timeline_nystrom_out.log 2>&1; head -26 gpurun_out/timeline_nystrom_out.log

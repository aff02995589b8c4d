#!/bin/bash
set -x
mkdir -p gpurun_out
TM_EXTRA_NVCC_FLAGS=-DTM_TIMELINE python transmil-deepgraft_b200/build.py --force > gpurun_out/build_tl.log 2>&1
timeout 120 python tools/timeline_attn3v.py > gpurun_out/timeline_attn3v.log 2>&1; head -30 gpurun_out/timeline_attn3v.log

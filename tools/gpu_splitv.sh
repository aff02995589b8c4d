#!/bin/bash
# A/B of projecting v beside the pinv chain (TM_SPLIT_V=0 = one qkv launch before the fork), then the GPU suite.
set -x
mkdir -p gpurun_out
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --clock-hold 0"
for rep in 1 2; do
TM_SPLIT_V=0 timeout 200 $B > gpurun_out/sv_off_$rep.json 2>> gpurun_out/sv.err
timeout 200 $B > gpurun_out/sv_on_$rep.json 2>> gpurun_out/sv.err
done
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/sv_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["ms_per_step"], d.get("stage_pass_ms_per_step"), d["gpu_launches"])
    except Exception as e: print(f, "failed", e)
P
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log

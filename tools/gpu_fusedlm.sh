#!/bin/bash
# A/B of the landmark sums fused into the q,k projection epilogue (TM_FUSED_LM=0 = separate pass), then the GPU suite.
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemm_tc_gpu.py -m gpu -x -q > gpurun_out/pytest_gemm.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gemm.log
tail -15 gpurun_out/pytest_gemm.log
B="python bench.py --steps 30 --warmup 5 --no-e2e --clock-hold 0"
for rep in 1 2; do
TM_FUSED_LM=0 timeout 200 $B > gpurun_out/fl_off_$rep.json 2>> gpurun_out/fl.err
timeout 200 $B > gpurun_out/fl_on_$rep.json 2>> gpurun_out/fl.err
done
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/fl_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["ms_per_step"], d.get("stage_pass_ms_per_step"), d["gpu_launches"], d.get("parity"), {k:v["ms_per_launch"] for k,v in d["stages"].items() if k in ("qkv","landmarks")})
    except Exception as e: print(f, "failed", e)
P
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log

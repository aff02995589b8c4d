#!/bin/bash
# Cluster-multicast GEMM variant (TM_GEMM_CLUSTER=1): correctness first (short timeouts: a hang must not hold the box), then A/B.
set -x
mkdir -p gpurun_out
TM_GEMM_CLUSTER=1 timeout 90 python -m pytest tests/test_gemm_tc_gpu.py -m gpu -x -q > gpurun_out/pytest_cl.log 2>&1; rc=$?; echo "pytest exit $rc" >> gpurun_out/pytest_cl.log
tail -6 gpurun_out/pytest_cl.log
if [ $rc -ne 0 ]; then exit 1; fi
for c in 0 1; do TM_GEMM_CLUSTER=$c timeout 60 python tools/time_gemms.py 2>&1 | sed "s/^/cluster=$c /" >> gpurun_out/gemm_cl.log; done
cat gpurun_out/gemm_cl.log
TM_GEMM_CLUSTER=1 timeout 200 python -m pytest tests/test_parity_gpu.py tests/test_f16_features_gpu.py -m gpu -x -q > gpurun_out/pytest_cl2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_cl2.log
tail -4 gpurun_out/pytest_cl2.log
B="python bench.py --steps 30 --warmup 5 --no-e2e --clock-hold 0"
for rep in 1 2; do
TM_GEMM_CLUSTER=0 timeout 100 $B > gpurun_out/cl_off_$rep.json 2>> gpurun_out/cl.err
TM_GEMM_CLUSTER=1 timeout 100 $B > gpurun_out/cl_on_$rep.json 2>> gpurun_out/cl.err
done
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/cl_o*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["ms_per_step"], d.get("stage_pass_ms_per_step"), d.get("parity"), {k:v["ms_per_launch"] for k,v in d["stages"].items() if k in ("qkv","fc1_gemm1")})
    except Exception as e: print(f, "failed", e)
P

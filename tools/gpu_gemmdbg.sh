#!/bin/bash
# GEMM timing on its own (tools/time_gemms.py; TM_GEMM_DBG switches parts of the TMA-store epilogue off), tests, bench.
set -x
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gemm_tc_gpu.py tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/pytest_gemm.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gemm.log
tail -3 gpurun_out/pytest_gemm.log
for d in ${DBG_MODES:-0}; do TM_GEMM_DBG=$d timeout 60 python tools/time_gemms.py >> gpurun_out/gemm_dbg.log 2>&1; done
cat gpurun_out/gemm_dbg.log
timeout 100 python bench.py --steps 30 --warmup 5 --no-e2e --clock-hold 0 > gpurun_out/b.json 2> gpurun_out/b.err
python -c "
import json; d=json.loads(open('gpurun_out/b.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['stage_pass_ms_per_step'], d.get('parity')); print({k:v['ms_per_launch'] for k,v in d['stages'].items()})"

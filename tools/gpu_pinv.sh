#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 90 python tools/timeline_pinv.py > gpurun_out/pinv_check.log 2>&1; echo "pinv check exit $?"; cat gpurun_out/pinv_check.log | tail -8
TM_PINV_KERNEL=tma timeout 90 python tools/timeline_pinv.py > gpurun_out/pinv_check_tma.log 2>&1; echo "pinv(tma) check exit $?"; tail -5 gpurun_out/pinv_check_tma.log
grep -q "max|Z" gpurun_out/pinv_check.log || exit 1
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['stage_pass_ms_per_step']); print({k:(v['ms_per_launch'],v['frac']) for k,v in d['stages'].items()})"

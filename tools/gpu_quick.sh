#!/bin/bash
# quick GPU check: pinv kernel sanity, the whole GPU suite, one bench line
set -x
mkdir -p gpurun_out
timeout 120 python tools/timeline_pinv.py > gpurun_out/pinv_check.log 2>&1; echo "pinv check exit $?"; cat gpurun_out/pinv_check.log
grep -q "max|Z" gpurun_out/pinv_check.log || exit 1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 $BENCH_ARGS > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['stage_pass_ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches'], d.get('parity')); print({k:(v['ms_per_launch'],v['frac']) for k,v in d['stages'].items()})"

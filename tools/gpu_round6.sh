#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 120 python tools/timeline_pinv.py > gpurun_out/pinv_check.log 2>&1; echo "pinv check exit $?"; cat gpurun_out/pinv_check.log
grep -q "max|Z" gpurun_out/pinv_check.log || exit 1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['stage_pass_ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches']); print({k:(v['ms_per_launch'],v['frac']) for k,v in d['stages'].items()})"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --clock-hold 0"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none -s 126 -c 46 -f -o gpurun_out/full $CMD > gpurun_out/ncu2.log 2>&1
ncu -i gpurun_out/full.ncu-rep --page raw --csv > gpurun_out/full_raw.csv 2>/dev/null
ls -la gpurun_out
SZ=$(stat -c %s gpurun_out/full.ncu-rep); if [ "$SZ" -gt 30000000 ]; then rm gpurun_out/full.ncu-rep; fi

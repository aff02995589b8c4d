#!/bin/bash
# the three generations of the pinv cluster kernel side by side, then the parity suite and a bench line on the default one
set -x
mkdir -p gpurun_out
for K in remote bulk tma; do
  TM_PINV_KERNEL=$K timeout 90 python tools/timeline_pinv.py > gpurun_out/pinv_check_$K.log 2>&1; echo "pinv($K) check exit $?"; tail -4 gpurun_out/pinv_check_$K.log
done
grep -q "max|Z" gpurun_out/pinv_check_remote.log || exit 1
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['stage_pass_ms_per_step']); print({k:(v['ms_per_launch'],v['frac']) for k,v in d['stages'].items()})"
TM_EXTRA_NVCC_FLAGS=-DTM_PINV_TIMELINE python transmil-deepgraft_b200/build.py --force > gpurun_out/build_tl.log 2>&1
timeout 120 python tools/timeline_pinv.py > gpurun_out/timeline_push.log 2>&1; head -13 gpurun_out/timeline_push.log

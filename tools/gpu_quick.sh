#!/bin/bash
# Quick confirmation of a build: the GPU suite and a short bench line (no e2e, no clock hold).
set -x
mkdir -p gpurun_out
timeout 150 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 60 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --clock-hold 0 > gpurun_out/b.json 2> gpurun_out/b.err
python -c "
import json; d=json.loads(open('gpurun_out/b.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['gpu_launches'])"

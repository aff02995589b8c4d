#!/bin/bash
# Bags data-parallel on 2 GPUs exactly as the driver launches it.
set -x
mkdir -p gpurun_out
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/dp2.json 2> gpurun_out/dp2.err; echo "dp2 exit $?"
tail -2 gpurun_out/dp2.err | cut -c1-300
python -c "
import json; d=json.loads(open('gpurun_out/dp2.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['clocks'])"

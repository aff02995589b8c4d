#!/bin/bash
# multi-GPU call: bench.py under torchrun at N = $1 GPUs (all sub-objects), the multi-GPU parity tests
N=${1:-2}
set -x
mkdir -p gpurun_out
nvidia-smi -L
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 \
    bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err ); echo "bench exit $?"
tail -5 gpurun_out/bench_n$N.err
python - <<PY
import json
l=json.loads([x for x in open('gpurun_out/bench_n$N.json') if x.startswith('{')][-1])
for k in ('value','ms_per_step','e2e','parity','train_step','cfg5','token_shard'):
    print(k, json.dumps(l.get(k))[:1800])
PY
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -q > gpurun_out/pytest_sharded_n$N.log 2>&1; echo "sharded tests exit $?"
tail -8 gpurun_out/pytest_sharded_n$N.log

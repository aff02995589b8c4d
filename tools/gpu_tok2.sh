#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded_gpu.py tests/test_ingest_gpu.py -m gpu -x -q > gpurun_out/pytest_sharded.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_sharded.log; tail -12 gpurun_out/pytest_sharded.log
for N in 1 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --shard tokens --patches 200000 --steps 20 --warmup 4 > gpurun_out/tok_$N.json 2> gpurun_out/tok_$N.err; echo "tokens $N exit $?"; tail -3 gpurun_out/tok_$N.err | cut -c1-300
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/tok_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d.get('parity'), d['gpu_launches'])
    except Exception as e: print(f, 'ERR', e)
PY

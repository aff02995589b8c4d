#!/bin/bash
# gpurun --gpus 8: scaling of both multi-GPU modes + the multi-rank sharded parity test
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q -k "multi_gpu" > gpurun_out/pytest_multi.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi.log; tail -4 gpurun_out/pytest_multi.log
for N in 1 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/scale_bags_$N.json 2> gpurun_out/scale_bags_$N.err; echo "bags $N exit $?"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --shard tokens --patches 200000 --steps 10 --warmup 3 > gpurun_out/scale_tokens_$N.json 2> gpurun_out/scale_tokens_$N.err; echo "tokens $N exit $?"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$N bench.py --gpus $N --shard tokens --patches 200000 --features 2048 --steps 10 --warmup 3 > gpurun_out/scale_tokens2048_$N.json 2> gpurun_out/scale_tokens2048_$N.err; echo "tokens2048 $N exit $?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/scale_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d.get('parity'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/scale_tokens_8.err

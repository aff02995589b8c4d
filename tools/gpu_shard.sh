#!/bin/bash
# token-sharded bag on N = $1 GPUs: parity tests (peer-memory transport eager + graph, NCCL transport), then the strong-scaling
# point of BASELINE config 4 (200k x 2048) and of the 200k x 1024 bag.  Every step under its own timeout.
N=${1:-2}
set -x
mkdir -p gpurun_out
nvidia-smi -L | head -2
nvidia-smi topo -m 2>/dev/null | head -12
timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q > gpurun_out/pytest_sharded_n$N.log 2>&1; echo "sharded tests exit $?"
tail -15 gpurun_out/pytest_sharded_n$N.log
for D in 2048 1024; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 \
      bench.py --gpus $N --shard tokens --n 200000 --features $D --steps 20 --warmup 3 > gpurun_out/shard_n${N}_$D.json 2> gpurun_out/shard_n${N}_$D.err
  echo "shard bench D=$D exit $?"; tail -3 gpurun_out/shard_n${N}_$D.err
  python - <<PY
import json
try:
    l=json.loads([x for x in open('gpurun_out/shard_n${N}_$D.json') if x.startswith('{')][-1])
    t=l['token_shard']
    print({k:t.get(k) for k in ('ms_per_step','single_gpu_ms_per_step','speedup_vs_single_gpu','variants_ms_per_step','variants_launches_per_step_rank0','exchange_timeouts','graph_equals_eager_bits','max_abs_dlogit_peer_vs_nccl','parity','e2e')})
except Exception as e:
    print('no line', e)
PY
done

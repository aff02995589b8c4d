#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['stage_pass_ms_per_step'], d['e2e']['ms_per_step'], d['clocks'], d.get('parity')); print({k:(v['ms_per_launch'],v['frac']) for k,v in d['stages'].items()})"
for N in 1 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$N bench.py --gpus $N --shard tokens --patches 200000 --steps 10 --warmup 3 > gpurun_out/bench_tok$N.json 2> gpurun_out/bench_tok$N.err; echo "tok$N exit $?"; tail -c 700 gpurun_out/bench_tok$N.json; tail -3 gpurun_out/bench_tok$N.err
done
TM_EXTRA_NVCC_FLAGS=-DTM_PINV_TIMELINE python transmil-deepgraft_b200/build.py --force > gpurun_out/build_tl.log 2>&1
python tools/timeline_pinv.py > gpurun_out/timeline_pinv.log 2>&1; cat gpurun_out/timeline_pinv.log

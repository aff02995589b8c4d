#!/bin/bash
# gpurun --gpus 2: GPU parity suite (incl. sharded), bench (bag-dp N=1), token-sharded bench N=1,2, ncu of PPEG.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print(d['ms_per_step'], d['e2e'], d['clocks'], d.get('parity')); print({k:(v['ms_per_launch'],v['frac']) for k,v in d['stages'].items()})"
python bench.py --shard tokens --n 200000 --steps 10 --warmup 3 > gpurun_out/bench_tok1.json 2> gpurun_out/bench_tok1.err; echo "tok1 exit $?"; tail -c 1500 gpurun_out/bench_tok1.json; tail -5 gpurun_out/bench_tok1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --shard tokens --n 200000 --steps 10 --warmup 3 > gpurun_out/bench_tok2.json 2> gpurun_out/bench_tok2.err; echo "tok2 exit $?"; tail -c 1500 gpurun_out/bench_tok2.json; tail -5 gpurun_out/bench_tok2.err
python bench.py --n 200000 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_200k_1gpu.json 2>> gpurun_out/bench.err; tail -c 800 gpurun_out/bench_200k_1gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_dp2.json 2> gpurun_out/bench_dp2.err; echo "dp2 exit $?"; tail -c 600 gpurun_out/bench_dp2.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --clock-hold 0"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:ppeg_kernel -c 2 -f -o gpurun_out/ppeg $CMD > gpurun_out/ncu_ppeg.log 2>&1
ncu -i gpurun_out/ppeg.ncu-rep --page raw --csv > gpurun_out/ppeg_raw.csv 2>/dev/null
ls -la gpurun_out

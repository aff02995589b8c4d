#!/bin/bash
# The one GPU-lease script: every measurement under profiles/ comes from one of these modes.  Each step runs under its own
# `timeout`; outputs land in gpurun_out/ (copied into profiles/ by hand once read).
#   gpurun --timeout 1500 -- 'bash tools/gpu_evidence.sh tests'          full GPU suite (risky kernels first, own time-outs)
#   gpurun --timeout 900  -- 'bash tools/gpu_evidence.sh bench'          both bench arms + smoke, as the driver runs them
#   gpurun --timeout 1200 -- 'bash tools/gpu_evidence.sh ncu'            forward: launch list + one --set full capture (CSV)
#   gpurun --timeout 900  -- 'bash tools/gpu_evidence.sh train'          training step: numbers + ncu launch lists
#   gpurun --timeout 900  -- 'bash tools/gpu_evidence.sh tools'          D6 torch eager, cfg5, files->logits, sweep, budgets
#   gpurun --gpus N --timeout 1200 -- 'bash tools/gpu_evidence.sh multi N'   bench.py under torchrun + sharded parity tests
#   gpurun --gpus N --timeout 1200 -- 'bash tools/gpu_evidence.sh shard N'   token-sharded bag: tests + strong-scaling points (shardbench N: the points only)
#   gpurun --timeout 400 -- 'bash tools/gpu_evidence.sh pinv'             pinv cluster kernel: parity test + time per generation
#   gpurun --gpus N --timeout 600  -- 'bash tools/gpu_evidence.sh shardprof N [patches]'   per-stage times of the sharded forward
MODE=${1:-tests}
N=${2:-2}
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/smi.txt
show() { python - "$@" <<'PY'
import json, sys
path, keys = sys.argv[1], sys.argv[2:]
try:
    l = json.loads([x for x in open(path) if x.startswith('{')][-1])
    for k in keys:
        print(k, json.dumps(l.get(k))[:1800])
except Exception as e:
    print('no JSON line in', path, e)
PY
}
case $MODE in
tests)
  timeout 300 python -m pytest tests/test_bgemm_tc_gpu.py tests/test_backward_fast_gpu.py -m gpu -q > gpurun_out/pytest_bgemm.log 2>&1; echo "bgemm exit $?" | tee -a gpurun_out/pytest_bgemm.log
  tail -5 gpurun_out/pytest_bgemm.log
  timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_bgemm_tc_gpu.py --deselect tests/test_backward_fast_gpu.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
  tail -30 gpurun_out/pytest_gpu.log
  ;;
bench)
  timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench.err
  ( time timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2>> gpurun_out/bench.err ); echo "bench exit $?"
  tail -5 gpurun_out/bench.err
  show gpurun_out/bench.json value ms_per_step e2e parity roofline cpu_baseline torch_eager train_step cfg5 token_shard stages
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
  ;;
ncu)
  CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras --clock-hold 0"
  $CMD > gpurun_out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
  SKIP=$(python - <<'P'
import csv
rows = [r for r in csv.reader(l for l in open('gpurun_out/launches.csv') if l.startswith('"'))]
h = rows[0]; k = h.index('Kernel Name'); i = h.index('ID')
ids = [int(r[i]) for r in rows[1:] if 'prep_operands' in r[k]]
print(ids[2] if len(ids) > 2 else 0)
P
)
  echo "skip $SKIP"
  $CMD > gpurun_out/plain2.log 2>&1 &&
  timeout 700 ncu --set full --import-source on --clock-control none -s $SKIP -c 58 -f -o gpurun_out/full $CMD > gpurun_out/ncu2.log 2>&1
  ncu -i gpurun_out/full.ncu-rep --page raw --csv > gpurun_out/full_raw.csv 2>/dev/null
  SZ=$(stat -c %s gpurun_out/full.ncu-rep); if [ "$SZ" -gt 30000000 ]; then rm gpurun_out/full.ncu-rep; fi
  ;;
train)
  timeout 300 python tools/bench_train.py --steps 5 --warmup 3 > gpurun_out/train_64x1000.json 2> gpurun_out/train.err; echo "train exit $?"
  timeout 300 python tools/bench_train.py --bags 1 --patches 20000 --steps 5 --warmup 3 > gpurun_out/train_1x20000.json 2>> gpurun_out/train.err
  cat gpurun_out/train_*.json; tail -5 gpurun_out/train.err
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/train_launches.csv \
      python tools/bench_train.py --steps 1 --warmup 1 > gpurun_out/train_ncu.log 2>&1; echo "train ncu exit $?"
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/train_launches_20k.csv \
      python tools/bench_train.py --bags 1 --patches 20000 --steps 1 --warmup 1 > gpurun_out/train_ncu20k.log 2>&1; echo "train ncu 20k exit $?"
  python tools/launch_summary.py gpurun_out/train_launches.csv > gpurun_out/train_launch_summary.txt 2>&1; head -40 gpurun_out/train_launch_summary.txt
  # one --set full capture of the training step's top kernel (a handful of launches from the middle of the second step)
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:bgemm_tc_kernel -s 400 -c 8 -f -o gpurun_out/train_bgemm_full \
      python tools/bench_train.py --steps 1 --warmup 1 > gpurun_out/train_ncu_full.log 2>&1; echo "train ncu full exit $?"
  ncu -i gpurun_out/train_bgemm_full.ncu-rep --page raw --csv > gpurun_out/train_bgemm_full_raw.csv 2>/dev/null
  SZ=$(stat -c %s gpurun_out/train_bgemm_full.ncu-rep 2>/dev/null || echo 0); if [ "$SZ" -gt 30000000 ]; then rm gpurun_out/train_bgemm_full.ncu-rep; fi
  ;;
tools)
  timeout 400 python tools/bench_torch_eager.py --n 50000 --features 1024 > gpurun_out/torch_eager_50k_1024.jsonl 2> gpurun_out/eager.err; echo "eager exit $?"
  timeout 300 python tools/bench_torch_eager.py --n 20000 --features 2048 > gpurun_out/torch_eager_20k_2048.jsonl 2>> gpurun_out/eager.err
  timeout 300 python tools/bench_cfg5.py --features 2048 > gpurun_out/cfg5_2048.json 2> gpurun_out/cfg5.err; echo "cfg5 exit $?"
  timeout 300 python tools/bench_files.py > gpurun_out/files_to_logits.jsonl 2> gpurun_out/files.err; echo "files exit $?"
  timeout 300 python tools/stage_budget.py > gpurun_out/stage_budget.txt 2> gpurun_out/stage_budget.err; echo "stage budget exit $?"; tail -40 gpurun_out/stage_budget.txt
  timeout 400 python tools/sweep.py > gpurun_out/sweep.jsonl 2> gpurun_out/sweep.err; echo "sweep exit $?"
  ;;
multi)
  ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 \
      bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err ); echo "bench exit $?"
  tail -5 gpurun_out/bench_n$N.err
  show gpurun_out/bench_n$N.json value ms_per_step e2e parity train_step cfg5 token_shard
  timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -q > gpurun_out/pytest_sharded_n$N.log 2>&1; echo "sharded tests exit $?"
  tail -8 gpurun_out/pytest_sharded_n$N.log
  ;;
shard|shardbench)
  if [ $MODE = shard ]; then
    nvidia-smi topo -m 2>/dev/null | head -12
    timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q > gpurun_out/pytest_sharded_n$N.log 2>&1; echo "sharded tests exit $?"
    tail -15 gpurun_out/pytest_sharded_n$N.log
  fi
  for D in 2048 1024; do
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 \
        bench.py --gpus $N --shard tokens --patches 200000 --features $D --steps 20 --warmup 3 > gpurun_out/shard_n${N}_$D.json 2> gpurun_out/shard_n${N}_$D.err
    echo "shard bench D=$D exit $?"; tail -3 gpurun_out/shard_n${N}_$D.err
    show gpurun_out/shard_n${N}_$D.json value ms_per_step token_shard
  done
  ;;
pinv)
  timeout 120 python -m pytest tests/test_pinv_gpu.py -m gpu -x -q 2>&1 | tail -15
  timeout 120 python tools/pinv_check.py 2>&1 | tail -5
  ;;
shardprof)
  P=${3:-200000}
  for D in 2048 1024; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29557 \
        tools/shard_profile.py --patches $P --features $D > gpurun_out/shard_profile_n${N}_${P}_$D.txt 2> gpurun_out/shard_profile.err
    echo "shard profile D=$D exit $?"; tail -3 gpurun_out/shard_profile.err; cat gpurun_out/shard_profile_n${N}_${P}_$D.txt
  done
  ;;
*) echo "unknown mode $MODE"; exit 2 ;;
esac
ls -la gpurun_out | tail -30

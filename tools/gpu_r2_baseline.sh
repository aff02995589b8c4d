#!/bin/bash
# Round 2, first GPU call (one B200): the new tcgen05 backward GEMM, then everything round 1 left unmeasured.
#   gpurun --timeout 1500 -- 'bash tools/gpu_r2_baseline.sh'
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/smi.txt
nproc >> gpurun_out/smi.txt
timeout 240 python -m pytest tests/test_bgemm_tc_gpu.py -m gpu -q -x > gpurun_out/pytest_bgemm.log 2>&1; echo "bgemm exit $?" | tee -a gpurun_out/pytest_bgemm.log
tail -25 gpurun_out/pytest_bgemm.log
if ! grep -q "bgemm exit 0" gpurun_out/pytest_bgemm.log; then export TM_BGEMM=simt; echo "bgemm_tc failed: the rest runs with TM_BGEMM=simt"; fi
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_bgemm_tc_gpu.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_train.py --steps 3 --warmup 2 > gpurun_out/train_64x1000.json 2> gpurun_out/train.err; echo "train exit $?"
timeout 300 python tools/bench_train.py --bags 1 --patches 20000 --steps 3 --warmup 2 > gpurun_out/train_1x20000.json 2>> gpurun_out/train.err
timeout 300 python tools/bench_train.py --engine simt --steps 2 --warmup 1 > gpurun_out/train_64x1000_simt.json 2>> gpurun_out/train.err
cat gpurun_out/train_*.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/train_launches.csv \
    python tools/bench_train.py --steps 1 --warmup 1 > gpurun_out/train_ncu.log 2>&1; echo "train ncu exit $?"
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
timeout 400 python tools/bench_torch_eager.py --n 50000 --features 1024 > gpurun_out/torch_eager_50k_1024.jsonl 2> gpurun_out/eager.err; echo "eager exit $?"
timeout 300 python tools/bench_torch_eager.py --n 20000 --features 2048 > gpurun_out/torch_eager_20k_2048.jsonl 2>> gpurun_out/eager.err
timeout 300 python tools/bench_cfg5.py --features 2048 > gpurun_out/cfg5_2048.json 2> gpurun_out/cfg5.err; echo "cfg5 exit $?"
timeout 300 python tools/bench_files.py > gpurun_out/files_to_logits.jsonl 2> gpurun_out/files.err; echo "files exit $?"
ls -la gpurun_out

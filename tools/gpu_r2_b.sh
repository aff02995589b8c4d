#!/bin/bash
# Round 2, second GPU call: TMA feed of the backward GEMM (incl. MN-major tf32 operands), backward suite per engine, training step.
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_bgemm_tc_gpu.py tests/test_backward_fast_gpu.py -m gpu -q > gpurun_out/pytest_bgemm.log 2>&1; echo "bgemm exit $?" | tee -a gpurun_out/pytest_bgemm.log
tail -40 gpurun_out/pytest_bgemm.log
if grep -q "tma" gpurun_out/pytest_bgemm.log && grep -q "failed" gpurun_out/pytest_bgemm.log; then export TM_BGEMM_LOADER=cpasync; echo "TMA feed failed: the rest runs with TM_BGEMM_LOADER=cpasync"; fi
timeout 600 python -m pytest tests/test_zz_backward_gpu.py tests/test_zzz_gradcam_hooks_gpu.py tests/test_zzz_stem768_backward_gpu.py -m gpu -q > gpurun_out/pytest_bwd.log 2>&1; echo "bwd exit $?" | tee -a gpurun_out/pytest_bwd.log
tail -30 gpurun_out/pytest_bwd.log
timeout 300 python tools/bench_train.py --steps 3 --warmup 2 > gpurun_out/train_64x1000.json 2> gpurun_out/train.err; echo "train exit $?"
timeout 300 python tools/bench_train.py --bags 1 --patches 20000 --steps 3 --warmup 2 > gpurun_out/train_1x20000.json 2>> gpurun_out/train.err
cat gpurun_out/train_*.json; tail -5 gpurun_out/train.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/train_launches.csv \
    python tools/bench_train.py --steps 1 --warmup 1 > gpurun_out/train_ncu.log 2>&1; echo "train ncu exit $?"
ls -la gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/train_launches_20k.csv \
    python tools/bench_train.py --bags 1 --patches 20000 --steps 1 --warmup 1 > gpurun_out/train_ncu20k.log 2>&1; echo "train ncu 20k exit $?"

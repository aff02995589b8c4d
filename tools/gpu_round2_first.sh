#!/bin/bash
# First GPU call of the next round (one B200): everything that was written after round 1's GPU budget was spent and has only
# run on the host so far -- the tests sorted last (MDMIL, file ingest, 768-d stem backward, GradCAM hooks), then the whole GPU
# suite, the bench, and the two measurement tools that have no number yet (config 5 variable-length bags, stock-PyTorch baseline).
#   gpurun --timeout 1500 -- 'bash tools/gpu_round2_first.sh'
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zzz_mdmil_gpu.py tests/test_zzz_h5_ingest_gpu.py tests/test_zzz_stem768_backward_gpu.py \
    tests/test_zzz_gradcam_hooks_gpu.py tests/test_zzz_sharp_softmax_gpu.py tests/test_zzz_batch64_gpu.py -s -m gpu -q > gpurun_out/pytest_new.log 2>&1; echo "new tests exit $?" | tee -a gpurun_out/pytest_new.log
tail -15 gpurun_out/pytest_new.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
timeout 300 python tools/bench_cfg5.py --features 2048 > gpurun_out/cfg5_2048.json 2> gpurun_out/cfg5.err; echo "cfg5 exit $?"
timeout 300 python tools/bench_cfg5.py --features 1024 --f16 > gpurun_out/cfg5_1024_f16.json 2>> gpurun_out/cfg5.err
timeout 400 python tools/bench_torch_eager.py --n 50000 --features 1024 > gpurun_out/torch_eager_50k_1024.jsonl 2> gpurun_out/eager.err; echo "eager exit $?"
timeout 300 python tools/bench_torch_eager.py --n 20000 --features 2048 > gpurun_out/torch_eager_20k_2048.jsonl 2>> gpurun_out/eager.err
ls -la gpurun_out
timeout 400 python tools/bench_files.py > gpurun_out/files_to_logits.jsonl 2> gpurun_out/files.err; echo "files exit $?"

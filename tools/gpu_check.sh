#!/bin/bash
# New tests of this build and a plain run of the small-forward script.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pdl_gpu.py tests/test_f16_features_gpu.py -x -q > gpurun_out/pytest_new.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_new.log
tail -5 gpurun_out/pytest_new.log
timeout 200 python tools/sanitize_fwd.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
# compute-sanitizer is closed on this pool ("runs under it have left GPUs needing a reset"): the bit-equivalence test of
# tests/test_pdl_gpu.py and the oracle parity tests are the race / bounds evidence instead.

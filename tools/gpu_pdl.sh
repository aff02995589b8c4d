#!/bin/bash
# A/B of programmatic dependent launch: same build, TM_PDL=0 vs variants of which launches around the two-stream fork use it.
set -x
mkdir -p gpurun_out
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --clock-hold 0"
for rep in 1 2; do
TM_PDL=0 timeout 200 $B > gpurun_out/pdl_off_$rep.json 2>> gpurun_out/pdl.err
for x in 29 61 93 157 285 509 31; do
TM_PDL_X=$x timeout 200 $B > gpurun_out/pdl_x${x}_$rep.json 2>> gpurun_out/pdl.err
done
done
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/pdl_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["ms_per_step"], d.get("stage_pass_ms_per_step"))
    except Exception as e: print(f, "failed", e)
P

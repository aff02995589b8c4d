#!/bin/bash
# Round-end evidence run on one B200: full GPU suite, both bench arms, ncu launch list, one `--set full` capture of a
# warm forward exported as CSV on the box (the .ncu-rep is deleted when it would not fit the 64 MiB return limit), sweep.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench.err
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2>> gpurun_out/bench.err; echo "bench exit $?"
python -c "
import json; d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['stage_pass_ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches'], d.get('parity')); print({k:(v['ms_per_launch'],v['frac']) for k,v in d['stages'].items()})"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --clock-hold 0"
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
timeout 600 ncu --set full --clock-control none -s $SKIP -c 58 -f -o gpurun_out/full $CMD > gpurun_out/ncu2.log 2>&1
ncu -i gpurun_out/full.ncu-rep --page raw --csv > gpurun_out/full_raw.csv 2>/dev/null
SZ=$(stat -c %s gpurun_out/full.ncu-rep); if [ "$SZ" -gt 30000000 ]; then rm gpurun_out/full.ncu-rep; fi
timeout 400 python tools/sweep.py > gpurun_out/sweep.jsonl 2> gpurun_out/sweep.err; echo "sweep exit $?"
ls -la gpurun_out

#!/bin/bash
# bench.py with every sub-object on one GPU (and the reference arm), as the driver runs it
set -x
mkdir -p gpurun_out
( time timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err ); echo "bench exit $?"
tail -5 gpurun_out/bench.err
python - <<'PY'
import json
l=json.load(open('gpurun_out/bench.json'))
for k in ('value','ms_per_step','e2e','parity','roofline','cpu_baseline','torch_eager','train_step','cfg5','token_shard'):
    print(k, json.dumps(l.get(k))[:1500])
print(json.dumps(l.get('stages'))[:3000])
PY

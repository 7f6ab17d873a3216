#!/bin/bash
# A/B runs of bench.py under environment switches (GPU box): usage tools/ab_bench.sh <tag> "<ENV=.. ENV=..>" ...
TAG=$1; shift
i=0
mkdir -p gpurun_out
for envs in "$@"; do
  env $envs python bench.py --steps 5 --warmup 3 --parity off --no-cpu --e2e-steps 3 > gpurun_out/${TAG}_ab$i.json 2> gpurun_out/${TAG}_ab$i.err
  python - gpurun_out/${TAG}_ab$i.json "$envs" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], '|', round(d['value']/1e9,2), 'G/s', round(d['ms_per_step'],2), 'ms e2e', round(d['e2e']['value']/1e9,2), d['kernel_ms'], 'ok' if d['conservation']['ok'] else 'BAD')
    for k,v in d['per_k'].items(): print('   k',k, v['ms'], 'retries', v['table_retries'])
except Exception as e:
    print(sys.argv[2], 'FAILED', e)
PY
  i=$((i+1))
done

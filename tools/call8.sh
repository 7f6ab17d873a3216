#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
run() { tag=$1; shift; env $ENVS python bench.py --steps 3 --warmup 2 --parity off --no-cpu --e2e-steps 1 "$@" > $OUT/c8_$tag.json 2> $OUT/c8_$tag.err
python - $OUT/c8_$tag.json "$tag $ENVS $*" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], '|', round(d['value']/1e9,2), 'G/s', round(d['ms_per_step'],2), 'ms')
    for k,v in d['per_k'].items(): print('   k',k, v['ms'], 'retries', v['table_retries'], 'P', v['partitions'])
except Exception as e:
    print(sys.argv[2], 'FAILED', e)
PY
}
ENVS="KC_X=0" run p94k --partitions 94000
ENVS="KC_X=0" run p23k --partitions 23500
ENVS="KC_BATCH_READS=4194304" run b4m
ENVS="KC_BATCH_READS=1048576" run b1m

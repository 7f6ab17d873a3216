#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_kcount_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/c3_pytest.log 2>&1; tail -3 $OUT/c3_pytest.log
bash tools/ab_bench.sh c3 "KC_X=0" "KCOUNT_B200_LIB=/root/repo/build_variants/libkcount_noq.so"
for cfg in C1 C5 C4; do
  python bench.py --config $cfg --steps 5 --warmup 3 --parity off --no-cpu --e2e-steps 3 > $OUT/c3_$cfg.json 2> $OUT/c3_$cfg.err
  python - $OUT/c3_$cfg.json $cfg <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], '|', round(d['value']/1e9,2), 'G/s', round(d['ms_per_step'],2), 'ms e2e', round(d['e2e']['value']/1e9,2), d['kernel_ms'], 'ok' if d['conservation']['ok'] else 'BAD')
    for k,v in d['per_k'].items(): print('   k',k, v['ms'], 'retries', v['table_retries'])
except Exception as e:
    print(sys.argv[2], 'FAILED', e)
PY
done

#!/bin/bash
bash tools/ab_bench.sh c10 "KC_X=0" "KCOUNT_B200_LIB=/root/repo/build_variants/libkcount_t512.so" "KCOUNT_B200_LIB=/root/repo/build_variants/libkcount_d16.so"
python - <<'PY'
import json
for i in range(3):
    try:
        d=json.loads(open(f"gpurun_out/c10_ab{i}.json").read().strip().splitlines()[-1])
        print(i, {k:(v['survivors'],v['distinct'],v['partitions'],v['table_retries']) for k,v in d['per_k'].items()})
    except Exception as e: print(i,'failed',e)
PY
KCOUNT_B200_LIB=/root/repo/build_variants/libkcount_t512.so python -m pytest tests/test_kcount_gpu.py -m gpu -x -q 2>&1 | tail -3

#!/bin/bash
# A/B + count-kernel capture (scratch script for one gpurun call)
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_kcount_gpu.py -m gpu -x -q > $OUT/c2_pytest_default.log 2>&1; tail -2 $OUT/c2_pytest_default.log
KC_REC_MIN_WORDS=8 python -m pytest tests/test_kcount_gpu.py -m gpu -x -q > $OUT/c2_pytest_rec32.log 2>&1; tail -2 $OUT/c2_pytest_rec32.log
bash tools/ab_bench.sh c2 "KC_X=0" "KC_REC_MIN_WORDS=8" 
timeout 900 ncu --set full --clock-control none --import-source on -k regex:count_kernel -c 8 -f -o /tmp/c2_full python tools/prof_run.py 13333334 21,33 > $OUT/c2_ncu_full.log 2>&1
echo "full capture rc=$?"
ncu -i /tmp/c2_full.ncu-rep --page raw --csv > $OUT/c2_count_raw.csv 2>/dev/null
ncu -i /tmp/c2_full.ncu-rep --page source --csv --print-source cuda,sass > $OUT/c2_count_src.csv 2>/dev/null
gzip -f $OUT/c2_count_src.csv
ls -la $OUT | tail -8

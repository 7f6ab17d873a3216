#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 600 ncu --set full --clock-control none --import-source on -k regex:count_kernel --launch-skip 2 -c 1 -f -o /tmp/c4_full python tools/prof_run.py 13333334 21 > $OUT/c4_ncu_full.log 2>&1
echo "full capture rc=$?"
ncu -i /tmp/c4_full.ncu-rep --page raw --csv > $OUT/c4_count_raw.csv 2>/dev/null
ncu -i /tmp/c4_full.ncu-rep --page source --csv --print-source cuda,sass > $OUT/c4_count_src.csv 2>/dev/null
gzip -f $OUT/c4_count_src.csv

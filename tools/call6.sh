#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_kcount_gpu.py -m gpu -x -q > $OUT/c6_pytest.log 2>&1; tail -3 $OUT/c6_pytest.log
bash tools/ab_bench.sh c6 "KC_X=0"
KC_TRACE=1 python tools/hostprof.py 13333334 > $OUT/c6_hostprof.log 2> $OUT/c6_hostprof.err
tail -12 $OUT/c6_hostprof.log

#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_kcount_gpu.py -m gpu -x -q > $OUT/c5_pytest.log 2>&1; tail -3 $OUT/c5_pytest.log
bash tools/ab_bench.sh c5 "KC_X=0"

#!/bin/bash
# Evidence for profiles/ (run on the GPU box through gpurun, one GPU): usage tools/collect_profiles.sh <tag>
#   1. bench.py without a profiler (the numbers that count)
#   2. the ncu launch list of the same command (shares per kernel; absolute times are cold-cache and serialised)
#   3. one `ncu --set full` capture of the count / parse / fill kernels at the bench workload, exported as CSV pages
set -u
TAG=${1:-vX}
OUT=gpurun_out
mkdir -p $OUT
python bench.py --steps ${STEPS:-10} --warmup 3 ${BENCH_EXTRA:-} > $OUT/${TAG}_bench_n1.json 2> $OUT/${TAG}_bench_n1.err || exit 1
tail -c 600 $OUT/${TAG}_bench_n1.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --steps 3 --warmup 3 --parity off --no-cpu --e2e-steps 1 > $OUT/${TAG}_ncu_bench.log 2>&1
echo "launch list rc=$?"
python tools/prof_run.py 13333334 21,33 > $OUT/${TAG}_prof_run.log 2>&1 || exit 1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'count_kernel|parse_kernel|fill_kernel' \
  --launch-skip 0 -c 16 -f -o /tmp/${TAG}_full python tools/prof_run.py 13333334 21,33 > $OUT/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > $OUT/${TAG}_full_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_full.ncu-rep --page source --csv --print-source cuda,sass > $OUT/${TAG}_full_src.csv 2>/dev/null
ncu -i /tmp/${TAG}_full.ncu-rep --page source --csv --print-source cuda > $OUT/${TAG}_full_text.csv 2>/dev/null
gzip -f $OUT/${TAG}_full_src.csv $OUT/${TAG}_full_text.csv
[ $(stat -c %s /tmp/${TAG}_full.ncu-rep) -lt 25000000 ] && cp /tmp/${TAG}_full.ncu-rep $OUT/
ls -la $OUT | tail -12

#!/bin/bash
# Evidence for profiles/ (run on the GPU box through gpurun, one GPU): usage tools/collect_profiles.sh <tag>
#   1. bench.py without a profiler (the numbers that count)
#   2. the ncu launch list of the same command (shares per kernel; absolute times are cold-cache and serialised)
#   3. `ncu --set full` captures at the bench workload, exported as CSV pages: the count kernel's launches (probes + main,
#      k=21 and k=33) and the first parse / fill launches; profiles/traffic.json is regenerated from the raw pages
set -u
TAG=${1:-vX}
OUT=gpurun_out
mkdir -p $OUT
python bench.py --steps ${STEPS:-10} --warmup 3 ${BENCH_EXTRA:-} > $OUT/${TAG}_bench_n1.json 2> $OUT/${TAG}_bench_n1.err || exit 1
tail -c 600 $OUT/${TAG}_bench_n1.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --steps 3 --warmup 3 --parity off --no-cpu --e2e-steps 1 > $OUT/${TAG}_ncu_bench.log 2>&1
echo "launch list rc=$?"
python tools/prof_run.py 13333334 21,33 > $OUT/${TAG}_prof_run.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:count_kernel -c 8 -f -o /tmp/${TAG}_count \
  python tools/prof_run.py 13333334 21,33 > $OUT/${TAG}_ncu_count.log 2>&1
echo "count capture rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'parse_kernel|fill_kernel' -c 4 -f -o /tmp/${TAG}_parse \
  python tools/prof_run.py 13333334 21 > $OUT/${TAG}_ncu_parse.log 2>&1
echo "parse/fill capture rc=$?"
for w in count parse; do
  ncu -i /tmp/${TAG}_$w.ncu-rep --page raw --csv > $OUT/${TAG}_${w}_raw.csv 2>/dev/null
  ncu -i /tmp/${TAG}_$w.ncu-rep --page source --csv --print-source cuda,sass > $OUT/${TAG}_${w}_src.csv 2>/dev/null
  gzip -f $OUT/${TAG}_${w}_src.csv
done
python tools/make_traffic.py $TAG $OUT/${TAG}_count_raw.csv $OUT/${TAG}_parse_raw.csv && cp profiles/traffic.json $OUT/${TAG}_traffic.json
ls -la $OUT | tail -14

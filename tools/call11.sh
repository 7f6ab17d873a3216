#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus 4 --config C3 --second --steps 5 --warmup 3 --e2e-steps 5 > $OUT/c11_bench_c3_n4.json 2> $OUT/c11_bench_c3_n4.err; echo "bench c3 n4 rc=$?"
$TR --master-port 29522 tools/hostprof_e2e_dist.py > $OUT/c11_hostprof_e2e_n4.log 2>&1; echo "hostprof rc=$?"
grep -E "^\[copy\] rank 0|rank 0 rep 3" $OUT/c11_hostprof_e2e_n4.log | cut -c1-700
python - <<'PY'
import json
for f in ("gpurun_out/c11_bench_c3_n4.json",):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value']/1e9,2),'G/s',round(d['ms_per_step'],2),'ms e2e',round(d['e2e']['value']/1e9,2), round(d['e2e']['ms_per_step'],1), 'parity',d['parity'] and (d['parity']['checked'],d['parity']['equal']), 'cons',d['conservation']['ok'], 'gen_s', d['gen_s'])
        for k,v in d['per_k'].items(): print('   k',k,v['ms'],'retries',v['table_retries'])
        print('   exch',d['exchange']); print('   second', d.get('c2_weak_scaled'))
    except Exception as e: print(f,'FAILED',e)
PY

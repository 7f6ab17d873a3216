#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 4 --steps 10 --warmup 3 --e2e-steps 5 > $OUT/r2_v10_bench_n4.json 2> $OUT/r2_v10_bench_n4.err; echo "n4 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 2 --steps 10 --warmup 3 --e2e-steps 5 > $OUT/r2_v10_bench_n2.json 2> $OUT/r2_v10_bench_n2.err; echo "n2 rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r2_v10_bench_n4.json","gpurun_out/r2_v10_bench_n2.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value']/1e9,2),'G/s',round(d['ms_per_step'],2),'ms e2e',round(d['e2e']['value']/1e9,2), round(d['e2e']['ms_per_step'],1), 'parity',d['parity'] and (d['parity']['checked'],d['parity']['equal']), 'cons',d['conservation']['ok'])
        for k,v in d['per_k'].items(): print('   k',k,v['ms'],'retries',v['table_retries'])
        print('   exch',{k:(round(v['gbs_per_dir'],1),v['ms_comm'],v['ms_not_hidden'],v['rounds']) for k,v in d['exchange'].items()})
        print('   ceiling', {k:v for k,v in d.get('e2e_copy_ceiling').items() if 'gbs' in k or 'floor_ms_per' in k})
    except Exception as e: print(f,'FAILED',e)
PY

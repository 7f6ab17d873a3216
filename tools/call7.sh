#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_kcount_gpu.py -m gpu -x -q -k "nccl" > $OUT/c7_pytest_nccl.log 2>&1; tail -3 $OUT/c7_pytest_nccl.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > $OUT/c7_bench_n2.json 2> $OUT/c7_bench_n2.err; echo "bench n2 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --config C3 --reads 8000000 --steps 3 --warmup 3 --e2e-steps 3 > $OUT/c7_bench_c3_n2.json 2> $OUT/c7_bench_c3_n2.err; echo "bench c3 n2 rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/c7_bench_n2.json","gpurun_out/c7_bench_c3_n2.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value']/1e9,2),'G/s',round(d['ms_per_step'],2),'ms e2e',round(d['e2e']['value']/1e9,2), round(d['e2e']['ms_per_step'],1), 'parity',d['parity'] and (d['parity']['checked'],d['parity']['equal']), 'cons',d['conservation']['ok'])
        for k,v in d['per_k'].items(): print('   k',k,v['ms'],'retries',v['table_retries'])
        print('   exch',d['exchange'])
    except Exception as e: print(f,'FAILED',e)
PY

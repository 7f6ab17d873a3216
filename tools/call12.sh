#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
(nvidia-smi topo -m; echo; numactl -H 2>&1 | head -20; for n in /sys/devices/system/node/node*; do echo $n $(cat $n/cpulist); done; cat /proc/self/status | grep -i "allowed_list") > $OUT/c12_topology.log 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { tag=$1; shift
 env $ENVS $TR --master-port $PORT bench.py --gpus 2 --config C3 --reads 12000000 --steps 3 --warmup 2 --e2e-steps 2 --parity off > $OUT/c12_$tag.json 2> $OUT/c12_$tag.err
 python - $OUT/c12_$tag.json "$tag $ENVS" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2],'|',round(d['value']/1e9,2),'G/s',round(d['ms_per_step'],2),'ms e2e',round(d['e2e']['ms_per_step'],1))
    for k,v in d['per_k'].items(): print('   k',k,v['ms'])
    print('   exch',{k:(round(v['gbs_per_dir'],1),v['ms_comm'],v['ms_not_hidden']) for k,v in d['exchange'].items()})
except Exception as e: print(sys.argv[2],'FAILED',e)
PY
}
PORT=29531 ENVS="KC_X=0" run base
PORT=29532 ENVS="NCCL_MIN_P2P_NCHANNELS=32 NCCL_MAX_P2P_NCHANNELS=32" run ch32
PORT=29533 ENVS="NCCL_MIN_P2P_NCHANNELS=8 NCCL_MAX_P2P_NCHANNELS=8" run ch8
grep -i "channel\|P2P" $OUT/c12_base.err | head -5

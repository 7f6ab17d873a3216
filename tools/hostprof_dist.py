"""torchrun worker: wall-clock phases of one kcount round per rank for several parse batch sizes (exchange overlap study)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import bench
from mhm2_proxy_b200 import kcount
from mhm2_proxy_b200.dist import share_unique_id

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
uid = share_unique_id(rank, kcount.KmerCounter.comm_unique_id)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000000
tb, to = bench.gen_reads(n, seed=1 + rank)
d_b = tb.cuda(); d_o = to.cuda(); torch.cuda.synchronize()
for batch in (0, 1 << 20, 1 << 22, 1 << 24):
    for rep in range(3):
        for k in (21, 33):
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            kc = kcount.KmerCounter(k, num_shards=world, shard_id=rank, device=local, expected_instances=n * (150 - k - 1) * world,
                                    batch_reads=batch)
            kc.comm_init(uid)
            t1 = time.perf_counter()
            kc.add_reads_device(d_b.data_ptr(), d_o.data_ptr(), n, d_b.numel(), 150)
            t2 = time.perf_counter()
            kc.finish(); t3 = time.perf_counter()
            st = kc.stats()
            kc.close(); t4 = time.perf_counter()
            if rep == 2:
                print(f"rank {rank} batch {batch} k={k} create {1e3*(t1-t0):.1f} add {1e3*(t2-t1):.1f} (kern {st['ms_parse']:.1f}: count {st['ms_parse_count']:.1f} "
                      f"fill {st['ms_parse_fill']:.1f}) finish {1e3*(t3-t2):.1f} (exch {st['ms_exchange']:.1f} kern {st['ms_finish']:.1f}, count_kernel "
                      f"{st['ms_insert']:.1f}) destroy {1e3*(t4-t3):.1f} total {1e3*(t4-t0):.1f} ms", flush=True)
dist.destroy_process_group()

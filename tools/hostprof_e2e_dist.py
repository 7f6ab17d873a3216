"""torchrun worker (N ranks, one per GPU): where does the end-to-end arm's time go as N grows?
  1. the box's copy ceiling: every rank copies a pinned buffer host->device, device->host and both at once, all ranks at the
     same time (plain cudaMemcpyAsync through torch) -> GB/s per rank and in aggregate;
  2. wall-clock phases of the e2e step per rank (kc_reads_create, kc_add_reads_resident, kc_finish_to_host per k).
usage: python -m torch.distributed.run --nproc-per-node N tools/hostprof_e2e_dist.py [reads per GPU] [copy MB]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import bench
from mhm2_proxy_b200 import kcount
from mhm2_proxy_b200.dist import share_unique_id

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else bench.WORKLOADS["C2"]["reads"]
mb = int(sys.argv[2]) if len(sys.argv) > 2 else 1024


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def allsum(v):
    if world == 1:
        return v
    t = torch.tensor([v], dtype=torch.float64, device="cuda")
    dist.all_reduce(t)
    return float(t.item())


# ---- 1. copy ceiling ----
hb = torch.empty(mb << 20, dtype=torch.uint8, pin_memory=True)
hb2 = torch.empty(mb << 20, dtype=torch.uint8, pin_memory=True)
db = torch.empty(mb << 20, dtype=torch.uint8, device="cuda")
db2 = torch.empty(mb << 20, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
try:
    cpus = sorted(os.sched_getaffinity(0))
    aff = f"{cpus[0]}-{cpus[-1]} ({len(cpus)} cpus)"
except Exception:
    aff = "?"
for what in ("h2d", "d2h", "both"):
    for rep in range(3):
        barrier()
        t0 = time.perf_counter()
        if what in ("h2d", "both"):
            with torch.cuda.stream(s1):
                db.copy_(hb, non_blocking=True)
        if what in ("d2h", "both"):
            with torch.cuda.stream(s2):
                hb2.copy_(db2, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    gbs = (2 if what == "both" else 1) * (mb << 20) / dt / 1e9
    tot = allsum(gbs)
    print(f"[copy] rank {rank} (affinity {aff}) {what}: {gbs:.1f} GB/s" + (f" | all {world} ranks at once: {tot:.1f} GB/s" if rank == 0 else ""),
          flush=True)
del hb, hb2, db, db2

# ---- 2. phases of the e2e step ----
uid = share_unique_id(rank, kcount.KmerCounter.comm_unique_id) if world > 1 else None
wl = bench.Workload("C2", world, rank, reads=n)
hout = bench.HostOut(wl.n_reads * 6, 2)
for rep in range(4):
    barrier()
    t0 = time.perf_counter()
    res = kcount.ResidentReads(wl.tb.data_ptr(), wl.to.data_ptr(), wl.n_reads, device=local)
    t1 = time.perf_counter()
    line = f"[e2e] rank {rank} rep {rep}: reads_create {1e3 * (t1 - t0):.1f}"
    for k in wl.ks:
        ta = time.perf_counter()
        kc = kcount.KmerCounter(k, num_shards=world, shard_id=rank, device=local, expected_instances=wl.n_reads * (150 - k - 1))
        if world > 1:
            kc.comm_init(uid)
        tb_ = time.perf_counter()
        kc.add_reads_resident(res)
        tc = time.perf_counter()
        nl = k // 32 + 1
        ns = kc.finish_into(hout.keys.data_ptr(), hout.counts.data_ptr(), hout.left.data_ptr(), hout.right.data_ptr(),
                            min(hout.cap, hout.keys.numel() // nl))
        td = time.perf_counter()
        st = kc.stats()
        kc.close()
        te = time.perf_counter()
        line += (f" | k={k}: create {1e3 * (tb_ - ta):.1f} add {1e3 * (tc - tb_):.1f} (parse wall {st['ms_parse']:.1f}) finish_to_host "
                 f"{1e3 * (td - tc):.1f} (exch wait {st['ms_exchange']:.1f}, count {st['ms_insert']:.1f}, d2h {ns * (8 * nl + 4) / 1e6:.0f} MB) "
                 f"destroy {1e3 * (te - td):.1f}")
    res.close()
    t2 = time.perf_counter()
    if rep >= 2:
        print(line + f" | step {1e3 * (t2 - t0):.1f} ms", flush=True)
if world > 1:
    dist.destroy_process_group()

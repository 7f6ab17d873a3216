import sys, time, os
sys.path.insert(0, '/root/repo')
import torch, numpy as np
import bench
from mhm2_proxy_b200 import kcount
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000000
tb, to = bench.gen_reads(n, seed=1)
d_b = tb.cuda(); d_o = to.cuda(); torch.cuda.synchronize()
for rep in range(3):
    for k in (21, 33):
        t0 = time.perf_counter()
        kc = kcount.KmerCounter(k, expected_instances=n * (150 - k - 1))
        t1 = time.perf_counter()
        kc.add_reads_device(d_b.data_ptr(), d_o.data_ptr(), n, d_b.numel(), 150)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        kc.finish(); torch.cuda.synchronize(); t3 = time.perf_counter()
        st = kc.stats()
        kc.close(); t4 = time.perf_counter()
        print(f"k={k} create {1e3*(t1-t0):.1f} add {1e3*(t2-t1):.1f} (kern {st['ms_parse']:.1f}: count {st['ms_parse_count']:.1f} fill {st['ms_parse_fill']:.1f}) finish {1e3*(t3-t2):.1f} (kern {st['ms_finish']:.1f}, count_kernel {st['ms_insert']:.1f}) destroy {1e3*(t4-t3):.1f}  total {1e3*(t4-t0):.1f} ms  -> {st['instances_parsed']/(t4-t0)/1e9:.2f} G/s", flush=True)

# end-to-end arm: pinned host reads in, survivors out to pinned host buffers
hout = bench.HostOut(n)
for rep in range(2):
    for k in (21, 33):
        t0 = time.perf_counter()
        kc = kcount.KmerCounter(k, expected_instances=n * (150 - k - 1))
        kc.add_reads_ptr(tb.data_ptr(), to.data_ptr(), n)
        t1 = time.perf_counter()
        kc.finish(); t2 = time.perf_counter()
        ns = kc.survivors_into(hout.keys.data_ptr(), hout.counts.data_ptr(), hout.left.data_ptr(), hout.right.data_ptr())
        t3 = time.perf_counter()
        st = kc.stats(); kc.close()
        print(f"e2e k={k} add(h2d+parse) {1e3*(t1-t0):.1f} (h2d {st['ms_h2d']:.1f}, parse kernels {st['ms_parse']:.1f}) finish {1e3*(t2-t1):.1f} d2h {1e3*(t3-t2):.1f} ({ns} survivors) total {1e3*(t3-t0):.1f} ms -> {st['instances_parsed']/(t3-t0)/1e9:.2f} G/s", flush=True)

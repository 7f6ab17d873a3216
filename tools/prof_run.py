"""One small kcount round (for ncu): python tools/prof_run.py [reads] [k,k,...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from mhm2_proxy_b200 import kcount
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000000
ks = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [21]
tb, to = bench.gen_reads(n, seed=1)
d_b = tb.cuda(); d_o = to.cuda(); torch.cuda.synchronize()
for k in ks:
    kc = kcount.KmerCounter(k, expected_instances=n * (150 - k - 1))
    kc.add_reads_device(d_b.data_ptr(), d_o.data_ptr(), n, d_b.numel(), 150)
    kc.finish()
    st = kc.stats()
    print(k, {x: st[x] for x in ("instances_parsed", "distinct", "survivors", "supermers_out", "num_partitions", "ms_parse_count", "ms_parse_fill", "ms_insert")})
    kc.close()

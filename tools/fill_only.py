"""Parse + fill only (no count stage): python tools/fill_only.py [reads] [k,k,..]  — for fill-kernel timing experiments"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from mhm2_proxy_b200 import kcount
n = int(sys.argv[1]) if len(sys.argv) > 1 else 13333334
ks = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [21, 33]
tb, to = bench.gen_reads(n, seed=1)
d_b = tb.cuda(); d_o = to.cuda(); torch.cuda.synchronize()
for rep in range(3):
    for k in ks:
        kc = kcount.KmerCounter(k, expected_instances=n * (150 - k - 1))
        kc.add_reads_device(d_b.data_ptr(), d_o.data_ptr(), n, d_b.numel(), 150)
        st = kc.stats()
        print(rep, k, {x: round(st[x], 2) for x in ("ms_parse_count", "ms_parse_fill", "ms_parse")}, st["supermers_out"], flush=True)
        kc.close()

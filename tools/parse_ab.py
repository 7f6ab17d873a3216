"""A/B of library builds (KCOUNT_B200_LIB=...): parse + count kernel times at bench size, device-resident."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from mhm2_proxy_b200 import kcount
n = int(sys.argv[1]) if len(sys.argv) > 1 else 13333334
tb, to = bench.gen_reads(n, seed=1)
d_b = tb.cuda(); d_o = to.cuda(); torch.cuda.synchronize()
for k in (21, 33):
    for rep in range(3):
        kc = kcount.KmerCounter(k, expected_instances=n * (150 - k - 1))
        kc.add_reads_device(d_b.data_ptr(), d_o.data_ptr(), n, d_b.numel(), 150)
        kc.finish()
        st = kc.stats()
        kc.close()
    print(f"{os.environ.get('KCOUNT_B200_LIB', 'default')} k={k}: parse_kernel {st['ms_parse_count']:.2f} fill_kernel {st['ms_parse_fill']:.2f} parse {st['ms_parse']:.2f} count_kernel {st['ms_insert']:.2f} finish {st['ms_finish']:.2f} items {st['num_partitions']} survivors {st['survivors']}", flush=True)

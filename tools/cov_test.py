"""Does the count stage's cost per instance depend on coverage (more reads over the same community)?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from mhm2_proxy_b200 import kcount
for n in (13333334, 26666668):
    tb, to = bench.gen_reads(n, seed=1)
    d_b = tb.cuda(); d_o = to.cuda(); torch.cuda.synchronize()
    for k in (21, 33):
        for rep in range(2):
            kc = kcount.KmerCounter(k, expected_instances=n * (150 - k - 1))
            kc.add_reads_device(d_b.data_ptr(), d_o.data_ptr(), n, d_b.numel(), 150)
            kc.finish()
            st = kc.stats()
            kc.close()
        print(f"reads {n} k={k}: P {st['num_partitions']} instances {st['instances_parsed']} distinct {st['distinct']} survivors {st['survivors']} "
              f"count_kernel {st['ms_insert']:.2f} ms = {st['ms_insert'] * 1e6 / st['instances_parsed']:.4f} ns/instance, parse {st['ms_parse']:.2f} ms, retries {st['table_retries']}",
              flush=True)
    del d_b, d_o, tb, to
    torch.cuda.empty_cache()

#!/usr/bin/env python
"""bench.py — kcount k-mers inserted/sec on B200 (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W]          our arm (CUDA library through the C ABI)
  python bench.py --impl reference [...]                       the reference's own CPU kcount on the host cores

A step = one pass of the hot path (analyze_kmers: parse -> [exchange] -> insert -> filter) over the synthetic read set,
once per k of the workload (the reference runs one kcount round per k, main.cpp:169). At N=1 the workload is
BASELINE.json configs[1]: ~2 Gbase of synthetic 150 bp paired reads, k=21 and k=33, on 1xB200.
`value`: inputs already resident in HBM (kc_add_reads_device). `e2e`: HOST (pinned) buffers through kc_add_reads, H2D copies
and the D2H read of the surviving table inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "kcount k-mers inserted/sec"
UNIT = "kmers/s"
READ_LEN = 150
C2_READS = 13_333_334  # ~2 Gbase of 150 bp reads (6.67 M pairs)
C2_KS = (21, 33)
GENOME_BP = 60_000_000  # ~33x mean depth
N_GENOMES = 40
ERR = 0.005


def b_alg(k, L=READ_LEN):
    """Algorithmic bytes per inserted instance for parse+insert: read the packed read once, read the key,
    read-modify-write three u16 counters."""
    W = 8 * (k // 32 + 1)
    return L / (L - k - 1) + W + 12


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def gen_reads(n_reads, seed, pinned=True):
    """Synthetic PackedRead bytes + offsets; host memory is pinned (torch) so that H2D copies run at full PCIe speed."""
    import torch
    from mhm2_proxy_b200 import synth
    com = synth.make_community(GENOME_BP, n_genomes=N_GENOMES, sigma=1.0, seed=0xB200)
    total = n_reads * READ_LEN
    pin = pinned and torch.cuda.is_available()
    tb = torch.empty(total, dtype=torch.uint8, pin_memory=pin)
    rb, ro = synth.make_reads(n_reads, com, read_len=READ_LEN, error_rate=ERR, seed=seed, out=tb.numpy())
    to = torch.empty(n_reads + 1, dtype=torch.int64, pin_memory=pin)
    to.numpy()[:] = ro.astype(np.int64)
    return tb, to


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.proc, self.lines = gpu, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhm": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def reference_rate(tb, to, n_sample, ks, cores):
    """The reference's own CPU kcount (oracle/_ref: kcount_cpu.cpp + kmer_dht.cpp + kmer.cpp + hash_funcs.c compiled
    unmodified) on the first n_sample reads, one emulated UPC++ rank per core, GASNet replaced by shared-memory files.
    Falls back to the C restatement (kind 'port', 1 core) where the reference build is absent."""
    from oracle import harness
    rb = tb.numpy()[: n_sample * READ_LEN]
    ro = to.numpy()[: n_sample + 1].astype(np.uint64)
    inst = secs = 0.0
    if harness.have_ref():
        kind = "reference"
        for k in ks:
            r = harness.run_multirank(k, rb, ro, cores)
            inst += r["instances"]
            secs += r["seconds"]
    else:
        kind, cores = "port", 1
        for k in ks:
            t0 = time.time()
            r = harness.run("port", k, rb, ro)
            secs += time.time() - t0
            inst += r.stats["instances"]
    return {"value": inst / secs, "unit": UNIT, "cores": cores, "kind": kind, "seconds": secs,
            "sample": f"first {n_sample} reads of the workload ({int(inst)} instances over k={list(ks)}); "
                      f"{cores} emulated UPC++ rank(s), one per core, exchange through /dev/shm files"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_sample = args.ref_reads or min(args.reads, 40_000 * cores)
    tb, to = gen_reads(n_sample, seed=0xB200_0002, pinned=False)
    for _ in range(args.warmup if args.ref_warm else 0):
        reference_rate(tb, to, n_sample, args.ks, cores)
    t0 = time.time()
    rates = [reference_rate(tb, to, n_sample, args.ks, cores) for _ in range(args.steps)]
    wall = time.time() - t0
    inst = sum(r["value"] * r["seconds"] for r in rates)
    secs = sum(r["seconds"] for r in rates)
    v = inst / secs
    cb = dict(rates[-1]); cb["value"] = v
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup if args.ref_warm else 0, "ms_per_step": 1000 * secs / args.steps, "wall_s": wall,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(args, n_sample, 1),
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(args, reads_per_gpu, n_gpus):
    return {"workload": f"C2: ~{reads_per_gpu * READ_LEN / 1e9:.2f} Gbase synthetic 150bp paired reads per GPU, k={list(args.ks)}, "
                        f"kcount on {n_gpus}xB200", "reads_per_gpu": reads_per_gpu, "read_len": READ_LEN, "k": list(args.ks),
            "genome_bp": GENOME_BP, "n_genomes": N_GENOMES, "error_rate": ERR, "dmin_thres": 2,
            "l2_flush": "inputs (>= 2 GB) and supermer buffers are far larger than the 126 MB L2; nothing is reused across steps",
            "parallelism": f"reads split across {n_gpus} GPU(s), table hash-sharded by quick_hash(minimizer) % {n_gpus}"}

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
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "25",
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
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup if args.ref_warm else 0, "ms_per_step": 1000 * secs / args.steps, "wall_s": wall,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(args, args.reads, args.gpus),
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    })


def workload_config(args, reads_per_gpu, n_gpus):
    tag = "C2" if (reads_per_gpu == C2_READS and tuple(args.ks) == tuple(C2_KS)) else \
        "C3 (one GPU's eighth of 50 Gbase)" if (tuple(args.ks) == (31,) and reads_per_gpu >= 41_000_000) else "custom"
    return {"workload": f"{tag}: ~{reads_per_gpu * READ_LEN / 1e9:.2f} Gbase synthetic 150bp paired reads per GPU, k={list(args.ks)}, "
                        f"kcount on {n_gpus}xB200", "reads_per_gpu": reads_per_gpu, "read_len": READ_LEN, "k": list(args.ks),
            "genome_bp": GENOME_BP, "n_genomes": N_GENOMES, "error_rate": ERR, "dmin_thres": 2,
            "l2_flush": "inputs (>= 2 GB) and supermer buffers are far larger than the 126 MB L2; nothing is reused across steps",
            "parallelism": f"reads split across {n_gpus} GPU(s), table hash-sharded by quick_hash(minimizer) % {n_gpus}"}


class HostOut:
    """Pinned host buffers that receive the surviving table in the end-to-end arm (allocated once, like an application would)."""

    def __init__(self, n_reads):
        import torch
        self.cap = max(1 << 20, n_reads * 6)  # survivors are bounded by the distinct genomic k-mers, far below this
        self.keys = torch.empty(self.cap * 2, dtype=torch.int64, pin_memory=True)
        self.counts = torch.empty(self.cap, dtype=torch.int16, pin_memory=True)
        self.left = torch.empty(self.cap, dtype=torch.uint8, pin_memory=True)
        self.right = torch.empty(self.cap, dtype=torch.uint8, pin_memory=True)


def one_pass(kcount, k, args, rank, world, dev, mode, tb, to, d_b, d_o, uid, hout=None, resident=None):
    """One kcount round at one k. mode 'device': reads already in HBM; 'host': the step's reads come from pinned host buffers
    (uploaded once per step by kc_reads_create, shared by the k rounds as the reference shares its PackedReads) and the table
    goes back to pinned host memory."""
    n = to.numel() - 1
    inst_hint = n * (READ_LEN - k - 1)
    kc = kcount.KmerCounter(k, num_shards=world, shard_id=rank, device=dev, expected_instances=inst_hint,
                            num_partitions=args.partitions)
    try:
        if world > 1:
            kc.comm_init(uid)
        if mode == "device":
            kc.add_reads_device(d_b.data_ptr(), d_o.data_ptr(), n, d_b.numel(), READ_LEN)
        elif resident is not None:
            kc.add_reads_resident(resident)
        else:
            kc.add_reads_ptr(tb.data_ptr(), to.data_ptr(), n)
        d2h = 0
        if mode == "host":
            # the call a drop-in user makes: finish and receive the table in host memory (kc_finish_to_host)
            nl = k // 32 + 1
            cap = min(hout.cap, hout.keys.numel() // nl)
            ns = kc.finish_into(hout.keys.data_ptr(), hout.counts.data_ptr(), hout.left.data_ptr(), hout.right.data_ptr(), cap)
            d2h = ns * (8 * nl + 4)
        else:
            kc.finish()
        st = kc.stats()
        st["d2h_bytes"] = d2h
        return st
    finally:
        kc.close()


def run_ours(args):
    import torch
    import torch.distributed as dist
    from mhm2_proxy_b200 import kcount
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the kcount path is CUDA only (use --impl reference for the CPU arm)")
    kcount.load_library()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = None
    if world > 1:
        from mhm2_proxy_b200.dist import share_unique_id
        uid = share_unique_id(rank, kcount.KmerCounter.comm_unique_id)

    n_reads = args.reads
    t0 = time.time()
    tb, to = gen_reads(n_reads, seed=0xB200_0002 + 7919 * rank)
    t_gen = time.time() - t0
    d_b = tb.cuda(non_blocking=True)
    d_o = to.cuda(non_blocking=True)
    torch.cuda.synchronize()

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    hout = HostOut(n_reads)

    def one_step(mode):
        """All k rounds over one set of reads. End-to-end mode: host -> device copy of the reads (once; HBM keeps them for
        every k), every round's table back to the host."""
        res = None
        if mode == "host" and not args.e2e_per_round_upload:
            res = kcount.ResidentReads(tb.data_ptr(), to.data_ptr(), to.numel() - 1, device=local)
        try:
            return [one_pass(kcount, k, args, rank, world, local, mode, tb, to, d_b, d_o, uid, hout, res) for k in args.ks]
        finally:
            if res is not None:
                res.close()

    def timed(mode, steps, warmup, sampler=None):
        for _ in range(warmup):
            one_step(mode)
        sync()
        if sampler:
            sampler.start()
        stats = []
        # CUDA events on torch's stream bracket the steps: every library call returns only after its own streams have
        # drained (the ABI is host-synchronous), so the first event sits on an idle device and the second is recorded
        # after the last kernel and copy of the last step have completed; the wall clock is kept as a cross-check
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_start = time.perf_counter()
        ev0.record()
        for _ in range(steps):
            stats.extend(one_step(mode))
        ev1.record()
        sync()
        wall = time.perf_counter() - t_start
        el = ev0.elapsed_time(ev1) * 1e-3
        timed.wall[mode] = wall
        clocks = sampler.stop() if sampler else None
        if world > 1:
            t = torch.tensor([el], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            el = float(t.item())
        return el, stats, clocks

    timed.wall = {}
    sampler = ClockSampler(local) if rank == 0 else None
    el, stats, clocks = timed("device", args.steps, args.warmup, sampler)
    inst_local = sum(s["instances_parsed"] for s in stats)
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    el2, stats2, _ = timed("host", e2e_steps, 1)
    inst2_local = sum(s["instances_parsed"] for s in stats2)
    if world > 1:
        t = torch.tensor([inst_local, inst2_local], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        inst_all, inst2_all = float(t[0].item()), float(t[1].item())
    else:
        inst_all, inst2_all = float(inst_local), float(inst2_local)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # per-kernel device times (CUDA events on the library's stream, summed over the timed steps of rank 0)
    kern = {}
    for name, key in (("parse_kernel", "ms_parse_count"), ("fill_kernel", "ms_parse_fill"), ("count_kernel", "ms_insert")):
        kern[name] = sum(s[key] for s in stats)
    launches = {"parse_kernel": sum(s["n_parse_launches"] for s in stats) // 2,
                "fill_kernel": sum(s["n_parse_launches"] for s in stats) // 2,
                "count_kernel": sum(s["n_insert_launches"] for s in stats)}
    dom = max((n for n in kern if n in launches), key=lambda n: kern[n])
    peak, peak_src = hbm_peak()
    per_k = {}
    alg_bytes = 0.0
    for k in args.ks:
        sk = [s for s, kk in zip(stats, list(args.ks) * args.steps) if kk == k]
        ik = sum(s["instances_parsed"] for s in sk)
        alg_bytes += ik * b_alg(k)
        per_k[str(k)] = {"instances_per_pass": sk[0]["instances_parsed"], "distinct": sk[0]["distinct"], "survivors": sk[0]["survivors"],
                         "partitions": sk[0]["num_partitions"], "supermers": sk[0]["supermers_out"],
                         "instances_counted_rank0": sk[0]["instances_inserted"], "supermers_counted_rank0": sk[0]["supermers_in"],
                         "max_probe": sk[0]["max_probe"],
                         "supermer_bytes": sk[0]["supermer_bytes_out"], "table_retries": sk[0]["table_retries"],
                         "table_group_bytes_max": sk[0]["table_bytes_max"],
                         "ms": {kk: round(sum(s[kk] for s in sk) / len(sk), 3) for kk in
                                ("ms_parse_count", "ms_parse_fill", "ms_parse", "ms_exchange", "ms_insert", "ms_finish")},
                         "b_alg": round(b_alg(k), 2)}
    achieved = alg_bytes / (kern[dom] * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(dom)
        except Exception:
            traffic = None
    cpu = None
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        cpu = reference_rate(tb, to, min(n_reads, args.ref_reads or 40_000 * cores), args.ks, cores)
    value = inst_all / el
    h2d = (tb.numel() + to.numel() * 8) * (len(args.ks) if args.e2e_per_round_upload else 1)
    d2h = sum(s["d2h_bytes"] for s in stats2) / e2e_steps
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000 * el / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic", "config": workload_config(args, n_reads, world),
        "timing": "CUDA events around the K steps (torch.cuda.synchronize + barrier on both sides; the library calls are "
                  "host-synchronous), max over ranks; per-kernel times are CUDA events on the library's stream",
        "wall_ms_per_step": round(1000 * timed.wall["device"] / args.steps, 3),
        "e2e": {"value": inst2_all / el2, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(d2h),
                "steps": e2e_steps, "ms_per_step": 1000 * el2 / e2e_steps,
                "path": ("kc_add_reads per k round" if args.e2e_per_round_upload else
                         "kc_reads_create once per step (pinned host reads -> HBM, shared by the k rounds)") +
                        " + kc_add_reads_resident + kc_finish_to_host (table -> pinned host memory) per k round"},
        "gpu_launches": int(sum(s["gpu_launches"] for s in stats)),
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "launches": launches[dom],
                     "avg_launch_ms": kern[dom] / max(1, launches[dom]),
                     "algorithmic_bytes": "B_alg(k) = L/(L-k-1) + W + 12 per instance x instances of the timed steps (DESIGN.md)",
                     "pipeline_frac": (alg_bytes / el / 1e9) / peak},
        "kernel_ms": {k_: round(v, 3) for k_, v in kern.items()}, "per_k": per_k,
        "cpu_baseline": cpu, "clocks": clocks, "gen_s": round(t_gen, 1),
    }
    emit(out)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, line)
    else:
        sys.stdout.write(line.decode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reads", type=int, default=C2_READS, help="reads per GPU")
    ap.add_argument("--ks", type=lambda s: tuple(int(x) for x in s.split(",")), default=C2_KS)
    ap.add_argument("--partitions", type=int, default=0)
    ap.add_argument("--e2e-steps", dest="e2e_steps", type=int, default=2)
    ap.add_argument("--e2e-per-round-upload", dest="e2e_per_round_upload", action="store_true",
                    help="end-to-end arm: copy the reads host->device again for every k round (kc_add_reads) instead of once per step")
    ap.add_argument("--ref-reads", dest="ref_reads", type=int, default=0, help="reads in the CPU sample (default 40k per core)")
    ap.add_argument("--ref-warm", dest="ref_warm", action="store_true", help="also run warm-up passes in --impl reference")
    ap.add_argument("--no-cpu", dest="no_cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    # libraries (NCCL banner, torchrun notices) may write to fd 1: keep stdout for the ONE JSON line
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

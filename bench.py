#!/usr/bin/env python
"""bench.py — kcount k-mers inserted/sec on B200 (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C1..C5]     our arm (CUDA library through the C ABI)
  python bench.py --impl reference [...]                                      the reference's own CPU kcount on the host cores

A step = one pass of the hot path (analyze_kmers: parse -> [exchange] -> insert -> filter) over the synthetic read set,
once per k of the workload (the reference runs one kcount round per k, main.cpp:169). Workloads = BASELINE.json configs
(generators: SURVEY.md §8d):
  C1  arcticsynth-shaped, 2 M reads, k=31            C2  ~2 Gbase per GPU, k=21 and k=33   (default for --gpus 1, 2, 4)
  C3  50 Gbase over 8 GPUs = 41.67 M reads per GPU, k=31, community scaled with N          (default for --gpus 8)
  C4  C2 reads at k=99 + previous-round contigs     C5  skewed community (Zipf, repeat, poly-A), k=31
`value`: reads already resident in HBM (kc_add_reads_device; excludes the H2D copy). `e2e`: HOST (pinned) buffers, H2D copies
and the D2H read of the surviving table inside the timed region. `parity`: the table digest (kc_table_digest) of the very
workload that was timed against the same digest of the reference's own kcount_cpu.cpp (oracle/_ref), outside the timed
region: at N=1 over the whole read set (or a stated prefix when the host lacks the memory), at N>1 over a fixed read set
split N ways and exchanged over NCCL, per owner shard.
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

# per-GPU sizes; C3's community grows with the number of GPUs so that every shard holds the same share of it (weak scaling)
WORKLOADS = {
    "C1": dict(reads=2_000_000, ks=(31,), genome_bp=100_000_000, n_genomes=32, err=0.01, seed=0xB200_0001,
               what="arcticsynth-shaped: 32 genomes / 100 Mbp, log-normal abundances, 1 % error"),
    "C2": dict(reads=13_333_334, ks=(21, 33), genome_bp=60_000_000, n_genomes=40, err=0.005, seed=0xB200_0002,
               what="~2 Gbase per GPU from a 60 Mbp community (~33x), 0.5 % error"),
    "C3": dict(reads=41_666_667, ks=(31,), genome_bp=187_500_000, n_genomes=37.5, err=0.01, seed=0xB200_0003, scale_genome=True,
               what="50 Gbase / 1.5 Gbp / 300 genomes over 8 GPUs = 6.25 Gbase and 187.5 Mbp of community per GPU, 1 % error"),
    "C4": dict(reads=13_333_334, ks=(99,), genome_bp=60_000_000, n_genomes=40, err=0.005, seed=0xB200_0004, contigs=True,
               what="C2-style reads at k=99 (Kmer<128>) + previous-round contigs cut from the genomes (depths 2..200, duplicates, "
                    "extension-conflicting copies)"),
    "C5": dict(reads=13_333_334, ks=(31,), genome_bp=60_000_000, n_genomes=1000, err=0.01, seed=0xB200_0005, zipf=1.2, skew=True,
               what="Zipf(1.2) abundances over 1000 genomes + 0.1 % reads from a 2 kbp repeat + poly-A reads (saturating counters)"),
}


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


def workload_config(name, n_reads, ks, n_gpus, genome_bp, n_genomes, ctg=None):
    w = WORKLOADS[name]
    c = {"workload": f"{name}: {w['what']}; {n_reads} synthetic {READ_LEN}bp paired reads "
                     f"(~{n_reads * READ_LEN / 1e9:.2f} Gbase) per GPU, k={list(ks)}, kcount on {n_gpus}xB200",
         "reads_per_gpu": n_reads, "read_len": READ_LEN, "k": list(ks), "genome_bp": genome_bp,
         "n_genomes": n_genomes, "error_rate": w["err"], "dmin_thres": 2,
         "l2_flush": "inputs (>= 0.3 GB) and supermer buffers are far larger than the 126 MB L2; nothing is reused across steps",
         "parallelism": f"reads split across {n_gpus} GPU(s), table hash-sharded by quick_hash(minimizer) % {n_gpus}"}
    if ctg is not None:
        c["contigs_per_gpu"] = len(ctg[2])
        c["contig_bases_per_gpu"] = int(ctg[1][-1])
    return c


def community_size(name, n_gpus):
    w = WORKLOADS[name]
    scale = n_gpus if w.get("scale_genome") else 1
    return int(w["genome_bp"] * scale), max(1, int(round(w["n_genomes"] * scale)))


class Workload:
    """One rank's synthetic input: PackedRead bytes + offsets in (pinned) host memory, optionally contigs."""

    def __init__(self, name, n_gpus, rank, reads=None, ks=None, pinned=True):
        import torch
        from mhm2_proxy_b200 import synth
        w = dict(WORKLOADS[name])
        self.name, self.w = name, w
        self.n_reads = int(reads or w["reads"])
        self.ks = tuple(ks or w["ks"])
        self.genome_bp, self.n_genomes = community_size(name, n_gpus)
        t0 = time.time()
        self.com = synth.make_community(self.genome_bp, n_genomes=self.n_genomes, sigma=1.0, seed=w["seed"] & 0xFFFF,
                                        zipf=w.get("zipf"))
        pin = pinned and torch.cuda.is_available()
        self.tb = torch.empty(self.n_reads * READ_LEN, dtype=torch.uint8, pin_memory=pin)
        seed = w["seed"] + 7919 * rank
        if w.get("skew"):
            _, ro = synth.make_skewed_reads(self.n_reads, self.com, read_len=READ_LEN, error_rate=w["err"], seed=seed, out=self.tb.numpy())
        else:
            _, ro = synth.make_reads(self.n_reads, self.com, read_len=READ_LEN, error_rate=w["err"], seed=seed, out=self.tb.numpy())
        self.to = torch.empty(self.n_reads + 1, dtype=torch.int64, pin_memory=pin)
        self.to.numpy()[:] = ro.astype(np.int64)
        self.ctg = None
        if w.get("contigs"):  # every rank cuts the same contigs and takes its slice by index
            cs, co, cd = synth.make_contig_tiling(self.com, self.ks[0], seed=w["seed"] & 0xFFFF)
            c0, c1 = len(cd) * rank // n_gpus, len(cd) * (rank + 1) // n_gpus
            self.ctg = (np.ascontiguousarray(cs[int(co[c0]):int(co[c1])]), (co[c0:c1 + 1] - co[c0]).astype(np.uint64), cd[c0:c1].copy())
        self.gen_s = time.time() - t0

    def ctg_instances(self, k):
        if self.ctg is None:
            return 0
        return int(np.maximum(np.diff(self.ctg[1].astype(np.int64)) - k - 1, 0).sum())

    def config(self, n_gpus):
        return workload_config(self.name, self.n_reads, self.ks, n_gpus, self.genome_bp, self.n_genomes, self.ctg)


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
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
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


def reference_rate(wl, n_sample, ks, cores):
    """The reference's own CPU kcount (oracle/_ref: kcount_cpu.cpp + kmer_dht.cpp + kmer.cpp + hash_funcs.c compiled
    unmodified) on the first n_sample reads, one emulated UPC++ rank per core, GASNet replaced by shared-memory files.
    Falls back to the C restatement (kind 'port', 1 core) where the reference build is absent."""
    from oracle import harness
    rb = wl.tb.numpy()[: n_sample * READ_LEN]
    ro = wl.to.numpy()[: n_sample + 1].astype(np.uint64)
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


def default_config(n_gpus):
    return "C3" if n_gpus >= 8 else "C2"


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    name = args.config or default_config(args.gpus)
    n_reads = args.reads or WORKLOADS[name]["reads"]
    n_sample = args.ref_reads or min(n_reads, 40_000 * cores)
    wl = Workload(name, args.gpus, 0, reads=n_sample, ks=args.ks, pinned=False)
    for _ in range(args.warmup if args.ref_warm else 0):
        reference_rate(wl, n_sample, wl.ks, cores)
    t0 = time.time()
    rates = [reference_rate(wl, n_sample, wl.ks, cores) for _ in range(args.steps)]
    wall = time.time() - t0
    inst = sum(r["value"] * r["seconds"] for r in rates)
    secs = sum(r["seconds"] for r in rates)
    v = inst / secs
    cb = dict(rates[-1]); cb["value"] = v
    # the config line describes the full workload of our arm; `sample` says what was timed
    gbp, ng = community_size(name, args.gpus)
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup if args.ref_warm else 0, "ms_per_step": 1000 * secs / args.steps, "wall_s": wall,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(name, n_reads, wl.ks, args.gpus, gbp, ng),
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    })


class HostOut:
    """Pinned host buffers that receive the surviving table in the end-to-end arm (allocated once, like an application would)."""

    def __init__(self, rows, n_longs_max):
        import torch
        self.cap = max(1 << 20, int(rows))
        self.keys = torch.empty(self.cap * n_longs_max, dtype=torch.int64, pin_memory=True)
        self.counts = torch.empty(self.cap, dtype=torch.int16, pin_memory=True)
        self.left = torch.empty(self.cap, dtype=torch.uint8, pin_memory=True)
        self.right = torch.empty(self.cap, dtype=torch.uint8, pin_memory=True)


def one_pass(kcount, k, args, rank, world, dev, mode, wl, d_b, d_o, uid, hout=None, resident=None, want_digest=False, n_reads=None):
    """One kcount round at one k. mode 'device': reads already in HBM; 'host': the step's reads come from pinned host buffers
    (uploaded once per step by kc_reads_create, shared by the k rounds as the reference shares its PackedReads) and the table
    goes back to pinned host memory. Contigs (C4) always come from host memory, as Contigs do in the reference."""
    n = n_reads if n_reads is not None else wl.n_reads
    inst_hint = n * (READ_LEN - k - 1) + wl.ctg_instances(k)
    kc = kcount.KmerCounter(k, num_shards=world, shard_id=rank, device=dev, expected_instances=inst_hint,
                            num_partitions=args.partitions)
    try:
        if world > 1:
            kc.comm_init(uid)
        if mode == "device":
            kc.add_reads_device(d_b.data_ptr(), d_o.data_ptr(), n, n * READ_LEN, READ_LEN)
        elif resident is not None:
            kc.add_reads_resident(resident)
        else:
            kc.add_reads_ptr(wl.tb.data_ptr(), wl.to.data_ptr(), n)
        if wl.ctg is not None:
            if world > 1:
                kc.flush()  # kcount.cpp:148: the reads pass is flushed (collective) before the contig pass
            kc.add_contigs(*wl.ctg)
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
        if want_digest:
            st["digest"] = kc.digest()
        return st
    finally:
        kc.close()


def parity_check(kcount, args, rank, world, local, wl, d_b, d_o, uid, dist):
    """Compare the CUDA path's table digest with the reference's (oracle/_ref, the reference's own sources) on this workload.
    N=1: the whole read set when the host has the memory for the reference's tables, else a stated prefix. N>1: a fixed read
    set (the first reads of every rank's share) exchanged over NCCL, per owner shard. Outside the timed region."""
    import psutil
    from oracle import harness
    out = {"checked": False, "equal": None, "oracle": "reference" if harness.have_ref() else None}
    if not harness.have_ref():
        out["note"] = "oracle/_ref/libkcount_ref.so absent"
        return out
    cores = os.cpu_count() or 1
    if world == 1:
        n = wl.n_reads if not args.parity_reads else min(wl.n_reads, args.parity_reads)
    else:
        n = min(wl.n_reads, (args.parity_reads or 1_000_000) // world)
    per_k = {}
    equal = True
    t_ref = 0.0
    for k in wl.ks:
        inst = n * world * (READ_LEN - k - 1)
        W = 8 * (k // 32 + 1)
        frac = args.parity_distinct_frac
        # memory of the reference's tables: 3 x num_elems slots of W + 20 bytes (kcount_cpu.cpp:431-438)
        need = 3 * frac * inst * (W + 20) + 2.5 * n * world * READ_LEN
        avail = psutil.virtual_memory().available
        if world == 1 and need > 0.6 * avail:
            n_fit = int(n * 0.6 * avail / need)
            out["note"] = (f"host has {avail / 2**30:.0f} GiB available; the reference's tables for all {n} reads need "
                           f"{need / 2**30:.0f} GiB: parity on the first {n_fit} reads")
            n = n_fit
            inst = n * (READ_LEN - k - 1)
        st = one_pass(kcount, k, args, rank, world, local, "device", wl, d_b, d_o, uid, want_digest=True, n_reads=n)
        mine = list(st["digest"]) + [st["instances_parsed"], st["instances_inserted"]]
        rb = wl.tb.numpy()[: n * READ_LEN]
        cs = wl.ctg
        if world > 1:
            # rank 0 needs every rank's digest and the very bytes every rank parsed (a few MB each)
            box = [None] * world
            dist.all_gather_object(box, (mine, np.array(rb), cs))
            rows = [b_[0] for b_ in box]
            if rank != 0:
                continue
            rb = np.concatenate([b_[1] for b_ in box])
            ctgs = [b_[2] for b_ in box if b_[2] is not None]
            cs = None
            if ctgs:
                co = [np.zeros(1, dtype=np.uint64)]
                for c in ctgs:
                    co.append(c[1][1:] + co[-1][-1])
                cs = (np.concatenate([c[0] for c in ctgs]), np.concatenate(co), np.concatenate([c[2] for c in ctgs]))
        else:
            rows = [mine]
        ro = np.arange(n * world + 1, dtype=np.uint64) * np.uint64(READ_LEN)
        P = world * max(1, min(cores, 64) // world)
        kw = {}
        if cs is not None:
            kw = dict(ctg_seqs=cs[0], ctg_offsets=cs[1], ctg_depths=cs[2])
        t0 = time.time()
        ref = harness.run_multirank_stream(k, rb, ro, P, distinct_frac=frac, chunk_reads=args.parity_chunk, **kw)
        t_ref += time.time() - t0
        want = harness.combine_digests(ref["digests"], world)
        got = [tuple(r[:4]) for r in rows]
        ok = (got == [tuple(w_) for w_ in want] and ref["dropped"] == 0 and ref["singleton_overrides"] == 0
              and sum(r[4] for r in rows) == sum(r[5] for r in rows) == ref["instances"])
        equal = equal and ok
        per_k[str(k)] = {"equal": ok, "survivors": sum(g[0] for g in got), "reference_survivors": ref["survivors"],
                         "instances": ref["instances"], "instances_inserted_gpu": sum(r[5] for r in rows),
                         "reference_dropped": ref["dropped"], "reference_singleton_overrides": ref["singleton_overrides"],
                         "reference_ranks": P, "reference_kmers_per_s": ref["instances"] / ref["seconds"],
                         "digest_gpu": [f"{g[2]:016x}/{g[3]:016x}" for g in got], "digest_reference": [f"{w_[2]:016x}/{w_[3]:016x}" for w_ in want]}
    if rank != 0:
        return None
    out.update({"checked": True, "equal": equal, "reads": n * world, "reads_of_workload": wl.n_reads * world,
                "transport": "nccl" if world > 1 else "none (one shard)", "per_k": per_k, "reference_seconds": round(t_ref, 1),
                "digest": "kc_table_digest: n, sum(count), sum and xor over rows of fmix64-chained (key words, count, left, right)"})
    return out


def copy_ceiling(world, dist, mb=256):
    """What the box gives every rank when all ranks copy at once (plain cudaMemcpyAsync from / to pinned memory): host->device,
    device->host, and both directions together, GB/s per rank (the minimum over ranks). The end-to-end arm moves h2d + d2h bytes
    per step through exactly this path; with N ranks on one host the figures fall (shared PCIe switches / host memory)."""
    import torch
    hb = torch.empty(mb << 20, dtype=torch.uint8, pin_memory=True)
    hb2 = torch.empty(mb << 20, dtype=torch.uint8, pin_memory=True)
    db = torch.empty(mb << 20, dtype=torch.uint8, device="cuda")
    db2 = torch.empty(mb << 20, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    out = {}
    for what in ("h2d", "d2h", "both"):
        best = 0.0
        for _ in range(3):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            if what in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    db.copy_(hb, non_blocking=True)
            if what in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    hb2.copy_(db2, non_blocking=True)
            torch.cuda.synchronize()
            best = max(best, (mb << 20) / (time.perf_counter() - t0) / 1e9)  # per direction
        if world > 1:
            t = torch.tensor([best], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            best = float(t.item())
        out[what + "_gbs_per_rank_per_dir"] = round(best, 1)
    return out


def dropin_seam(wl, ks):
    """The real seam: the reference's own analyze_kmers (kcount.cpp, unmodified) over PackedReads with
    integration/kcount_b200.cpp as its backend, timed from 'reads in PackedReads' to 'local_kmers filled'
    (oracle/_ref/libkcount_dropin.so = reference sources + our binding)."""
    import ctypes as C
    from oracle import harness
    if not harness.have_dropin():
        return None
    lib = C.CDLL(harness.DROPIN_SO)
    lib.odrop_time.restype = C.c_double
    lib.odrop_time.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.odrop_seam_seconds.argtypes = [C.c_void_p]
    rb = wl.tb.numpy()
    ro = wl.to.numpy().astype(np.uint64)
    secs, inst, per_k = 0.0, 0, {}
    for k in ks:
        nk = np.zeros(1, dtype=np.uint64)
        t = lib.odrop_time(k, 33, 2, rb.ctypes.data, ro.ctypes.data, wl.n_reads, nk.ctypes.data)
        if t < 0:
            return None
        sp = np.zeros(3)
        lib.odrop_seam_seconds(sp.ctypes.data_as(C.c_void_p))
        secs += t
        inst += wl.n_reads * (READ_LEN - k - 1)
        per_k[str(k)] = {"seconds": round(t, 3), "local_kmers": int(nk[0]),
                         "seconds_reference_count_kmers_loop": round(t - float(sp.sum()), 3),  # PackedRead::unpack + tolower, one thread
                         "seconds_binding_staging": round(float(sp[0]), 3), "seconds_inside_library": round(float(sp[1]), 3),
                         "seconds_filling_local_kmers": round(float(sp[2]), 3)}
    return {"value": inst / secs, "unit": UNIT, "seconds": round(secs, 3), "per_k": per_k,
            "path": "reference analyze_kmers<MAX_K> (count_kmers unpacks every PackedRead to ASCII on one host thread, "
                    "SeqBlockInserter::process_seq stages it, kc_add_reads_ascii, kc_finish, survivors inserted one by one into "
                    "the reference's bytell_hash_map local_kmers)"}


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

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def measure(name, steps, warmup, e2e_steps, e2e_warmup, with_parity, sampler=None):
        wl = Workload(name, world, rank, reads=args.reads, ks=args.ks)
        d_b = wl.tb.cuda(non_blocking=True)
        d_o = wl.to.cuda(non_blocking=True)
        torch.cuda.synchronize()
        hout = None  # sized from the tables the device-resident passes produce (an application knows its genome size)

        def one_step(mode):
            """All k rounds over one set of reads. End-to-end mode: host -> device copy of the reads (once; HBM keeps them for
            every k), every round's table back to the host."""
            res = None
            if mode == "host" and not args.e2e_per_round_upload:
                res = kcount.ResidentReads(wl.tb.data_ptr(), wl.to.data_ptr(), wl.n_reads, device=local)
            try:
                return [one_pass(kcount, k, args, rank, world, local, mode, wl, d_b, d_o, uid, hout, res) for k in wl.ks]
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
            clocks = sampler.stop() if sampler else None
            if world > 1:
                t = torch.tensor([el], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                el = float(t.item())
            return el, stats, clocks, wall

        parity = None
        if with_parity and world > 1:  # before timing: the NCCL data plane must produce the reference's per-owner tables
            parity = parity_check(kcount, args, rank, world, local, wl, d_b, d_o, uid, dist)
        el, stats, clocks, wall = timed("device", steps, warmup, sampler)
        hout = HostOut(1.1 * max(s["survivors"] for s in stats) + (1 << 20), max(k // 32 + 1 for k in wl.ks))
        el2, stats2, _, _ = timed("host", e2e_steps, e2e_warmup)
        # invariants of every timed pass, summed over ranks: nothing parsed may be missing from the tables
        sums = [sum(s["instances_parsed"] for s in stats), sum(s["instances_inserted"] for s in stats),
                sum(s["instances_parsed"] for s in stats2), sum(s["instances_inserted"] for s in stats2),
                sum(s["supermer_bytes_sent"] for s in stats), sum(s["supermer_bytes_recv"] for s in stats)]
        if world > 1:
            t = torch.tensor(sums, dtype=torch.float64, device="cuda")
            dist.all_reduce(t)
            sums = [float(x) for x in t.tolist()]
        if with_parity and world == 1:
            parity = parity_check(kcount, args, rank, world, local, wl, d_b, d_o, uid, dist)
        nks_ = len(wl.ks)
        h2d = (wl.tb.numel() + wl.to.numel() * 8) * (nks_ if args.e2e_per_round_upload else 1)
        if wl.ctg is not None:
            h2d += (len(wl.ctg[0]) + 8 * len(wl.ctg[1]) + 8 * len(wl.ctg[2])) * nks_
        res = dict(wl=wl, el=el, stats=stats, clocks=clocks, wall=wall, el2=el2, stats2=stats2, sums=sums, parity=parity,
                   steps=steps, e2e_steps=e2e_steps, h2d=h2d, config=wl.config(world))
        del d_b, d_o, hout
        return res

    name = args.config or default_config(world)
    ceiling = copy_ceiling(world, dist)
    sampler = ClockSampler(local) if rank == 0 else None
    e2e_steps = max(1, args.e2e_steps)
    m = measure(name, args.steps, args.warmup, e2e_steps, max(3, min(args.warmup, 3)), args.parity != "off", sampler)
    second = None
    if (world >= 8 and name == "C3" and not args.config and not args.no_second) or (args.second and name != "C2"):
        # the C2 weak-scaled line of the N = 1, 2, 4 runs, kept beside the 8-GPU config (the C3 buffers are released first)
        m["wl"].tb = m["wl"].to = None
        torch.cuda.empty_cache()
        second = measure("C2", max(3, args.steps // 2), 3, 3, 2, False)
        second["wl"].tb = second["wl"].to = None
    wl = m["wl"]
    stats, stats2, el, el2, sums = m["stats"], m["stats2"], m["el"], m["el2"], m["sums"]
    inst_all, inst_ins, inst2_all, inst2_ins = sums[0], sums[1], sums[2], sums[3]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    nks = len(wl.ks)
    # per-kernel device times (CUDA events on the library's stream, summed over the timed steps of rank 0)
    kern = {}
    for kname, key in (("parse_kernel", "ms_parse_count"), ("fill_kernel", "ms_parse_fill"), ("count_kernel", "ms_insert")):
        kern[kname] = sum(s[key] for s in stats)
    launches = {"parse_kernel": sum(s["n_parse_launches"] for s in stats) // 2,
                "fill_kernel": sum(s["n_parse_launches"] for s in stats) // 2,
                "count_kernel": sum(s["n_insert_launches"] for s in stats)}
    dom = max((n for n in kern if n in launches), key=lambda n: kern[n])
    peak, peak_src = hbm_peak()
    per_k = {}
    alg_bytes = 0.0       # whole pipeline: B_alg(k) per instance
    alg_bytes_dom = 0.0   # the dominant kernel's own share of B_alg
    exch = {}
    for ki, k in enumerate(wl.ks):
        sk = [s for i, s in enumerate(stats) if i % nks == ki]
        ik = sum(s["instances_parsed"] for s in sk)
        ic = sum(s["instances_inserted"] for s in sk)
        W = 8 * (k // 32 + 1)
        alg_bytes += ik * b_alg(k)
        # split of B_alg over the kernels: parse + fill read the packed read, the count kernel reads the key and
        # read-modify-writes the counters; the count kernel is charged for the instances IT counted (its shard's)
        alg_bytes_dom += ic * (W + 12) if dom == "count_kernel" else ik * (READ_LEN / (READ_LEN - k - 1)) / 2
        per_k[str(k)] = {"instances_per_pass": sk[0]["instances_parsed"], "distinct": sk[0]["distinct"], "survivors": sk[0]["survivors"],
                         "partitions": sk[0]["num_partitions"], "supermers": sk[0]["supermers_out"],
                         "instances_counted_rank0": sk[0]["instances_inserted"], "supermers_counted_rank0": sk[0]["supermers_in"],
                         "max_probe": sk[0]["max_probe"],
                         "supermer_bytes": sk[0]["supermer_bytes_out"], "table_retries": sk[0]["table_retries"],
                         "table_group_bytes_max": sk[0]["table_bytes_max"],
                         "ms": {kk: round(sum(s[kk] for s in sk) / len(sk), 3) for kk in
                                ("ms_parse_count", "ms_parse_fill", "ms_parse", "ms_exchange", "ms_comm", "ms_insert", "ms_finish")},
                         "b_alg": round(b_alg(k), 2)}
        if world > 1:
            sent = sum(s["supermer_bytes_sent"] for s in sk) / len(sk)
            recv = sum(s["supermer_bytes_recv"] for s in sk) / len(sk)
            ms_comm = sum(s["ms_comm"] for s in sk) / len(sk)
            gbs = max(sent, recv) / (ms_comm * 1e-3) / 1e9 if ms_comm > 0 else None
            exch[str(k)] = {"bytes_sent": int(sent), "bytes_recv": int(recv), "ms_comm": round(ms_comm, 3),
                            "ms_not_hidden": round(sum(s["ms_exchange"] for s in sk) / len(sk), 3),
                            "rounds": int(sum(s["exchange_rounds"] for s in sk) / len(sk)),
                            "gbs_per_dir": gbs, "frac_of_900": gbs / 900.0 if gbs else None}
    achieved = alg_bytes_dom / (kern[dom] * 1e-3) / 1e9
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
        cpu = reference_rate(wl, min(wl.n_reads, args.ref_reads or 40_000 * cores), wl.ks, cores)
    dropin = None
    if world == 1 and args.dropin:
        dropin = dropin_seam(wl, wl.ks)
    value = inst_all / el
    steps, e2es = m["steps"], m["e2e_steps"]
    h2d = m["h2d"]
    d2h = sum(s["d2h_bytes"] for s in stats2) / e2es
    conserved = (inst_all == inst_ins) and (inst2_all == inst2_ins) and (sums[4] == sums[5])
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": args.warmup,
        "ms_per_step": 1000 * el / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic", "config": m["config"],
        "value_excludes": "H2D of the reads and D2H of the table (reads resident in HBM when the timed region starts; `e2e` includes both)",
        "timing": "CUDA events around the K steps (torch.cuda.synchronize + barrier on both sides; the library calls are "
                  "host-synchronous), max over ranks; per-kernel times are CUDA events on the library's stream",
        "wall_ms_per_step": round(1000 * m["wall"] / steps, 3),
        "e2e": {"value": inst2_all / el2, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(d2h),
                "steps": e2es, "ms_per_step": 1000 * el2 / e2es,
                "path": ("kc_add_reads per k round" if args.e2e_per_round_upload else
                         "kc_reads_create once per step (pinned host reads -> HBM, shared by the k rounds)") +
                        " + kc_add_reads_resident + kc_finish_to_host (table -> pinned host memory) per k round"},
        "e2e_copy_ceiling": dict(ceiling, **{
            "what": f"all {world} rank(s) copying at once, pinned memory, GB/s per rank and direction (min over ranks)",
            "floor_ms_per_step": round(1e3 * max(h2d / (ceiling["h2d_gbs_per_rank_per_dir"] * 1e9), d2h / (ceiling["d2h_gbs_per_rank_per_dir"] * 1e9)), 2),
            "floor_ms_note": "the larger of the step's upload and download at those rates; the copies overlap the kernels and each other only in part"}),
        "e2e_dropin": dropin,
        "gpu_launches": int(sum(s["gpu_launches"] for s in stats)),
        "parity": m["parity"],
        "conservation": {"ok": conserved, "instances_parsed": inst_all, "instances_inserted": inst_ins,
                         "exchange_bytes_sent": sums[4], "exchange_bytes_received": sums[5],
                         "what": "summed over ranks and timed steps: every parsed instance is in some shard's table, every byte sent arrived"},
        "exchange": exch or None,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "launches": launches[dom],
                     "avg_launch_ms": kern[dom] / max(1, launches[dom]),
                     "algorithmic_bytes": "count_kernel is charged its own share of B_alg(k) = L/(L-k-1) + W + 12: the key read "
                                          "and the counter read-modify-writes, W + 12 bytes per instance it counted; the packed-read "
                                          "bytes L/(L-k-1) belong to parse_kernel + fill_kernel (DESIGN.md §4)",
                     "pipeline_frac": (alg_bytes / el / 1e9) / peak,
                     "pipeline_achieved": alg_bytes / el / 1e9},
        "kernel_ms": {k_: round(v, 3) for k_, v in kern.items()}, "per_k": per_k,
        "cpu_baseline": cpu, "clocks": m["clocks"], "gen_s": round(wl.gen_s, 1),
    }
    if second is not None:
        s = second
        out["c2_weak_scaled"] = {"value": s["sums"][0] / s["el"], "unit": UNIT, "ms_per_step": 1000 * s["el"] / s["steps"],
                                 "steps": s["steps"], "e2e_value": s["sums"][2] / s["el2"], "e2e_ms_per_step": 1000 * s["el2"] / s["e2e_steps"],
                                 "config": s["config"]["workload"],
                                 "conserved": s["sums"][0] == s["sums"][1] and s["sums"][4] == s["sums"][5]}
    emit(out)
    if world > 1:
        dist.destroy_process_group()
    bad = (m["parity"] and m["parity"].get("checked") and not m["parity"]["equal"]) or not conserved
    if bad:
        sys.stderr.write("bench.py: PARITY FAILURE — the timed tables differ from the reference's (see `parity` / `conservation`)\n")
        sys.exit(3)


_REAL_STDOUT = None


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, line)
    else:
        sys.stdout.write(line.decode())


def gen_reads(n_reads, seed, pinned=True):
    """C2-shaped reads with an explicit seed (tests and tools): PackedRead bytes + offsets as torch tensors."""
    import torch
    from mhm2_proxy_b200 import synth
    w = WORKLOADS["C2"]
    com = synth.make_community(w["genome_bp"], n_genomes=w["n_genomes"], sigma=1.0, seed=w["seed"] & 0xFFFF)
    pin = pinned and torch.cuda.is_available()
    tb = torch.empty(n_reads * READ_LEN, dtype=torch.uint8, pin_memory=pin)
    _, ro = synth.make_reads(n_reads, com, read_len=READ_LEN, error_rate=w["err"], seed=seed, out=tb.numpy())
    to = torch.empty(n_reads + 1, dtype=torch.int64, pin_memory=pin)
    to.numpy()[:] = ro.astype(np.int64)
    return tb, to


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=None, choices=sorted(WORKLOADS), help="BASELINE.json config (default: C2; C3 at --gpus 8)")
    ap.add_argument("--reads", type=int, default=0, help="reads per GPU (default: the config's)")
    ap.add_argument("--ks", type=lambda s: tuple(int(x) for x in s.split(",")), default=None)
    ap.add_argument("--partitions", type=int, default=0)
    ap.add_argument("--e2e-steps", dest="e2e_steps", type=int, default=10)
    ap.add_argument("--e2e-per-round-upload", dest="e2e_per_round_upload", action="store_true",
                    help="end-to-end arm: copy the reads host->device again for every k round (kc_add_reads) instead of once per step")
    ap.add_argument("--ref-reads", dest="ref_reads", type=int, default=0, help="reads in the CPU sample (default 40k per core)")
    ap.add_argument("--ref-warm", dest="ref_warm", action="store_true", help="also run warm-up passes in --impl reference")
    ap.add_argument("--no-cpu", dest="no_cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--parity", default="on", choices=["on", "off"], help="digest check against the reference build (outside the timed region)")
    ap.add_argument("--parity-reads", dest="parity_reads", type=int, default=0,
                    help="reads in the parity check (default: N=1 the whole workload, N>1 one million in total)")
    ap.add_argument("--parity-distinct-frac", dest="parity_distinct_frac", type=float, default=0.0,
                    help="sizes the reference's tables: num_elems = frac x instances per rank (default by config)")
    ap.add_argument("--parity-chunk", dest="parity_chunk", type=int, default=2_000_000)
    ap.add_argument("--dropin", action="store_true", help="also time the reference's own analyze_kmers through the binding (e2e_dropin)")
    ap.add_argument("--no-second", dest="no_second", action="store_true", help="--gpus 8 default run: skip the C2 weak-scaled second line")
    ap.add_argument("--second", action="store_true", help="also measure the C2 weak-scaled second line (what the --gpus 8 default run does)")
    args = ap.parse_args()
    if not args.parity_distinct_frac:
        name = args.config or default_config(args.gpus)
        # distinct / instances: ~0.13-0.19 at 33x and 0.5 % error (C2, C4); up to ~0.5 at 1 % error; ~0.7 at 3x (C1)
        args.parity_distinct_frac = {"C1": 0.8, "C2": 0.25, "C3": 0.55, "C4": 0.3, "C5": 0.55}[name]
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

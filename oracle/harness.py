"""TEST INFRASTRUCTURE ONLY — ctypes front end to the two CPU checkers.

* ``port``: oracle/libkcount_oracle.so, the plain-C restatement (kcount_oracle.c), always buildable.
* ``ref`` : oracle/_ref/libkcount_ref.so, the reference's own kcount_cpu.cpp/kmer_dht.cpp/kmer.cpp/hash_funcs.c compiled
  unmodified from /root/reference (ref_driver.cpp); built in the authoring container, travels to the GPU box prebuilt.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
The product package (mhm2_proxy_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libkcount_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libkcount_ref.so")


def build(verbose=False):
    """Compile the C restatement and, when /root/reference is present, the reference build."""
    out = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if verbose or out.returncode:
        print(out.stdout[-2000:], out.stderr[-4000:])
    if out.returncode:
        raise RuntimeError("oracle build failed")


def have_ref():
    return os.path.exists(REF_SO)


_libs = {}


def _lib(kind):
    if kind in _libs:
        return _libs[kind]
    path, pfx = (PORT_SO, "ko_") if kind == "port" else (REF_SO, "oref_")
    if not os.path.exists(path):
        if kind == "port":
            build()
        else:
            raise FileNotFoundError(path)
    lib = C.CDLL(path)
    f = lambda n: getattr(lib, pfx + n)
    run = f("run")
    run.restype = C.c_void_p
    common = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
    run.argtypes = common + ([C.c_int] if kind == "port" else [C.c_double, C.c_uint64, C.c_int])
    f("n_longs").argtypes = [C.c_void_p]
    f("n_longs").restype = C.c_int
    for n in ("num_survivors", "num_prefilter"):
        f(n).argtypes = [C.c_void_p]
        f(n).restype = C.c_uint64
    f("copy_survivors").argtypes = [C.c_void_p] * 6
    f("copy_prefilter").argtypes = [C.c_void_p] * 3
    f("stats").argtypes = [C.c_void_p] * 3
    f("free").argtypes = [C.c_void_p]
    f("kmer_prims").argtypes = [C.c_int, C.c_char_p, C.c_int] + [C.c_void_p] * 6
    f("kmer_prims").restype = C.c_int
    f("murmur64").argtypes = [C.c_void_p, C.c_uint32]
    f("murmur64").restype = C.c_uint64
    f("quick_hash").argtypes = [C.c_uint64]
    f("quick_hash").restype = C.c_uint64
    g = f("get_ext4") if kind == "port" else f("get_ext")
    g.argtypes = [C.c_uint16] * 5 + [C.c_int]
    g.restype = C.c_char
    _libs[kind] = (lib, pfx)
    return _libs[kind]


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleResult:
    """Survivors sorted by key words; prefilter rows (after the reads pass) sorted by key words too."""

    def __init__(self):
        self.n_longs = 0
        self.keys = self.counts = self.left = self.right = self.shard = None
        self.pf_keys = self.pf_vals = None
        self.stats = {}
        self.times = (0.0, 0.0, 0.0)


def run(kind, k, read_bytes, read_offsets, dmin_thres=2, num_shards=1, qual_offset=33, ctg_seqs=None, ctg_offsets=None,
        ctg_depths=None, want_prefilter=False, free_mem=None, num_elems=None):
    """Run kcount on the CPU. read_bytes: uint8 PackedRead bytes; read_offsets: uint64[n+1]."""
    lib, pfx = _lib(kind)
    f = lambda n: getattr(lib, pfx + n)
    rb = np.ascontiguousarray(read_bytes, dtype=np.uint8)
    ro = np.ascontiguousarray(read_offsets, dtype=np.uint64)
    n_reads = len(ro) - 1
    n_ctgs = 0
    cs = co = cd = None
    if ctg_offsets is not None and len(ctg_offsets) > 1:
        cs = np.ascontiguousarray(ctg_seqs, dtype=np.uint8)
        co = np.ascontiguousarray(ctg_offsets, dtype=np.uint64)
        cd = np.ascontiguousarray(ctg_depths, dtype=np.float64)
        n_ctgs = len(co) - 1
    args = [k, qual_offset, dmin_thres, num_shards, _p(rb), _p(ro), n_reads, _p(cs), _p(co), _p(cd), n_ctgs]
    if kind == "port":
        args.append(int(want_prefilter))
    else:
        lens = np.diff(ro.astype(np.int64))
        inst = int(np.maximum(lens - k - 1, 0).sum())
        if n_ctgs:
            inst += int(np.maximum(np.diff(co.astype(np.int64)) - k - 1, 0).sum())
        ne = int(num_elems if num_elems is not None else max(inst, 1000))
        w = 8 * (k // 32 + 1)
        # capacity = min(0.6*free/(W+20), 3*num_elems) (kcount_cpu.cpp:431-438): make memory non-binding by default
        fm = float(free_mem if free_mem is not None else 3.2 * ne * (w + 20) / 0.6)
        args += [fm, ne, int(want_prefilter)]
    h = f("run")(*args)
    if not h:
        raise RuntimeError(f"oracle {kind} run failed")
    try:
        res = OracleResult()
        nl = res.n_longs = f("n_longs")(h)
        n = f("num_survivors")(h)
        res.keys = np.zeros((n, nl), dtype=np.uint64)
        res.counts = np.zeros(n, dtype=np.uint16)
        res.left = np.zeros(n, dtype=np.uint8)
        res.right = np.zeros(n, dtype=np.uint8)
        res.shard = np.zeros(n, dtype=np.int32)
        f("copy_survivors")(h, _p(res.keys), _p(res.counts), _p(res.left), _p(res.right), _p(res.shard))
        if want_prefilter:
            npf = f("num_prefilter")(h)
            pk = np.zeros((npf, nl), dtype=np.uint64)
            pv = np.zeros((npf, 9), dtype=np.uint16)
            f("copy_prefilter")(h, _p(pk), _p(pv))
            order = np.lexsort([pk[:, j] for j in range(nl - 1, -1, -1)])
            res.pf_keys, res.pf_vals = pk[order], pv[order]
        st = np.zeros(8, dtype=np.uint64)
        tm = np.zeros(3, dtype=np.float64)
        f("stats")(h, _p(st), _p(tm))
        names = ["instances", "distinct_after_reads", "dropped", "singleton_overrides", "supermers", "supermer_bytes",
                 "ctg_instances", "purged"]
        res.stats = {nm: int(v) for nm, v in zip(names, st)}
        res.times = tuple(float(x) for x in tm)
        return res
    finally:
        f("free")(h)


def kmer_prims(kind, kmer_str, num_shards=1):
    """(words, rc_words, minimizer, slot_hash, owner, m) of an ASCII k-mer."""
    lib, pfx = _lib(kind)
    k = len(kmer_str)
    w = np.zeros(4, dtype=np.uint64)
    r = np.zeros(4, dtype=np.uint64)
    mn = np.zeros(1, dtype=np.uint64)
    hs = np.zeros(1, dtype=np.uint64)
    ow = np.zeros(1, dtype=np.int32)
    m = np.zeros(1, dtype=np.int32)
    nl = getattr(lib, pfx + "kmer_prims")(k, kmer_str.encode(), num_shards, _p(w), _p(r), _p(mn), _p(hs), _p(ow), _p(m))
    return w[:nl].copy(), r[:nl].copy(), int(mn[0]), int(hs[0]), int(ow[0]), int(m[0])


def murmur64(kind, words):
    lib, pfx = _lib(kind)
    a = np.ascontiguousarray(words, dtype=np.uint64)
    return int(getattr(lib, pfx + "murmur64")(_p(a), a.size * 8))


def quick_hash(kind, v):
    lib, pfx = _lib(kind)
    return int(getattr(lib, pfx + "quick_hash")(int(v)))


def get_ext(kind, a, c, g, t, count, dmin_thres=2):
    lib, pfx = _lib(kind)
    fn = getattr(lib, pfx + ("get_ext4" if kind == "port" else "get_ext"))
    return fn(a, c, g, t, count, dmin_thres).decode()


# ---- P-rank timed baseline: the reference's code, one process per emulated UPC++ rank (ref_driver.cpp phase_a/phase_b) ----

def _phase_main(argv):
    """Entry of one emulated rank (a separate `python oracle/harness.py phase ...` process). Prints one JSON line."""
    import json
    phase, k, P, me, prefix, dmin_thres, num_elems = argv[0], int(argv[1]), int(argv[2]), int(argv[3]), argv[4], int(argv[5]), int(argv[6])
    lib = C.CDLL(REF_SO)
    if phase == "a":
        rb = np.ascontiguousarray(np.load(prefix + "_bytes.npy", mmap_mode="r"))
        ro = np.ascontiguousarray(np.load(prefix + "_offs.npy", mmap_mode="r"))
        n = len(ro) - 1
        r0, r1 = n * me // P, n * (me + 1) // P
        lib.oref_phase_a.restype = C.c_double
        lib.oref_phase_a.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_char_p,
                                     C.c_void_p]
        inst = np.zeros(1, dtype=np.uint64)
        t = lib.oref_phase_a(k, 33, P, me, _p(rb), _p(ro), r0, r1, prefix.encode(), _p(inst))
        print(json.dumps([t, int(inst[0]), 0]))
        return
    lib.oref_phase_b.restype = C.c_double
    lib.oref_phase_b.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p]
    surv = np.zeros(1, dtype=np.uint64)
    dist = np.zeros(1, dtype=np.uint64)
    t = lib.oref_phase_b(k, dmin_thres, P, me, prefix.encode(), num_elems, _p(surv), _p(dist))
    print(json.dumps([t, int(surv[0]), int(dist[0])]))


def run_multirank(k, read_bytes, read_offsets, P, dmin_thres=2, workdir=None, timeout=1200):
    """Time the reference kcount with P emulated ranks on P processes. Returns dict(seconds, instances, survivors, ...).
    seconds = max over ranks of phase A + max over ranks of phase B (bulk-synchronous, like a barrier between them)."""
    import glob
    import json
    import sys
    import tempfile
    if not have_ref():
        raise FileNotFoundError(REF_SO)
    base = workdir or ("/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir())
    d = tempfile.mkdtemp(prefix="kcref_", dir=base)
    prefix = os.path.join(d, "x")
    try:
        np.save(prefix + "_bytes.npy", np.ascontiguousarray(read_bytes, dtype=np.uint8))
        np.save(prefix + "_offs.npy", np.ascontiguousarray(read_offsets, dtype=np.uint64))
        lens = np.diff(np.asarray(read_offsets).astype(np.int64))
        inst = int(np.maximum(lens - k - 1, 0).sum())
        num_elems = max(1000, int(inst * 1.3 / P))

        def phase(ph):
            procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "phase", ph, str(k), str(P), str(me), prefix,
                                       str(dmin_thres), str(num_elems)], stdout=subprocess.PIPE, text=True) for me in range(P)]
            out = []
            for pr in procs:
                o, _ = pr.communicate(timeout=timeout)
                if pr.returncode:
                    raise RuntimeError(f"reference rank process failed in phase {ph}")
                out.append(json.loads(o.strip().splitlines()[-1]))
            return out

        ra = phase("a")
        rbb = phase("b")
        ta, tb = max(r[0] for r in ra), max(r[0] for r in rbb)
        return dict(seconds=ta + tb, seconds_parse=ta, seconds_insert=tb, instances=sum(r[1] for r in ra),
                    survivors=sum(r[1] for r in rbb), distinct=sum(r[2] for r in rbb), ranks=P)
    finally:
        for f in glob.glob(os.path.join(d, "*")):
            os.remove(f)
        os.rmdir(d)


# ---- streaming P-rank run (full-size parity): persistent rank processes, reads fed in chunks (ref_driver.cpp RankWorker) ----

def _worker_main(argv):
    """One emulated rank as a long-lived process: JSON commands on stdin, one JSON answer per command on stdout."""
    import json
    import sys
    k, P, me, dmin_thres, num_elems, prefix = int(argv[0]), int(argv[1]), int(argv[2]), int(argv[3]), int(argv[4]), argv[5]
    lib = C.CDLL(REF_SO)
    lib.oref_w_open.restype = C.c_void_p
    lib.oref_w_open.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64]
    lib.oref_w_insert.restype = C.c_double
    lib.oref_w_insert.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
    lib.oref_w_finish.restype = C.c_double
    lib.oref_w_finish.argtypes = [C.c_void_p, C.c_void_p]
    lib.oref_w_free.argtypes = [C.c_void_p]
    lib.oref_phase_a.restype = C.c_double
    lib.oref_phase_a.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_char_p, C.c_void_p]
    lib.oref_phase_a_ctgs.restype = C.c_double
    lib.oref_phase_a_ctgs.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_char_p,
                                      C.c_void_p]
    rb = np.load(prefix + "_bytes.npy", mmap_mode="r")
    ro = np.load(prefix + "_offs.npy", mmap_mode="r")
    have_ctg = os.path.exists(prefix + "_cseqs.npy")
    if have_ctg:
        cs, co, cd = (np.load(prefix + "_cseqs.npy", mmap_mode="r"), np.load(prefix + "_coffs.npy", mmap_mode="r"),
                      np.load(prefix + "_cdepths.npy", mmap_mode="r"))
    w = lib.oref_w_open(k, dmin_thres, P, me, num_elems)
    if not w:
        print(json.dumps({"error": "open"}), flush=True)
        return
    print(json.dumps({"ok": 1}), flush=True)
    inst = np.zeros(1, dtype=np.uint64)
    for line in sys.stdin:
        cmd = json.loads(line)
        op = cmd["op"]
        if op == "a":  # parse my 1/P of reads [lo, hi) with rank_n() = P; supermers go to <chunk>_<me>_<target>.bin
            lo, hi = cmd["lo"], cmd["hi"]
            n = hi - lo
            r0, r1 = lo + n * me // P, lo + n * (me + 1) // P
            t = lib.oref_phase_a(k, 33, P, me, rb.ctypes.data, ro.ctypes.data, r0, r1, cmd["chunk"].encode(), _p(inst))
            print(json.dumps({"t": t, "inst": int(inst[0])}), flush=True)
        elif op == "actg":
            n = len(co) - 1
            c0, c1 = n * me // P, n * (me + 1) // P
            t = lib.oref_phase_a_ctgs(k, P, me, cs.ctypes.data, co.ctypes.data, cd.ctypes.data, c0, c1, cmd["chunk"].encode(), _p(inst))
            print(json.dumps({"t": t, "inst": int(inst[0])}), flush=True)
        elif op == "b":
            t = lib.oref_w_insert(w, cmd["chunk"].encode(), int(cmd.get("ctg", 0)))
            print(json.dumps({"t": t}), flush=True)
        elif op == "finish":
            out = np.zeros(8, dtype=np.uint64)
            t = lib.oref_w_finish(w, _p(out))
            print(json.dumps({"t": t, "out": [int(x) for x in out]}), flush=True)
            lib.oref_w_free(w)
            return


def run_multirank_stream(k, read_bytes, read_offsets, P, dmin_thres=2, ctg_seqs=None, ctg_offsets=None, ctg_depths=None,
                         chunk_reads=2_000_000, distinct_frac=0.5, workdir=None):
    """The reference kcount with P emulated ranks (one process each) over the WHOLE read set, reads streamed in chunks so
    that /dev/shm never holds more than one chunk's supermers. Returns per-rank digests of local_kmers
    (n, sum counts, sum mix, xor mix) — rank r owns quick_hash(minimizer) % P == r (kmer_dht.cpp:192-196) — plus the
    bulk-synchronous time (sum over chunks of max parse + max insert) and the reference's dropped / override counters.
    distinct_frac sizes every rank's table: num_elems = distinct_frac * instances / P (capacity is 3x that)."""
    import glob
    import json
    import sys
    import tempfile
    if not have_ref():
        raise FileNotFoundError(REF_SO)
    base = workdir or ("/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir())
    d = tempfile.mkdtemp(prefix="kcrefs_", dir=base)
    prefix = os.path.join(d, "x")
    procs = []
    try:
        ro = np.ascontiguousarray(read_offsets, dtype=np.uint64)
        np.save(prefix + "_bytes.npy", np.ascontiguousarray(read_bytes, dtype=np.uint8))
        np.save(prefix + "_offs.npy", ro)
        n_reads = len(ro) - 1
        inst = int(np.maximum(np.diff(ro.astype(np.int64)) - k - 1, 0).sum())
        have_ctg = ctg_offsets is not None and len(ctg_offsets) > 1
        if have_ctg:
            co = np.ascontiguousarray(ctg_offsets, dtype=np.uint64)
            np.save(prefix + "_cseqs.npy", np.ascontiguousarray(ctg_seqs, dtype=np.uint8))
            np.save(prefix + "_coffs.npy", co)
            np.save(prefix + "_cdepths.npy", np.ascontiguousarray(ctg_depths, dtype=np.float64))
            inst += int(np.maximum(np.diff(co.astype(np.int64)) - k - 1, 0).sum())
        num_elems = max(1000, int(inst * distinct_frac / P))
        for me in range(P):
            procs.append(subprocess.Popen([sys.executable, os.path.abspath(__file__), "worker", str(k), str(P), str(me), str(dmin_thres),
                                           str(num_elems), prefix], stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True, bufsize=1))

        def ask(cmd):
            line = json.dumps(cmd) + "\n"
            for pr in procs:
                pr.stdin.write(line)
                pr.stdin.flush()
            out = []
            for pr in procs:
                ans = pr.stdout.readline()
                if not ans:
                    raise RuntimeError("a reference rank worker died")
                out.append(json.loads(ans))
            return out

        for pr in procs:
            if "ok" not in json.loads(pr.stdout.readline() or '{"error": "died"}'):
                raise RuntimeError("reference rank worker failed to open its table")
        t_parse = t_ins = 0.0
        n_inst = 0
        ci = 0
        for lo in range(0, n_reads, chunk_reads):
            hi = min(n_reads, lo + chunk_reads)
            chunk = f"{prefix}_c{ci}"
            ci += 1
            ra = ask({"op": "a", "lo": lo, "hi": hi, "chunk": chunk})
            rb_ = ask({"op": "b", "chunk": chunk})
            t_parse += max(r["t"] for r in ra)
            t_ins += max(r["t"] for r in rb_)
            n_inst += sum(r["inst"] for r in ra)
        if have_ctg:
            chunk = f"{prefix}_ctg"
            ra = ask({"op": "actg", "chunk": chunk})
            rb_ = ask({"op": "b", "chunk": chunk, "ctg": 1})
            t_parse += max(r["t"] for r in ra)
            t_ins += max(r["t"] for r in rb_)
            n_inst += sum(r["inst"] for r in ra)
        rf = ask({"op": "finish"})
        t_ins += max(r["t"] for r in rf)
        outs = [r["out"] for r in rf]
        for pr in procs:
            pr.stdin.close()
            pr.wait(timeout=60)
        return dict(seconds=t_parse + t_ins, seconds_parse=t_parse, seconds_insert=t_ins, instances=n_inst, ranks=P,
                    digests=[tuple(o[:4]) for o in outs], survivors=sum(o[0] for o in outs), distinct=sum(o[4] for o in outs),
                    dropped=sum(o[5] for o in outs), singleton_overrides=sum(o[6] for o in outs))
    finally:
        for pr in procs:
            if pr.poll() is None:
                pr.kill()
        for f in glob.glob(os.path.join(d, "*")):
            os.remove(f)
        os.rmdir(d)


def combine_digests(digests, n_groups=1):
    """Per-rank digests of a P-rank run -> digests of n_groups shards, shard g = ranks r with r % n_groups == g (valid because
    n_groups divides P: x % P % n_groups == x % n_groups)."""
    M = (1 << 64) - 1
    out = []
    for g in range(n_groups):
        n = sc = sm = xm = 0
        for r in range(g, len(digests), n_groups):
            a, b, c, e = digests[r]
            n += a
            sc += b
            sm = (sm + c) & M
            xm ^= e
        out.append((n, sc, sm, xm))
    return out


def pack_reads_ref(seqs, quals, offsets, qual_offset=33):
    """FASTQ text -> PackedRead bytes through the reference's own PackedRead constructor (packed_reads.cpp:73-109)."""
    lib = C.CDLL(REF_SO)
    lib.oref_pack_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
    s = np.ascontiguousarray(seqs, dtype=np.uint8)
    q = np.ascontiguousarray(quals, dtype=np.uint8)
    o = np.ascontiguousarray(offsets, dtype=np.uint64)
    out = np.empty(int(o[-1]), dtype=np.uint8)
    if lib.oref_pack_reads(_p(s), _p(q), _p(o), len(o) - 1, qual_offset, _p(out)) != 0:
        raise RuntimeError("oref_pack_reads failed")
    return out


if __name__ == "__main__":
    import sys
    if len(sys.argv) > 1 and sys.argv[1] == "phase":
        _phase_main(sys.argv[2:])
    elif len(sys.argv) > 1 and sys.argv[1] == "worker":
        _worker_main(sys.argv[2:])


# ---- drop-in proof: the reference's own analyze_kmers with integration/kcount_b200.cpp as its backend (dropin_driver.cpp) ----
DROPIN_SO = os.path.join(HERE, "_ref", "libkcount_dropin.so")


def have_dropin():
    return os.path.exists(DROPIN_SO)


def run_dropin(k, read_bytes, read_offsets, dmin_thres=2, qual_offset=33, ctg_seqs=None, ctg_offsets=None, ctg_depths=None):
    """analyze_kmers<MAX_K> of the reference (kcount.cpp) over PackedReads + Contigs, kcount_b200.cpp behind it; returns the
    contents of kmer_dht->local_kmers sorted by key. Needs a GPU (the binding drives libkcount_b200.so)."""
    lib = C.CDLL(DROPIN_SO)
    lib.odrop_run.restype = C.c_void_p
    lib.odrop_run.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
    lib.odrop_n_longs.argtypes = [C.c_void_p]
    lib.odrop_num_survivors.argtypes = [C.c_void_p]
    lib.odrop_num_survivors.restype = C.c_uint64
    lib.odrop_copy_survivors.argtypes = [C.c_void_p] * 5
    lib.odrop_free.argtypes = [C.c_void_p]
    rb = np.ascontiguousarray(read_bytes, dtype=np.uint8)
    ro = np.ascontiguousarray(read_offsets, dtype=np.uint64)
    cs = co = cd = None
    n_ctgs = 0
    if ctg_offsets is not None and len(ctg_offsets) > 1:
        cs = np.ascontiguousarray(ctg_seqs, dtype=np.uint8)
        co = np.ascontiguousarray(ctg_offsets, dtype=np.uint64)
        cd = np.ascontiguousarray(ctg_depths, dtype=np.float64)
        n_ctgs = len(co) - 1
    h = lib.odrop_run(k, qual_offset, dmin_thres, _p(rb), _p(ro), len(ro) - 1, _p(cs), _p(co), _p(cd), n_ctgs)
    if not h:
        raise RuntimeError("drop-in run failed")
    try:
        res = OracleResult()
        nl = res.n_longs = lib.odrop_n_longs(h)
        n = lib.odrop_num_survivors(h)
        res.keys = np.zeros((n, nl), dtype=np.uint64)
        res.counts = np.zeros(n, dtype=np.uint16)
        res.left = np.zeros(n, dtype=np.uint8)
        res.right = np.zeros(n, dtype=np.uint8)
        lib.odrop_copy_survivors(h, _p(res.keys), _p(res.counts), _p(res.left), _p(res.right))
        return res
    finally:
        lib.odrop_free(h)

"""ctypes binding of the C ABI in include/kcount_b200.h (libkcount_b200.so, built in-tree by __graft_entry__.build()).

This is the Python stand-in for the reference-side binding (INTEGRATION.md shows the C++ one). It is plumbing only:
every data-path step runs in the CUDA library; there is no CPU fallback and a missing library is a hard error.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("KCOUNT_B200_LIB") or os.path.join(_HERE, "libkcount_b200.so")  # override: A/B builds of the same ABI

KC_OK = 0
ERR_NAMES = {1: "INVALID", 2: "CUDA", 3: "STATE", 4: "NOMEM", 5: "BADCHAR", 6: "COMM", 7: "OVERFLOW"}


class KcConfig(C.Structure):
    _fields_ = [
        ("kmer_len", C.c_int32), ("qual_cutoff", C.c_int32), ("dmin_thres", C.c_int32), ("dyn_min_depth", C.c_double),
        ("num_shards", C.c_int32), ("shard_id", C.c_int32), ("device", C.c_int32), ("num_partitions", C.c_int32),
        ("expected_instances", C.c_uint64), ("keep_prefilter", C.c_int32), ("batch_reads", C.c_int32),
        ("reserved", C.c_int32 * 6),
    ]


class KcStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "num_seqs", "num_bases", "instances_parsed", "instances_inserted", "ctg_instances_inserted", "supermers_out",
        "supermer_bytes_out", "supermer_bytes_sent", "supermers_in", "distinct", "purged", "survivors",
        "table_retries", "max_probe", "num_partitions", "supermer_bytes_recv", "exchange_rounds")] + [(n, C.c_double) for n in (
            "ms_h2d", "ms_parse", "ms_exchange", "ms_insert", "ms_comm", "ms_total", "ms_parse_count", "ms_parse_fill",
            "ms_finish")] + [(n, C.c_uint64) for n in (
                "gpu_launches", "n_parse_launches", "n_insert_launches", "table_bytes_max")] + [
                    ("reserved", C.c_uint64 * 4)]


EXPORTS = [
    "kc_config_default", "kc_create", "kc_destroy", "kc_last_error", "kc_release_cached_memory", "kc_add_reads", "kc_add_reads_device", "kc_reads_create", "kc_reads_create_fastq", "kc_reads_copy_packed", "kc_add_reads_resident", "kc_reads_destroy", "kc_dump_kmers", "kc_add_reads_ascii", "kc_add_contigs", "kc_set_dmin_thres",
    "kc_comm_unique_id", "kc_comm_init", "kc_finish", "kc_finish_to_host", "kc_finish_local_group", "kc_num_survivors", "kc_copy_survivors", "kc_survivors_device",
    "kc_num_prefilter", "kc_copy_prefilter", "kc_get_stats", "kc_prim_kmers", "kc_prim_get_ext",
    "kc_flush", "kc_device_count", "kc_table_digest", "kc_debug_set_key_cap", "kc_struct_sizes",
    "kc_table_digest_static",
]

_lib = None


def load_library():
    """dlopen the CUDA library; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("KCOUNT_B200_LIB", LIB_PATH)  # A/B runs of differently compiled builds (tools/ab_bench.sh)
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(the kcount path is CUDA only; there is no CPU fallback)")
    lib = C.CDLL(path)
    vp, u64, i32 = C.c_void_p, C.c_uint64, C.c_int32
    lib.kc_config_default.argtypes = [C.POINTER(KcConfig), i32]
    lib.kc_config_default.restype = None
    lib.kc_create.argtypes = [C.POINTER(KcConfig), C.POINTER(vp)]
    lib.kc_destroy.argtypes = [vp]
    lib.kc_last_error.argtypes = [vp]
    lib.kc_last_error.restype = C.c_char_p
    lib.kc_add_reads.argtypes = [vp, vp, vp, u64]
    lib.kc_add_reads_device.argtypes = [vp, vp, vp, u64, u64, C.c_uint32]
    lib.kc_add_contigs.argtypes = [vp, vp, vp, vp, u64]
    lib.kc_add_reads_ascii.argtypes = [vp, vp, vp, u64]
    lib.kc_set_dmin_thres.argtypes = [vp, i32]
    lib.kc_comm_unique_id.argtypes = [vp]
    lib.kc_comm_init.argtypes = [vp, vp]
    lib.kc_finish.argtypes = [vp]
    lib.kc_reads_create.argtypes = [i32, vp, vp, u64, i32, C.POINTER(vp)]
    lib.kc_add_reads_resident.argtypes = [vp, vp]
    lib.kc_reads_create_fastq.argtypes = [i32, vp, vp, vp, u64, i32, i32, C.POINTER(vp)]
    lib.kc_reads_copy_packed.argtypes = [vp, vp]
    lib.kc_dump_kmers.argtypes = [vp, C.c_char_p]
    lib.kc_reads_destroy.argtypes = [vp]
    lib.kc_finish_to_host.argtypes = [vp, vp, vp, vp, vp, u64, C.POINTER(u64)]
    lib.kc_finish_local_group.argtypes = [C.POINTER(vp), i32]
    lib.kc_num_survivors.argtypes = [vp, C.POINTER(u64)]
    lib.kc_copy_survivors.argtypes = [vp, vp, vp, vp, vp]
    lib.kc_survivors_device.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.kc_num_prefilter.argtypes = [vp, C.POINTER(u64)]
    lib.kc_copy_prefilter.argtypes = [vp, vp, vp]
    lib.kc_get_stats.argtypes = [vp, C.POINTER(KcStats)]
    lib.kc_prim_kmers.argtypes = [i32, i32, C.c_char_p, u64, i32, vp, vp, vp, vp, vp]
    lib.kc_prim_get_ext.argtypes = [i32, vp, u64, i32, C.c_double, vp]
    lib.kc_flush.argtypes = [vp]
    lib.kc_device_count.argtypes = [C.POINTER(i32)]
    lib.kc_table_digest.argtypes = [vp, vp]
    lib.kc_debug_set_key_cap.argtypes = [vp, i32]
    lib.kc_table_digest_static.argtypes = [i32, vp, vp, vp, vp, u64, i32, vp, vp, C.c_char_p, u64]
    lib.kc_struct_sizes.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.kc_struct_sizes.restype = None
    a, b = C.c_uint32(), C.c_uint32()
    lib.kc_struct_sizes(C.byref(a), C.byref(b))
    if (a.value, b.value) != (C.sizeof(KcConfig), C.sizeof(KcStats)):
        raise RuntimeError(f"{LIB_PATH}: struct sizes {a.value}/{b.value} differ from this binding's {C.sizeof(KcConfig)}/"
                           f"{C.sizeof(KcStats)} (stale build?)")
    _lib = lib
    return lib


class KcountError(RuntimeError):
    pass


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class KmerCounter:
    """One shard of the k-mer table = the reference's KmerDHT<MAX_K> + SeqBlockInserter + HashTableInserter for one rank
    (kmer_dht.hpp:119-171, kcount.hpp:57-69), driven in the order analyze_kmers uses (kcount.cpp:140-157):
    add_reads* -> [add_contigs] -> finish -> survivors."""

    def __init__(self, kmer_len, dmin_thres=2, num_shards=1, shard_id=0, device=0, num_partitions=0, expected_instances=0,
                 keep_prefilter=False, batch_reads=0, dbg_key_cap=0, qual_cutoff=20, dyn_min_depth=0.9):
        self.lib = load_library()
        cfg = KcConfig()
        self.lib.kc_config_default(C.byref(cfg), kmer_len)
        cfg.dmin_thres, cfg.num_shards, cfg.shard_id, cfg.device = dmin_thres, num_shards, shard_id, device
        cfg.num_partitions, cfg.expected_instances = num_partitions, expected_instances
        cfg.keep_prefilter, cfg.batch_reads = int(keep_prefilter), batch_reads
        cfg.qual_cutoff, cfg.dyn_min_depth = qual_cutoff, dyn_min_depth
        self.k = kmer_len
        self.n_longs = kmer_len // 32 + 1
        self.h = C.c_void_p()
        rc = self.lib.kc_create(C.byref(cfg), C.byref(self.h))
        if rc != KC_OK:
            msg = self.lib.kc_last_error(self.h).decode() if self.h else ""
            if self.h:
                self.lib.kc_destroy(self.h)
                self.h = None
            raise KcountError(f"kc_create failed ({ERR_NAMES.get(rc, rc)}): {msg}")
        if dbg_key_cap:
            self._ck(self.lib.kc_debug_set_key_cap(self.h, dbg_key_cap), "kc_debug_set_key_cap")

    def _ck(self, rc, what):
        if rc != KC_OK:
            raise KcountError(f"{what} failed ({ERR_NAMES.get(rc, rc)}): {self.lib.kc_last_error(self.h).decode()}")

    def add_reads(self, read_bytes, read_offsets):
        """PackedRead bytes (uint8) + offsets (uint64[n+1]) in HOST memory (numpy, ideally backed by pinned memory)."""
        rb = np.ascontiguousarray(read_bytes, dtype=np.uint8)
        ro = np.ascontiguousarray(read_offsets, dtype=np.uint64)
        self._ck(self.lib.kc_add_reads(self.h, _ptr(rb), _ptr(ro), len(ro) - 1), "kc_add_reads")

    def add_reads_ptr(self, bytes_ptr, offsets_ptr, n_reads):
        """Same, raw HOST pointers (e.g. torch pinned tensors' data_ptr())."""
        self._ck(self.lib.kc_add_reads(self.h, C.c_void_p(bytes_ptr), C.c_void_p(offsets_ptr), n_reads), "kc_add_reads")

    def add_reads_device(self, d_bytes_ptr, d_offsets_ptr, n_reads, total_bytes, max_read_len):
        """DEVICE pointers already resident on this shard's GPU."""
        self._ck(self.lib.kc_add_reads_device(self.h, C.c_void_p(d_bytes_ptr), C.c_void_p(d_offsets_ptr), n_reads, total_bytes,
                                              max_read_len), "kc_add_reads_device")

    def add_contigs(self, seqs, offsets, depths):
        cs = np.ascontiguousarray(seqs, dtype=np.uint8)
        co = np.ascontiguousarray(offsets, dtype=np.uint64)
        cd = np.ascontiguousarray(depths, dtype=np.float64)
        self._ck(self.lib.kc_add_contigs(self.h, _ptr(cs), _ptr(co), _ptr(cd), len(co) - 1), "kc_add_contigs")

    def add_reads_ascii(self, seqs, offsets):
        """Reads as ASCII ACGTN, lower case = low quality (the form SeqBlockInserter::process_seq receives)."""
        cs = np.ascontiguousarray(seqs, dtype=np.uint8)
        co = np.ascontiguousarray(offsets, dtype=np.uint64)
        self._ck(self.lib.kc_add_reads_ascii(self.h, _ptr(cs), _ptr(co), len(co) - 1), "kc_add_reads_ascii")

    def set_dmin_thres(self, v):
        self._ck(self.lib.kc_set_dmin_thres(self.h, v), "kc_set_dmin_thres")

    def comm_init(self, unique_id_bytes):
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(unique_id_bytes))
        self._ck(self.lib.kc_comm_init(self.h, buf), "kc_comm_init")

    @staticmethod
    def comm_unique_id():
        lib = load_library()
        buf = (C.c_uint8 * 128)()
        rc = lib.kc_comm_unique_id(buf)
        if rc != KC_OK:
            raise KcountError(f"kc_comm_unique_id failed ({ERR_NAMES.get(rc, rc)})")
        return bytes(buf)

    def flush(self):
        """KmerDHT::flush_updates: collective; drains the exchange rounds of the adds so far (no-op on one shard)."""
        self._ck(self.lib.kc_flush(self.h), "kc_flush")

    def finish(self):
        self._ck(self.lib.kc_finish(self.h), "kc_finish")

    def digest(self):
        """kc_table_digest: (n, sum of counts, sum of row hashes mod 2^64, xor of row hashes) of this shard's table."""
        out = np.zeros(4, dtype=np.uint64)
        self._ck(self.lib.kc_table_digest(self.h, _ptr(out)), "kc_table_digest")
        return tuple(int(x) for x in out)

    def dump_kmers(self, path):
        """KmerDHT::dump_kmers: "KMER count left right" per surviving k-mer; .gz -> gzip."""
        self._ck(self.lib.kc_dump_kmers(self.h, str(path).encode()), "kc_dump_kmers")

    def add_reads_resident(self, reads):
        """Parse reads that a ResidentReads object holds (or is still uploading) in HBM."""
        self._ck(self.lib.kc_add_reads_resident(self.h, reads.r), "kc_add_reads_resident")

    def finish_into(self, keys_ptr, counts_ptr, left_ptr, right_ptr, capacity):
        """kc_finish_to_host: finish and deliver the table into caller-owned HOST buffers (raw pointers, `capacity` entries);
        the copy overlaps the count stage. Returns the number of survivors."""
        n = C.c_uint64()
        self._ck(self.lib.kc_finish_to_host(self.h, C.c_void_p(keys_ptr), C.c_void_p(counts_ptr), C.c_void_p(left_ptr),
                                            C.c_void_p(right_ptr), capacity, C.byref(n)), "kc_finish_to_host")
        return n.value

    def num_survivors(self):
        n = C.c_uint64()
        self._ck(self.lib.kc_num_survivors(self.h, C.byref(n)), "kc_num_survivors")
        return n.value

    def survivors_into(self, keys_ptr, counts_ptr, left_ptr, right_ptr):
        """Copy the surviving table into caller-owned HOST buffers (raw pointers, e.g. pinned memory); returns n."""
        n = self.num_survivors()
        self._ck(self.lib.kc_copy_survivors(self.h, C.c_void_p(keys_ptr), C.c_void_p(counts_ptr), C.c_void_p(left_ptr),
                                            C.c_void_p(right_ptr)), "kc_copy_survivors")
        return n

    def survivors(self, sort=True):
        """(keys[n, n_longs] u64, counts u16, left u8 chars, right u8 chars); sorted by key words when sort=True."""
        n = self.num_survivors()
        keys = np.empty((n, self.n_longs), dtype=np.uint64)
        counts = np.empty(n, dtype=np.uint16)
        left = np.empty(n, dtype=np.uint8)
        right = np.empty(n, dtype=np.uint8)
        self._ck(self.lib.kc_copy_survivors(self.h, _ptr(keys), _ptr(counts), _ptr(left), _ptr(right)), "kc_copy_survivors")
        if sort and n:
            order = np.lexsort([keys[:, j] for j in range(self.n_longs - 1, -1, -1)])
            keys, counts, left, right = keys[order], counts[order], left[order], right[order]
        return keys, counts, left, right

    def survivors_device(self):
        p = [C.c_void_p() for _ in range(4)]
        self._ck(self.lib.kc_survivors_device(self.h, *[C.byref(x) for x in p]), "kc_survivors_device")
        return tuple(x.value for x in p)

    def prefilter(self, sort=True):
        n = C.c_uint64()
        self._ck(self.lib.kc_num_prefilter(self.h, C.byref(n)), "kc_num_prefilter")
        n = n.value
        keys = np.zeros((n, self.n_longs), dtype=np.uint64)
        vals = np.zeros((n, 9), dtype=np.uint16)
        self._ck(self.lib.kc_copy_prefilter(self.h, _ptr(keys), _ptr(vals)), "kc_copy_prefilter")
        if sort and n:
            order = np.lexsort([keys[:, j] for j in range(self.n_longs - 1, -1, -1)])
            keys, vals = keys[order], vals[order]
        return keys, vals

    def stats(self):
        st = KcStats()
        self._ck(self.lib.kc_get_stats(self.h, C.byref(st)), "kc_get_stats")
        return {n: getattr(st, n) for n, _ in KcStats._fields_ if n != "reserved"}

    def close(self):
        if getattr(self, "h", None):
            self.lib.kc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class ResidentReads:
    """kc_reads: PackedRead bytes + offsets uploaded once (asynchronously) and shared by the k rounds of a job. Takes raw HOST
    pointers (pinned memory copies fastest) or numpy arrays, which must stay alive until the first add_reads_resident."""

    def __init__(self, bytes_, offsets, n_reads=None, device=0, batch_reads=0):
        self.lib = load_library()
        if isinstance(bytes_, np.ndarray):
            self._keep = (np.ascontiguousarray(bytes_, dtype=np.uint8), np.ascontiguousarray(offsets, dtype=np.uint64))
            n_reads = len(self._keep[1]) - 1
            bp, op = _ptr(self._keep[0]), _ptr(self._keep[1])
        else:
            bp, op = C.c_void_p(bytes_), C.c_void_p(offsets)
        self.r = C.c_void_p()
        rc = self.lib.kc_reads_create(device, bp, op, n_reads, batch_reads, C.byref(self.r))
        if rc != KC_OK:
            self.close()
            raise KcountError(f"kc_reads_create failed ({ERR_NAMES.get(rc, rc)})")

    @classmethod
    def from_fastq(cls, seqs, quals, offsets, qual_offset=33, device=0, batch_reads=0):
        """kc_reads_create_fastq: concatenated sequence and quality lines (uint8 arrays) are packed into PackedRead bytes on
        the device."""
        self = cls.__new__(cls)
        self.lib = load_library()
        self._keep = (np.ascontiguousarray(seqs, dtype=np.uint8), np.ascontiguousarray(quals, dtype=np.uint8),
                      np.ascontiguousarray(offsets, dtype=np.uint64))
        self.total = int(self._keep[2][-1])
        self.r = C.c_void_p()
        rc = self.lib.kc_reads_create_fastq(device, _ptr(self._keep[0]), _ptr(self._keep[1]), _ptr(self._keep[2]),
                                            len(self._keep[2]) - 1, qual_offset, batch_reads, C.byref(self.r))
        if rc != KC_OK:
            self.close()
            raise KcountError(f"kc_reads_create_fastq failed ({ERR_NAMES.get(rc, rc)})")
        return self

    def packed(self):
        """The PackedRead bytes as they sit in HBM (kc_reads_copy_packed)."""
        out = np.empty(int(self._keep[-1][-1]), dtype=np.uint8)
        rc = self.lib.kc_reads_copy_packed(self.r, _ptr(out))
        if rc != KC_OK:
            raise KcountError(f"kc_reads_copy_packed failed ({ERR_NAMES.get(rc, rc)})")
        return out

    def close(self):
        if self.r:
            self.lib.kc_reads_destroy(self.r)
            self.r = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def _fmix64(x):
    x = x ^ (x >> np.uint64(33))
    x = x * np.uint64(0xff51afd7ed558ccd)
    x = x ^ (x >> np.uint64(33))
    x = x * np.uint64(0xc4ceb9fe1a85ec53)
    return x ^ (x >> np.uint64(33))


def table_digest(keys, counts, left, right):
    """The digest kc_table_digest computes on the device, for a table held in numpy arrays (any row order)."""
    keys = np.asarray(keys, dtype=np.uint64).reshape(len(counts), -1)
    with np.errstate(over="ignore"):
        x = np.full(len(counts), 0x9E3779B97F4A7C15, dtype=np.uint64)
        for j in range(keys.shape[1]):
            x = _fmix64(x ^ keys[:, j])
        c = np.asarray(counts).astype(np.uint64)
        x = _fmix64(x ^ (c | (np.asarray(left).astype(np.uint64) << np.uint64(16)) | (np.asarray(right).astype(np.uint64) << np.uint64(24))))
        sm = int(x.sum(dtype=np.uint64)) if len(x) else 0
    return (len(counts), int(c.sum()), sm, int(np.bitwise_xor.reduce(x)) if len(x) else 0)


DIGEST_STATIC_MAX_ROWS = 1 << 28


def table_digest_static(device, d_keys, d_counts, d_left, d_right, n, n_longs, d_out4, stream=0):
    """kc_table_digest_static: allocation-free digest of a table in DEVICE memory (raw pointers); the result lands in the
    4 x u64 DEVICE buffer d_out4 once the stream has run. Raises KcountError with the library's message on rejection."""
    lib = load_library()
    err = C.create_string_buffer(256)
    rc = lib.kc_table_digest_static(device, C.c_void_p(d_keys), C.c_void_p(d_counts), C.c_void_p(d_left), C.c_void_p(d_right), n,
                                    n_longs, C.c_void_p(d_out4), C.c_void_p(stream), err, 256)
    if rc != KC_OK:
        raise KcountError(f"kc_table_digest_static failed ({ERR_NAMES.get(rc, rc)}): {err.value.decode()}")


def device_count():
    n = C.c_int32()
    rc = load_library().kc_device_count(C.byref(n))
    if rc != KC_OK:
        raise KcountError(f"kc_device_count failed ({ERR_NAMES.get(rc, rc)})")
    return n.value


def finish_local_group(counters):
    """Finish shards 0..n-1 that all live in this process (kc_finish_local_group)."""
    lib = load_library()
    arr = (C.c_void_p * len(counters))(*[c.h for c in counters])
    rc = lib.kc_finish_local_group(arr, len(counters))
    if rc != KC_OK:
        msgs = "; ".join(lib.kc_last_error(c.h).decode() for c in counters)
        raise KcountError(f"kc_finish_local_group failed ({ERR_NAMES.get(rc, rc)}): {msgs}")


def analyze_kmers(kmer_len, read_bytes, read_offsets, dmin_thres=2, ctg_seqs=None, ctg_offsets=None, ctg_depths=None, **kw):
    """Single-shard convenience with the call order of the reference's analyze_kmers<MAX_K> (kcount.cpp:140-157):
    count_kmers -> [add_ctg_kmers] -> finish_updates. Returns (keys, counts, left, right, stats)."""
    with KmerCounter(kmer_len, dmin_thres=dmin_thres, **kw) as kc:
        kc.add_reads(read_bytes, read_offsets)
        if ctg_offsets is not None and len(ctg_offsets) > 1:
            kc.add_contigs(ctg_seqs, ctg_offsets, ctg_depths)
        kc.finish()
        out = kc.survivors()
        pf = kc.prefilter() if kw.get("keep_prefilter") else None
        return out + (kc.stats(), pf)


def prim_kmers(kmers, num_shards=1, device=0):
    """Device primitives on a list of equal-length ASCII k-mers: words, rc words, minimizer, slot hash, owner."""
    lib = load_library()
    k = len(kmers[0])
    n = len(kmers)
    nl = k // 32 + 1
    buf = "".join(kmers).encode()
    w = np.zeros((n, nl), dtype=np.uint64)
    r = np.zeros((n, nl), dtype=np.uint64)
    mn = np.zeros(n, dtype=np.uint64)
    hs = np.zeros(n, dtype=np.uint64)
    ow = np.zeros(n, dtype=np.int32)
    rc = lib.kc_prim_kmers(device, k, buf, n, num_shards, _ptr(w), _ptr(r), _ptr(mn), _ptr(hs), _ptr(ow))
    if rc != KC_OK:
        raise KcountError(f"kc_prim_kmers failed ({rc})")
    return w, r, mn, hs, ow


def prim_get_ext(rows5, dmin_thres=2, dyn_min_depth=0.9, device=0):
    lib = load_library()
    a = np.ascontiguousarray(rows5, dtype=np.uint16)
    out = np.zeros(len(a), dtype=np.uint8)
    rc = lib.kc_prim_get_ext(device, _ptr(a), len(a), dmin_thres, dyn_min_depth, _ptr(out))
    if rc != KC_OK:
        raise KcountError(f"kc_prim_get_ext failed ({rc})")
    return out

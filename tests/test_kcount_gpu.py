"""Parity tests proper: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded inputs.
Bar: bit-exact (identical surviving k-mer set, counts, extension chars; identical pre-filter counters)."""
import numpy as np
import pytest

from mhm2_proxy_b200 import synth

pytestmark = pytest.mark.gpu

KS = [21, 31, 32, 33, 55, 64, 77, 96, 99]


def _oracle(k, rb, ro, **kw):
    from oracle import harness
    kind = "ref" if harness.have_ref() and kw.pop("prefer_ref", False) else "port"
    kw.pop("prefer_ref", None)
    return harness.run(kind, k, rb, ro, **kw)


def _gpu(k, rb, ro, **kw):
    from mhm2_proxy_b200 import kcount
    return kcount.analyze_kmers(k, rb, ro, **kw)


def _assert_same(gpu, ref, what=""):
    keys, counts, left, right = gpu[:4]
    assert len(keys) == len(ref.keys), f"{what}: {len(keys)} survivors vs oracle {len(ref.keys)}"
    assert np.array_equal(keys, ref.keys), f"{what}: surviving k-mer sets differ"
    bad = np.nonzero(counts != ref.counts)[0]
    assert bad.size == 0, f"{what}: {bad.size} counts differ, first at {bad[:3]}: {counts[bad[:3]]} vs {ref.counts[bad[:3]]}"
    bad = np.nonzero((left != ref.left) | (right != ref.right))[0]
    assert bad.size == 0, f"{what}: {bad.size} extension chars differ, first {bad[:3]}"


@pytest.fixture(scope="module")
def community():
    return synth.make_community(150000, n_genomes=3, seed=5)


def test_primitives_match_oracle(built):
    from mhm2_proxy_b200 import kcount
    from oracle import harness
    rng = np.random.default_rng(3)
    for k in [15, 21, 31, 32, 33, 55, 63, 64, 65, 77, 96, 99, 127]:
        kmers = ["".join("ACGT"[c] for c in rng.integers(0, 4, k)) for _ in range(64)]
        kmers += ["A" * k, "T" * k, "ACGT" * (k // 4) + "ACGT"[: k % 4], "N" * k, ("ACGTN" * 30)[:k]]
        w, r, mn, hs, ow = kcount.prim_kmers(kmers, num_shards=7)
        for i, s in enumerate(kmers):
            ew, er, emn, ehs, eow, _ = harness.kmer_prims("port", s, num_shards=7)
            assert np.array_equal(w[i], ew) and np.array_equal(r[i], er), (k, s)
            assert int(mn[i]) == emn, (k, s, hex(int(mn[i])), hex(emn))
            assert int(hs[i]) == ehs and int(ow[i]) == eow, (k, s)


def test_get_ext_matches_oracle(built):
    from mhm2_proxy_b200 import kcount
    from oracle import harness
    rng = np.random.default_rng(4)
    rows = []
    for count in list(range(0, 300)) + [999, 1000, 1001, 9999, 10000, 20000, 30000, 65534, 65535]:
        for _ in range(6):
            e = rng.integers(0, max(2, count // 2 + 2), 4)
            if rng.random() < 0.3:
                e[rng.integers(0, 4)] = e[rng.integers(0, 4)]  # ties
            rows.append(list(np.minimum(e, 65535)) + [count])
    rows = np.asarray(rows, dtype=np.uint16)
    for thres in (1, 2, 3):
        got = kcount.prim_get_ext(rows, dmin_thres=thres)
        exp = np.array([ord(harness.get_ext("port", *[int(x) for x in row], dmin_thres=thres)) for row in rows], dtype=np.uint8)
        assert np.array_equal(got, exp), np.nonzero(got != exp)[0][:5]


@pytest.mark.parametrize("k", KS)
@pytest.mark.parametrize("parts", [1, 16])
def test_reads_parity(built, community, k, parts):
    rb, ro = synth.make_reads(8000, community, error_rate=0.01, seed=100 + k, n_rate=2e-3)
    gpu = _gpu(k, rb, ro, num_partitions=parts, keep_prefilter=True)
    ref = _oracle(k, rb, ro, want_prefilter=True)
    _assert_same(gpu, ref, f"k={k} parts={parts}")
    st, (pk, pv) = gpu[4], gpu[5]
    assert st["instances_inserted"] == ref.stats["instances"] == synth.count_instances(ro, k)
    assert st["distinct"] == ref.stats["distinct_after_reads"]
    assert np.array_equal(pk, ref.pf_keys), "pre-filter key sets differ"
    assert np.array_equal(pv, ref.pf_vals), "pre-filter counters differ"
    assert st["instances_inserted"] == st["instances_parsed"]  # nothing dropped


@pytest.mark.parametrize("k,cap", [(21, 48), (33, 200), (99, 1000)])
def test_subpass_split_parity(built, community, k, cap):
    """A partition with more distinct k-mers than one shared-memory table holds is re-run as hash sub-passes
    (dbg_key_cap forces that path hard: every partition is split many times); results must not change."""
    rb, ro = synth.make_reads(3000, community, error_rate=0.01, seed=31 + k, n_rate=1e-3)
    cs, co, cd = synth.make_contigs(community, k, 30, seed=k + 3, depth_max=6)
    gpu = _gpu(k, rb, ro, num_partitions=3, dbg_key_cap=cap, keep_prefilter=True, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
    ref = _oracle(k, rb, ro, want_prefilter=True, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
    _assert_same(gpu, ref, f"sub-pass k={k}")
    assert gpu[4]["table_retries"] > 0
    assert np.array_equal(gpu[5][0], ref.pf_keys) and np.array_equal(gpu[5][1], ref.pf_vals)


def test_auto_partitions(built, community):
    """num_partitions=0: the library sizes the partitions itself (non-power-of-two counts included)."""
    k = 31
    rb, ro = synth.make_reads(20000, community, error_rate=0.01, seed=77)
    for parts in (0, 5, 37, 1000):
        gpu = _gpu(k, rb, ro, num_partitions=parts)
        _assert_same(gpu, _oracle(k, rb, ro), f"parts={parts}")


@pytest.mark.parametrize("k", [21, 33, 99])
def test_ragged_reads_parity(built, community, k):
    rb, ro = synth.make_reads(6000, community, read_len=180, error_rate=0.02, seed=7 + k, n_rate=1e-3, var_len=True)
    gpu = _gpu(k, rb, ro, num_partitions=4, batch_reads=1500)  # several internal batches
    _assert_same(gpu, _oracle(k, rb, ro), f"ragged k={k}")


@pytest.mark.parametrize("k", [21, 31, 55, 99])
@pytest.mark.parametrize("thres", [2, 3])
def test_contig_pass_parity(built, community, k, thres):
    rb, ro = synth.make_reads(5000, community, error_rate=0.01, seed=300 + k)
    cs, co, cd = synth.make_contigs(community, k, 120, seed=k, max_len=3000, depth_max=6)
    gpu = _gpu(k, rb, ro, dmin_thres=thres, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd, num_partitions=8)
    ref = _oracle(k, rb, ro, dmin_thres=thres, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
    _assert_same(gpu, ref, f"ctg k={k} thres={thres}")
    assert gpu[4]["ctg_instances_inserted"] == ref.stats["ctg_instances"]


def test_contig_order_independence(built, community):
    """kcount_cpu.cpp:356-406 is a sequential state machine; survivors must not depend on contig order."""
    k = 31
    rb, ro = synth.make_reads(3000, community, error_rate=0.01, seed=41)
    cs, co, cd = synth.make_contigs(community, k, 80, seed=9, max_len=2000, depth_max=5, dup_frac=0.3, conflict_frac=0.1)
    base = _oracle(k, rb, ro, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
    rng = np.random.default_rng(0)
    for _ in range(3):
        perm = rng.permutation(len(cd))
        seqs = [cs[int(co[i]):int(co[i + 1])] for i in perm]
        co2 = np.zeros(len(perm) + 1, dtype=np.uint64)
        co2[1:] = np.cumsum([len(s) for s in seqs])
        cs2 = np.concatenate(seqs)
        ref = _oracle(k, rb, ro, ctg_seqs=cs2, ctg_offsets=co2, ctg_depths=cd[perm])
        assert np.array_equal(ref.keys, base.keys) and np.array_equal(ref.counts, base.counts)
        gpu = _gpu(k, rb, ro, ctg_seqs=cs2, ctg_offsets=co2, ctg_depths=cd[perm], num_partitions=4)
        _assert_same(gpu, base, "shuffled contigs")


def test_long_contig_is_tiled(built, community):
    k = 55
    codes = community[0]
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    seq = lut[codes[1000:1000 + 40000]]
    co = np.array([0, len(seq)], dtype=np.uint64)
    cd = np.array([7.0])
    rb, ro = synth.make_reads(500, community, seed=77)
    gpu = _gpu(k, rb, ro, ctg_seqs=seq, ctg_offsets=co, ctg_depths=cd)
    _assert_same(gpu, _oracle(k, rb, ro, ctg_seqs=seq, ctg_offsets=co, ctg_depths=cd), "40 kbp contig")


def test_long_reads_are_tiled(built, community):
    k = 31
    rb, ro = synth.make_reads(60, community, read_len=2500, error_rate=0.01, seed=5, paired=False)
    _assert_same(_gpu(k, rb, ro), _oracle(k, rb, ro), "2.5 kbp reads")


def test_saturation_and_heavy_hitters(built):
    """One k-mer seen > 65535 times: count and extension counters saturate at 65535 (kcount_cpu.cpp:152-155,348-350)."""
    k = 21
    unit = np.array([0, 1, 2, 3, 1], dtype=np.uint8)  # period 5: five distinct k-mers, ~25 instances each per read
    n = 3600
    reads = np.tile(np.tile(unit, 30)[:150], (n, 1))
    rb = (reads | (31 << 3)).astype(np.uint8).reshape(-1)
    ro = np.arange(n + 1, dtype=np.uint64) * 150
    gpu = _gpu(k, rb, ro, keep_prefilter=True, num_partitions=2)
    ref = _oracle(k, rb, ro, want_prefilter=True)
    assert ref.pf_vals.max() == 65535
    _assert_same(gpu, ref, "saturation")
    assert np.array_equal(gpu[5][1], ref.pf_vals)
    assert gpu[4]["table_retries"] > 0  # the u16 counters wrapped: those work items were re-run with u32 counters


def test_skewed_community_c5(built):
    """BASELINE config C5 in small: Zipf(1.2) abundances, reads from a short repeat and poly-A inserts (heavy-hitter
    k-mers: a few partitions far above the mean, the 'heavy' schedule and wide counters), 1 % errors."""
    k = 31
    com = synth.make_community(200000, n_genomes=50, seed=55, zipf=1.2)
    rb, ro = synth.make_reads(30000, com, error_rate=0.01, seed=56, n_rate=1e-3)
    rep = synth.make_community(2000, n_genomes=1, seed=57)
    rb2, ro2 = synth.make_reads(6000, rep, error_rate=0.01, seed=58)
    polya = np.full((800, 150), 0 | (31 << 3), dtype=np.uint8)  # poly-A reads: one k-mer, 118 instances per read
    polya[:, 70] = np.arange(800) % 4 | (31 << 3)               # one variable base: a handful of neighbours
    rb3 = polya.reshape(-1)
    rb_all = np.concatenate([rb, rb2, rb3])
    ro_all = np.concatenate([ro, ro2[1:] + ro[-1], np.arange(1, 801, dtype=np.uint64) * 150 + ro[-1] + ro2[-1]])
    ref = _oracle(k, rb_all, ro_all, want_prefilter=True)
    assert ref.pf_vals[:, 0].max() == 65535, "the input is meant to saturate a counter"
    for parts in (0, 64):
        gpu = _gpu(k, rb_all, ro_all, num_partitions=parts, keep_prefilter=True)
        _assert_same(gpu, ref, f"C5 parts={parts}")
        assert np.array_equal(gpu[5][0], ref.pf_keys) and np.array_equal(gpu[5][1], ref.pf_vals)
        assert gpu[4]["instances_inserted"] == gpu[4]["instances_parsed"]  # nothing dropped


@pytest.mark.parametrize("name", ["kcount_k21_reads", "kcount_k31_ctgs", "kcount_k33_reads", "kcount_k55_ctgs", "kcount_k99_ctgs"])
def test_against_committed_golden_tables(built, name):
    """The CUDA path against tests/golden/ (tables written by the reference's own kcount_cpu.cpp build, make_golden.py):
    no oracle in the loop."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", name + ".npz"))
    kw = {}
    if "ctg_seqs" in g:
        kw = dict(ctg_seqs=g["ctg_seqs"], ctg_offsets=g["ctg_offsets"], ctg_depths=g["ctg_depths"])
    keys, counts, left, right, st, pf = _gpu(int(g["k"]), g["read_bytes"], g["read_offsets"], keep_prefilter=True, **kw)
    assert np.array_equal(keys, g["keys"]) and np.array_equal(counts, g["counts"])
    assert np.array_equal(left, g["left"]) and np.array_equal(right, g["right"])
    assert np.array_equal(pf[0], g["pf_keys"]) and np.array_equal(pf[1], g["pf_vals"])


def test_edge_cases(built, community):
    from mhm2_proxy_b200 import kcount
    k = 31
    # no reads at all
    with kcount.KmerCounter(k) as kc:
        kc.finish()
        assert kc.num_survivors() == 0
    # reads shorter than k+2 contribute nothing (kcount.cpp:79, kcount_cpu.cpp:84-102)
    rb, ro = synth.make_reads(50, community, read_len=k + 1, seed=3)
    out = _gpu(k, rb, ro)
    assert len(out[0]) == 0 and out[4]["instances_inserted"] == 0
    # exactly k+2 bases: one instance per read; duplicates make it survive
    rb, ro = synth.make_reads(1, community, read_len=k + 2, seed=4, error_rate=0, lowq_tail_p=0)
    rb3, ro3 = np.tile(rb, 3), np.arange(4, dtype=np.uint64) * (k + 2)
    _assert_same(_gpu(k, rb3, ro3), _oracle(k, rb3, ro3), "k+2 reads")
    # the unpaired-read dummy mate "N" (merge_reads.cpp:313-314) is skipped
    rbn = np.concatenate([rb3, np.array([4 | (31 << 3)], dtype=np.uint8)])
    ron = np.concatenate([ro3, np.array([ro3[-1] + 1], dtype=np.uint64)])
    _assert_same(_gpu(k, rbn, ron), _oracle(k, rb3, ro3), "dummy N mate")
    # errors are reported, not thrown across the ABI as crashes
    with pytest.raises(kcount.KcountError):
        kcount.KmerCounter(0)
    with kcount.KmerCounter(k) as kc:
        bad = np.frombuffer(b"ACGTXACGT" * 10, dtype=np.uint8)
        with pytest.raises(kcount.KcountError):
            kc.add_contigs(bad, np.array([0, len(bad)], dtype=np.uint64), np.array([3.0]))


@pytest.mark.skipif(not __import__("oracle.harness", fromlist=["x"]).have_ref(), reason="reference build not present")
@pytest.mark.parametrize("k", [21, 33, 99])
def test_against_reference_build(built, community, k):
    """Same comparison against the reference's own kcount_cpu.cpp compiled from /root/reference (oracle/_ref)."""
    rb, ro = synth.make_reads(6000, community, error_rate=0.01, seed=900 + k, n_rate=1e-3)
    cs, co, cd = synth.make_contigs(community, k, 40, seed=k + 1)
    gpu = _gpu(k, rb, ro, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd, num_partitions=8)
    ref = _oracle(k, rb, ro, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd, prefer_ref=True)
    assert ref.stats["dropped"] == 0 and ref.stats["singleton_overrides"] == 0
    _assert_same(gpu, ref, f"reference build k={k}")


@pytest.mark.skipif(not __import__("oracle.harness", fromlist=["x"]).have_dropin(), reason="drop-in build not present")
@pytest.mark.parametrize("k,thres", [(21, 2), (33, 3), (55, 2), (99, 2)])
def test_dropin_through_reference_analyze_kmers(built, community, k, thres):
    """The drop-in boundary itself: the reference's OWN analyze_kmers<MAX_K> (kcount.cpp, compiled unmodified) driving
    integration/kcount_b200.cpp (SeqBlockInserter / HashTableInserter over the C ABI) must leave in kmer_dht->local_kmers
    exactly what the reference's CPU backend leaves there."""
    from oracle import harness
    rb, ro = synth.make_reads(5000, community, error_rate=0.01, seed=700 + k, n_rate=1e-3)
    cs, co, cd = synth.make_contigs(community, k, 40, seed=k + 11, depth_max=6)
    got = harness.run_dropin(k, rb, ro, dmin_thres=thres, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
    ref = _oracle(k, rb, ro, dmin_thres=thres, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd, prefer_ref=True)
    _assert_same((got.keys, got.counts, got.left, got.right), ref, f"drop-in k={k}")
    # reads only (no contig pass at all: fut_has_contigs is false, kcount.cpp:149)
    got = harness.run_dropin(k, rb, ro, dmin_thres=thres)
    _assert_same((got.keys, got.counts, got.left, got.right), _oracle(k, rb, ro, dmin_thres=thres), f"drop-in reads only k={k}")


def test_reads_ascii_equals_packed(built, community):
    """kc_add_reads_ascii (what process_seq receives: ASCII, lower case = low quality) == kc_add_reads (PackedRead bytes)."""
    from mhm2_proxy_b200 import kcount
    k = 31
    rb, ro = synth.make_reads(4000, community, error_rate=0.01, seed=9, n_rate=2e-3, var_len=True)
    lut = np.frombuffer(b"ACGTNNNN", dtype=np.uint8)
    asc = lut[rb & 7]
    asc = np.where((rb >> 3) < 20, asc | 0x20, asc).astype(np.uint8)
    with kcount.KmerCounter(k, num_partitions=4) as kc:
        kc.add_reads_ascii(asc, ro)
        kc.set_dmin_thres(2)
        kc.finish()
        got = kc.survivors()
    _assert_same(got, _oracle(k, rb, ro), "ascii reads")


def test_fastq_text_is_packed_on_the_device(built, community):
    """kc_reads_create_fastq == PackedRead's constructor byte for byte (incl. IUPAC codes, quality clamp at 31 and the
    unsigned-char wrap below the offset), at ragged batch boundaries; the table from it == the table from packed input;
    an illegal character is reported (the reference DIEs)."""
    from mhm2_proxy_b200 import kcount
    k = 31
    rb, ro = synth.make_reads(5000, community, error_rate=0.01, seed=21, n_rate=2e-3, var_len=True)
    rng = np.random.default_rng(3)
    iupac = np.frombuffer(b"NURYKMSWBDHV", dtype=np.uint8)
    seq = np.frombuffer(b"ACGTN", dtype=np.uint8)[rb & 7].copy()
    isn = (rb & 7) == 4
    seq[isn] = iupac[rng.integers(0, len(iupac), int(isn.sum()))]
    qual = ((rb >> 3).astype(np.int64) + 33)
    odd = rng.random(len(rb)) < 0.01  # qualities outside 0..31 after the offset
    qual[odd] = rng.integers(33 - 3, 33 + 60, int(odd.sum()))
    qual = qual.astype(np.uint8)
    from oracle import harness
    if not harness.have_ref():
        pytest.skip("reference build (oracle/_ref: packed_reads.cpp) not present")
    want = harness.pack_reads_ref(seq, qual, ro, 33)  # the reference's own PackedRead constructor
    with kcount.ResidentReads.from_fastq(seq, qual, ro, qual_offset=33, batch_reads=777) as rr:
        got = rr.packed()
        assert np.array_equal(got, want)
        with kcount.KmerCounter(k, num_partitions=6) as kc:
            kc.add_reads_resident(rr)
            kc.finish()
            _assert_same(kc.survivors(), _oracle(k, want, ro), "reads packed on the device")
    bad = seq.copy()
    bad[len(bad) // 2] = ord("x")
    with kcount.ResidentReads.from_fastq(bad, qual, ro, qual_offset=33) as rr, kcount.KmerCounter(k, num_partitions=6) as kc:
        with pytest.raises(kcount.KcountError, match="BADCHAR"):
            kc.add_reads_resident(rr)


def test_fastq_packing_against_committed_reference_bytes(built):
    """tests/golden/packed_reads.npz: bytes written by the reference's PackedRead constructor (make_golden.py); no oracle in
    the loop."""
    import os
    from mhm2_proxy_b200 import kcount
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "packed_reads.npz"))
    for batch in (0, 37):
        with kcount.ResidentReads.from_fastq(g["seqs"], g["quals"], g["offsets"], qual_offset=int(g["qual_offset"]), batch_reads=batch) as rr:
            assert np.array_equal(rr.packed(), g["packed"])


def test_table_digest_matches_numpy_and_reference(built, community):
    """kc_table_digest (device) == the numpy digest of the copied table == the digest of the reference's table."""
    from mhm2_proxy_b200 import kcount
    for k in (21, 55, 99):
        rb, ro = synth.make_reads(4000, community, error_rate=0.01, seed=900 + k)
        with kcount.KmerCounter(k) as kc:
            kc.add_reads(rb, ro)
            kc.finish()
            dg = kc.digest()
            tab = kc.survivors()
        assert dg == kcount.table_digest(*tab)
        ref = _oracle(k, rb, ro, prefer_ref=True)
        assert dg == kcount.table_digest(ref.keys, ref.counts, ref.left, ref.right)
        assert dg[0] == len(ref.keys) and dg[1] == int(ref.counts.astype(np.int64).sum())


def test_dump_kmers_matches_reference_format(built, community, tmp_path):
    """kc_dump_kmers writes what KmerDHT::dump_kmers writes (kmer_dht.cpp:238-266): "KMER count left right" per survivor,
    gzip when the name ends in .gz; compared as a set of lines with the oracle's table."""
    import gzip
    from mhm2_proxy_b200 import kcount
    for k in (21, 55):
        rb, ro = synth.make_reads(3000, community, error_rate=0.01, seed=50 + k)
        ref = _oracle(k, rb, ro)
        want = set()
        for row, c, le, ri in zip(ref.keys, ref.counts, ref.left, ref.right):
            kmer = "".join("ACGT"[(int(row[b >> 5]) >> (2 * (31 - (b & 31)))) & 3] for b in range(k))
            want.add(f"{kmer} {int(c)} {chr(le)} {chr(ri)}")
        with kcount.KmerCounter(k, num_partitions=3) as kc:
            kc.add_reads(rb, ro)
            kc.finish()
            kc.dump_kmers(tmp_path / f"kmers-{k}.txt.gz")
            kc.dump_kmers(tmp_path / f"kmers-{k}.txt")
        with gzip.open(tmp_path / f"kmers-{k}.txt.gz", "rt") as f:
            got_gz = f.read().splitlines()
        got = (tmp_path / f"kmers-{k}.txt").read_text().splitlines()
        assert len(got) == len(want) == len(got_gz)
        assert set(got) == want and set(got_gz) == want


def test_resident_reads_shared_by_rounds(built, community):
    """kc_reads_create uploads the PackedRead bytes once; two k rounds (and a round with long, tiled reads) parse them from
    HBM and give the tables kc_add_reads gives."""
    from mhm2_proxy_b200 import kcount
    rb, ro = synth.make_reads(9000, community, error_rate=0.01, seed=11, n_rate=1e-3, var_len=True)
    with kcount.ResidentReads(rb, ro, batch_reads=2500) as rr:
        for k in (21, 55):
            with kcount.KmerCounter(k, num_partitions=0) as kc:
                kc.add_reads_resident(rr)
                kc.finish()
                _assert_same(kc.survivors(), _oracle(k, rb, ro), f"resident reads k={k}")
    lb, lo = synth.make_reads(300, community, read_len=900, error_rate=0.01, seed=12)  # longer than one parse tile
    with kcount.ResidentReads(lb, lo, batch_reads=100) as rr, kcount.KmerCounter(33, num_partitions=5) as kc:
        kc.add_reads_resident(rr)
        kc.finish()
        _assert_same(kc.survivors(), _oracle(33, lb, lo), "resident long reads")


@pytest.mark.parametrize("k,parts", [(21, 4096), (33, 3000), (55, 8)])
def test_finish_to_host_equals_finish_plus_copy(built, community, k, parts):
    """kc_finish_to_host (count stage in slices, survivors copied to the host while the next slice is counted) delivers
    exactly the table of kc_finish + kc_copy_survivors; with too little room it reports OVERFLOW and the table stays
    available on the device."""
    from mhm2_proxy_b200 import kcount
    rb, ro = synth.make_reads(12000, community, error_rate=0.01, seed=300 + k, n_rate=1e-3)
    ref = _oracle(k, rb, ro)
    nl = k // 32 + 1
    cap = len(ref.counts) + 100
    keys = np.zeros((cap, nl), dtype=np.uint64)
    counts = np.zeros(cap, dtype=np.uint16)
    left = np.zeros(cap, dtype=np.uint8)
    right = np.zeros(cap, dtype=np.uint8)
    with kcount.KmerCounter(k, num_partitions=parts) as kc:
        kc.add_reads(rb, ro)
        n = kc.finish_into(keys.ctypes.data, counts.ctypes.data, left.ctypes.data, right.ctypes.data, cap)
        assert n == len(ref.counts)
        assert kc.num_survivors() == n
    order = np.lexsort(tuple(keys[:n, i] for i in range(nl - 1, -1, -1)))
    _assert_same((keys[:n][order], counts[:n][order], left[:n][order], right[:n][order]), ref, f"finish_to_host k={k}")
    small = max(1, n // 3)
    with kcount.KmerCounter(k, num_partitions=parts) as kc:
        kc.add_reads(rb, ro)
        with pytest.raises(kcount.KcountError, match="OVERFLOW"):
            kc.finish_into(keys.ctypes.data, counts.ctypes.data, left.ctypes.data, right.ctypes.data, small)
        _assert_same(kc.survivors(), ref, f"table after a short host buffer k={k}")


@pytest.mark.parametrize("n_shards,k", [(2, 21), (3, 33), (8, 31), (4, 99)])
def test_sharded_table_matches_oracle_shards(built, community, n_shards, k):
    """N shards in one process on one GPU (local transport): reads split across shards, supermers routed to the owner
    quick_hash(minimizer) % N (kmer_dht.cpp:192-196). Shard i must hold exactly the oracle's survivors with owner i."""
    from mhm2_proxy_b200 import kcount
    rb, ro = synth.make_reads(6000, community, error_rate=0.01, seed=500 + k, n_rate=1e-3)
    cs, co, cd = synth.make_contigs(community, k, 40, seed=k + 7, depth_max=6)
    ref = _oracle(k, rb, ro, num_shards=n_shards, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
    n = len(ro) - 1
    counters = [kcount.KmerCounter(k, num_shards=n_shards, shard_id=i, num_partitions=4, batch_reads=700) for i in range(n_shards)]
    try:
        for i, kc in enumerate(counters):
            r0, r1 = n * i // n_shards, n * (i + 1) // n_shards
            kc.add_reads(rb[int(ro[r0]):int(ro[r1])], ro[r0:r1 + 1] - ro[r0])
            c0, c1 = len(cd) * i // n_shards, len(cd) * (i + 1) // n_shards
            if c1 > c0:
                kc.add_contigs(cs[int(co[c0]):int(co[c1])], co[c0:c1 + 1] - co[c0], cd[c0:c1])
        kcount.finish_local_group(counters)
        total = 0
        for i, kc in enumerate(counters):
            keys, counts, left, right = kc.survivors()
            sel = ref.shard == i
            assert np.array_equal(keys, ref.keys[sel]), f"shard {i}: key set differs ({len(keys)} vs {sel.sum()})"
            assert np.array_equal(counts, ref.counts[sel]) and np.array_equal(left, ref.left[sel]) and np.array_equal(right, ref.right[sel])
            total += len(keys)
            st = kc.stats()
            assert st["supermer_bytes_sent"] > 0
        assert total == len(ref.keys)
    finally:
        for kc in counters:
            kc.close()


def test_nccl_exchange_two_gpus(built):
    """Real NCCL transport, one process per GPU under torchrun (needs >= 2 GPUs: `gpurun --gpus 2`)."""
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "tests", "nccl_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "NCCL-PARITY-OK" in out.stdout

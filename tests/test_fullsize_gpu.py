"""Size-independent properties at BASELINE.json's full C2 size (13.33 M reads, 2 Gbase), where the CPU oracle would take
minutes: the table must not depend on batching, partitioning, sharding or on how the reads reached the device, and must
satisfy the invariants of the reference's output (kcount_cpu.cpp:490-528)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]

N_READS = 13_333_334
K = 21


def _revcomp21(keys):
    """reverse complement of left-aligned 2-bit k-mers (k = 21) held in u64, numpy restatement of kmer.cpp:485-505"""
    x = keys.copy()
    m1, m2 = np.uint64(0x3333333333333333), np.uint64(0x0F0F0F0F0F0F0F0F)
    x = ((x >> np.uint64(2)) & m1) | ((x & m1) << np.uint64(2))
    x = ((x >> np.uint64(4)) & m2) | ((x & m2) << np.uint64(4))
    x = x.byteswap()  # 2-bit groups now in reverse order over all 32 positions
    x = ~x
    return x << np.uint64(2 * (32 - K))


def _sorted_table(keys, counts, left, right):
    o = np.argsort(keys[:, 0], kind="stable")
    return keys[o, 0], counts[o], left[o], right[o]


@pytest.fixture(scope="module")
def reads(built):
    import bench
    tb, to = bench.gen_reads(N_READS, seed=0xB200_0002)
    return tb.numpy(), to.numpy().astype(np.uint64)


def test_full_size_invariants_and_invariance(built, reads):
    from mhm2_proxy_b200 import kcount
    rb, ro = reads
    lens = np.diff(ro.astype(np.int64))
    expect_inst = int(np.maximum(lens - K - 1, 0).sum())

    with kcount.KmerCounter(K, expected_instances=expect_inst) as kc:
        kc.add_reads(rb, ro)
        kc.finish()
        st = kc.stats()
        a = _sorted_table(*kc.survivors(sort=False))
    assert st["instances_parsed"] == expect_inst == st["instances_inserted"]
    assert st["table_retries"] == 0
    assert st["survivors"] == len(a[0]) and st["distinct"] == st["survivors"] + st["purged"]
    keys, counts, left, right = a
    assert np.all(keys[1:] != keys[:-1]), "a k-mer appears twice in the table"
    assert counts.min() >= 2, "depth filter (kcount_cpu.cpp:497)"
    for ext in (left, right):
        assert np.isin(ext, np.frombuffer(b"ACGTFX", dtype=np.uint8)).all()
    assert not np.any((left == ord("X")) & (right == ord("X"))), "X/X k-mers are purged (kcount_cpu.cpp:519)"
    assert np.all(keys <= _revcomp21(keys)), "k-mers are canonical (kcount_cpu.cpp:326-331)"
    assert np.all(keys & np.uint64((1 << (64 - 2 * K)) - 1) == 0), "bits past k are zero (kmer.hpp:61-75)"

    # other batching, explicit partition count, reads resident in HBM: same table
    with kcount.ResidentReads(rb, ro, batch_reads=3_000_000) as rr, kcount.KmerCounter(K, num_partitions=50_000) as kc:
        kc.add_reads_resident(rr)
        kc.finish()
        b = _sorted_table(*kc.survivors(sort=False))
    for x, y, what in zip(a, b, ("k-mers", "counts", "left", "right")):
        assert np.array_equal(x, y), f"{what} depend on batching / partitioning"

    # two shards (quick_hash(minimizer) % 2, kmer_dht.cpp:192-196), each parsing half of the reads: disjoint, union == table
    n = len(ro) - 1
    half = n // 2
    shards = [kcount.KmerCounter(K, num_shards=2, shard_id=i, expected_instances=expect_inst // 2) for i in range(2)]
    try:
        for i, (r0, r1) in enumerate(((0, half), (half, n))):
            shards[i].add_reads(rb[int(ro[r0]):int(ro[r1])], ro[r0:r1 + 1] - ro[r0])
        kcount.finish_local_group(shards)
        parts = [s.survivors(sort=False) for s in shards]
    finally:
        for s in shards:
            s.close()
    assert 0.45 < len(parts[0][1]) / len(keys) < 0.55, "shards are balanced"
    c = _sorted_table(*[np.concatenate([p[j] for p in parts]) for j in range(4)])
    for x, y, what in zip(a, c, ("k-mers", "counts", "left", "right")):
        assert np.array_equal(x, y), f"{what} depend on sharding"


@pytest.mark.timeout(1500)
def test_c1_full_size_bit_exact_against_single_rank_reference(built):
    """BASELINE config C1 at its stated size: arcticsynth-shaped community (32 genomes / 100 Mbp, log-normal abundances,
    1 % error), 2 M reads, k = 31; the whole table bit-exact against the reference's kcount_cpu.cpp run as ONE rank
    (kcount_cpu.cpp:73-103,337-354,490-528), which must itself report no dropped insert and no singleton override."""
    import bench
    from mhm2_proxy_b200 import kcount
    from oracle import harness
    if not harness.have_ref():
        pytest.skip("reference build (oracle/_ref) not present")
    wl = bench.Workload("C1", 1, 0, pinned=False)
    rb, ro = wl.tb.numpy(), wl.to.numpy().astype(np.uint64)
    k = wl.ks[0]
    inst = int(np.maximum(np.diff(ro.astype(np.int64)) - k - 1, 0).sum())
    # 3x coverage: most k-mers are distinct. num_elems = 0.8 x instances -> the reference allocates 3x that (load ~ 0.3)
    ref = harness.run("ref", k, rb, ro, num_elems=int(0.8 * inst))
    # the reference logs seq.length() - 2 - k "kmers" per read (kcount_cpu.cpp:102) although its loop inserts one more
    assert ref.stats["dropped"] == 0 and ref.stats["singleton_overrides"] == 0 and ref.stats["instances"] == inst - wl.n_reads
    with kcount.KmerCounter(k, expected_instances=inst) as kc:
        kc.add_reads(rb, ro)
        kc.finish()
        st = kc.stats()
        keys, counts, left, right = kc.survivors()
        dg = kc.digest()
    assert st["instances_parsed"] == st["instances_inserted"] == inst
    assert st["distinct"] == ref.stats["distinct_after_reads"]
    assert len(keys) == len(ref.keys)
    assert np.array_equal(keys, ref.keys), "surviving k-mer sets differ"
    assert np.array_equal(counts, ref.counts) and np.array_equal(left, ref.left) and np.array_equal(right, ref.right)
    assert dg == kcount.table_digest(ref.keys, ref.counts, ref.left, ref.right)

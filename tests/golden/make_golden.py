"""Generates tests/golden/*.npz and kats.json from the REFERENCE's own code (oracle/_ref/libkcount_ref.so, compiled from
/root/reference by oracle/Makefile). Run in the authoring container only:  python tests/golden/make_golden.py
The fixtures pin the C restatement (oracle/kcount_oracle.c) and the CUDA path to the reference's behaviour."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from mhm2_proxy_b200 import synth  # noqa: E402
from oracle import harness as H  # noqa: E402

# the strings of the reference's own unit test (test/kmer-test.cpp:11-27,251-252 are generated the same way: poly-X,
# repeats and a random read); primitives have no expected values there, so we freeze the reference's outputs here
STRINGS = {
    "As": "A" * 130, "Cs": "C" * 130, "Gs": "G" * 130, "Ts": "T" * 130,
    "ACGTs": "ACGT" * 33, "TCGAs": "TCGA" * 33, "CAGTs": "CAGT" * 33,
}


def main():
    assert H.have_ref(), "reference build missing: make -C oracle ref"
    rng = np.random.default_rng(20261104)
    STRINGS["RandomRead"] = "".join("ACGT"[c] for c in rng.integers(0, 4, 150))
    STRINGS["WithN"] = STRINGS["RandomRead"][:40] + "N" + STRINGS["RandomRead"][41:100] + "NN" + STRINGS["RandomRead"][102:]
    kats = {"strings": STRINGS, "prims": [], "quick_hash": [], "murmur": [], "get_ext": []}
    for k in (15, 21, 31, 32, 33, 55, 63, 64, 77, 95, 96, 99, 127):
        for name, s in STRINGS.items():
            for off in (0, 1, 17):
                if off + k > len(s):
                    continue
                km = s[off:off + k]
                w, r, mn, hs, ow, m = H.kmer_prims("ref", km, num_shards=8)
                kats["prims"].append({"k": k, "str": name, "off": off, "words": [int(x) for x in w], "rc": [int(x) for x in r],
                                      "minimizer": mn, "hash": hs, "owner8": ow, "m": m})
    for v in [0, 1, 2, 0xFFFFFFFFFFFFFFFF, 0x0123456789ABCDEF] + [int(x) for x in rng.integers(0, 2 ** 63, 20)]:
        kats["quick_hash"].append([v, H.quick_hash("ref", v)])
    for nw in (1, 2, 3, 4):
        for _ in range(6):
            words = [int(x) for x in rng.integers(0, 2 ** 63, nw)]
            kats["murmur"].append([words, H.murmur64("ref", np.array(words, dtype=np.uint64))])
    for count in list(range(0, 64)) + [99, 100, 101, 1000, 9999, 10000, 30000, 65535]:
        for _ in range(4):
            e = [int(x) for x in np.minimum(rng.integers(0, max(2, count // 2 + 2), 4), 65535)]
            for thres in (1, 2, 3):
                kats["get_ext"].append([e, count, thres, H.get_ext("ref", *e, count, dmin_thres=thres)])
    json.dump(kats, open(os.path.join(HERE, "kats.json"), "w"))

    com = synth.make_community(30000, n_genomes=2, seed=21)
    for k in (21, 31, 33, 55, 99):
        rb, ro = synth.make_reads(1200, com, error_rate=0.01, seed=1000 + k, n_rate=2e-3)
        cs, co, cd = synth.make_contigs(com, k, 30, seed=2000 + k, max_len=1500, depth_max=6)
        for tag, ctg in (("reads", False), ("ctgs", True)):
            kw = dict(ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd) if ctg else {}
            res = H.run("ref", k, rb, ro, num_shards=8, want_prefilter=True, **kw)
            assert res.stats["dropped"] == 0 and res.stats["singleton_overrides"] == 0
            out = dict(k=k, read_bytes=rb, read_offsets=ro, keys=res.keys, counts=res.counts, left=res.left, right=res.right,
                       shard8=res.shard, pf_keys=res.pf_keys, pf_vals=res.pf_vals)
            if ctg:
                out.update(ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
            np.savez_compressed(os.path.join(HERE, f"kcount_k{k}_{tag}.npz"), **out)
            print(k, tag, len(res.counts), "survivors", len(res.pf_vals), "distinct")


def packed_reads_golden():
    """FASTQ text -> PackedRead bytes through the reference's own constructor (packed_reads.cpp:73-109, compiled unmodified
    into oracle/_ref): pins kc_reads_create_fastq. Includes IUPAC codes, qualities above the 31 clamp and
    below the offset (the unsigned-char wrap), ragged lengths."""
    com = synth.make_community(20000, n_genomes=1, seed=77)
    rb, ro = synth.make_reads(400, com, error_rate=0.01, seed=78, n_rate=5e-3, var_len=True)
    rng = np.random.default_rng(79)
    iupac = np.frombuffer(b"NURYKMSWBDHV", dtype=np.uint8)
    seq = np.frombuffer(b"ACGTN", dtype=np.uint8)[rb & 7].copy()
    isn = (rb & 7) == 4
    seq[isn] = iupac[rng.integers(0, len(iupac), int(isn.sum()))]
    qual = (rb >> 3).astype(np.int64) + 33
    odd = rng.random(len(rb)) < 0.02
    qual[odd] = rng.integers(33 - 3, 33 + 60, int(odd.sum()))
    qual = qual.astype(np.uint8)
    packed = H.pack_reads_ref(seq, qual, ro, 33)
    np.savez_compressed(os.path.join(HERE, "packed_reads.npz"), seqs=seq, quals=qual, offsets=ro, packed=packed, qual_offset=33)
    print("packed_reads", len(ro) - 1, "reads", len(packed), "bytes")


if __name__ == "__main__":
    if "--packed-only" not in sys.argv:
        main()
    packed_reads_golden()

"""Seeded synthetic inputs in the reference's own input format.

Reads are produced directly as PackedRead bytes (reference src/packed_reads.cpp:73-109: base code A0 C1 G2 T3 N4 in the
low 3 bits, min(qual - offset, 31) in the high 5 bits) plus an offsets array, i.e. the flattened form of
``vector<PackedRead>``. Contigs are ASCII + offsets + depths (reference src/contigs.hpp:57-66).
The same bytes feed the CUDA path and the CPU oracle.
"""
import numpy as np

HIGH_Q = 36  # Phred of a good base
LOW_Q = 15  # Phred of a low-quality tail base (< KCOUNT_QUAL_CUTOFF = 20)


def make_community(total_len, n_genomes=1, sigma=1.0, seed=0, zipf=None):
    """Random genomes (codes 0..3) concatenated; returns (codes, starts[n+1], abundance weights)."""
    rng = np.random.default_rng(seed)
    n_genomes = max(1, int(n_genomes))
    glen = max(1000, total_len // n_genomes)
    starts = np.arange(n_genomes + 1, dtype=np.int64) * glen
    codes = rng.integers(0, 4, size=int(starts[-1]), dtype=np.uint8)
    if zipf:
        w = 1.0 / np.arange(1, n_genomes + 1) ** float(zipf)
    elif sigma > 0 and n_genomes > 1:
        w = rng.lognormal(0.0, sigma, n_genomes)
    else:
        w = np.ones(n_genomes)
    return codes, starts, w / w.sum()


def make_reads(n_reads, community, read_len=150, error_rate=0.01, seed=1, n_rate=0.0, lowq_tail_p=0.3, lowq_tail=10,
               paired=True, var_len=False, chunk=1 << 19, out=None):
    """PackedRead bytes + offsets (uint64[n+1]) for ``n_reads`` reads sampled from ``community``.

    Pairs: mate 1 forward at pos, mate 2 the reverse complement ending at pos + insert (insert ~ N(350, 35)).
    Substitution errors uniform at ``error_rate``; with probability ``lowq_tail_p`` the last ``lowq_tail`` bases of a
    read get Phred 15 (exercises the quality mask); ``n_rate`` turns bases into N (code 4).
    ``var_len`` draws lengths in [read_len//3, read_len] for a ragged test input.
    """
    codes, starts, w = community
    rng = np.random.default_rng(seed)
    n_reads = int(n_reads)
    if var_len:
        lens = rng.integers(max(1, read_len // 3), read_len + 1, size=n_reads).astype(np.int64)
    else:
        lens = np.full(n_reads, read_len, dtype=np.int64)
    offsets = np.zeros(n_reads + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum(lens).astype(np.uint64)
    total = int(offsets[-1])
    data = out if out is not None else np.empty(total, dtype=np.uint8)
    glens = np.diff(starts)
    L = read_len
    for c0 in range(0, n_reads, chunk):
        c1 = min(n_reads, c0 + chunk)
        n = c1 - c0
        g = rng.choice(len(w), size=n, p=w) if len(w) > 1 else np.zeros(n, dtype=np.int64)
        ins = np.clip(rng.normal(350, 35, n).astype(np.int64), L, None) if paired else np.full(n, L, dtype=np.int64)
        span = np.maximum(ins, L)
        maxpos = np.maximum(glens[g] - span, 1)
        pos = (rng.random(n) * maxpos).astype(np.int64) + starts[g]
        is_m2 = (np.arange(c0, c1) & 1).astype(bool) if paired else np.zeros(n, dtype=bool)
        if paired:  # a pair shares genome, position and insert size
            pos[1::2] = pos[0::2][: len(pos[1::2])]
            ins[1::2] = ins[0::2][: len(ins[1::2])]
        begin = np.where(is_m2, pos + ins - L, pos)
        begin = np.clip(begin, 0, len(codes) - L)
        idx = begin[:, None] + np.arange(L, dtype=np.int64)[None, :]
        b = codes[idx]
        b[is_m2] = (3 - b[is_m2])[:, ::-1]
        if error_rate > 0:
            err = rng.random((n, L)) < error_rate
            sub = rng.integers(1, 4, size=(n, L), dtype=np.uint8)
            b = np.where(err, (b + sub) & 3, b).astype(np.uint8)
        if n_rate > 0:
            b = np.where(rng.random((n, L)) < n_rate, np.uint8(4), b).astype(np.uint8)
        q = np.full((n, L), min(HIGH_Q, 31), dtype=np.uint8)
        if lowq_tail_p > 0:
            tail = rng.random(n) < lowq_tail_p
            q[tail, L - lowq_tail:] = LOW_Q
        packed = (b | (q << 3)).astype(np.uint8)
        if var_len:
            ln = lens[c0:c1]
            mask = np.arange(L)[None, :] < ln[:, None]
            data[int(offsets[c0]):int(offsets[c1])] = packed[mask]
        else:
            data[int(offsets[c0]):int(offsets[c1])] = packed.reshape(-1)
    return data, offsets


def make_contigs(community, k, n_ctgs, seed=2, min_len=None, max_len=5000, depth_max=200, conflict_frac=0.02,
                 dup_frac=0.1, revcomp_frac=0.3):
    """Previous-round-style contigs cut from the true genomes: ASCII bytes, offsets, float64 depths.

    Includes exact duplicates / overlaps (same k-mers from several contigs), both strands, a few mutated copies whose
    k-mers disagree on an extension, and depths 0..depth_max (0 and 1 exercise the purge rules of
    kcount_cpu.cpp:356-406).
    """
    codes, starts, _ = community
    rng = np.random.default_rng(seed)
    min_len = min_len or (k + 2)
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    seqs, depths = [], []
    for i in range(int(n_ctgs)):
        if seqs and rng.random() < dup_frac:
            s = seqs[int(rng.integers(0, len(seqs)))].copy()
            if rng.random() < 0.5 and len(s) > k + 4:  # overlapping piece rather than full duplicate
                a = int(rng.integers(0, len(s) - k - 3))
                s = s[a:]
        else:
            ln = int(rng.integers(min_len, max_len + 1))
            p = int(rng.integers(0, max(1, len(codes) - ln)))
            s = codes[p:p + ln].copy()
        if rng.random() < conflict_frac and len(s) > 2:
            j = int(rng.integers(0, len(s)))
            s[j] = (s[j] + 1) & 3
        if rng.random() < revcomp_frac:
            s = (3 - s)[::-1].copy()
        seqs.append(s)
        depths.append(float(rng.integers(0, depth_max + 1)) + (0.5 if rng.random() < 0.3 else 0.0))
    offs = np.zeros(len(seqs) + 1, dtype=np.uint64)
    offs[1:] = np.cumsum([len(s) for s in seqs]).astype(np.uint64)
    data = lut[np.concatenate(seqs)] if seqs else np.zeros(0, dtype=np.uint8)
    return data, offs, np.asarray(depths, dtype=np.float64)


def count_instances(offsets, k):
    """Number of counted k-mer instances: sum over sequences of max(0, len - k - 1) (kcount_cpu.cpp:84-102,319)."""
    lens = np.diff(np.asarray(offsets).astype(np.int64))
    return int(np.maximum(lens - k - 1, 0).sum())


def make_contig_tiling(community, k, seed=3, min_len=500, max_len=50000, depth_lo=2, depth_hi=200, dup_frac=0.02,
                       conflict_frac=0.01, revcomp_frac=0.3):
    """BASELINE config C4's stand-in for the previous round's uutigs: the true genomes cut end to end into
    pieces of min_len..max_len bases with integer depths depth_lo..depth_hi; a few pieces are repeated (exact duplicates
    and overlapping tails: same k-mers from several contigs) and a few carry one substituted base (k-mers whose extensions
    disagree: the purge rules of kcount_cpu.cpp:356-406); some are reverse-complemented.
    Returns ASCII bytes, offsets (uint64[n+1]), depths (float64)."""
    codes, starts, _ = community
    rng = np.random.default_rng(seed)
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    cuts = []
    for g in range(len(starts) - 1):
        p, end = int(starts[g]), int(starts[g + 1])
        while p < end:
            ln = int(rng.integers(min_len, max_len + 1))
            if end - (p + ln) < min_len:
                ln = end - p
            cuts.append((p, p + ln))
            p += ln
    n0 = len(cuts)
    for _ in range(int(n0 * dup_frac)):
        a, b = cuts[int(rng.integers(0, n0))]
        if rng.random() < 0.5 and b - a > 2 * (k + 4):
            a += int(rng.integers(1, b - a - k - 3))
        cuts.append((a, b))
    lens = np.array([b - a for a, b in cuts], dtype=np.int64)
    offs = np.zeros(len(cuts) + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(lens).astype(np.uint64)
    data = np.empty(int(offs[-1]), dtype=np.uint8)
    flip = rng.random(len(cuts)) < revcomp_frac
    mutate = rng.random(len(cuts)) < conflict_frac
    for i, (a, b) in enumerate(cuts):
        s = codes[a:b]
        if mutate[i] and i >= n0:  # only repeated pieces: the original piece keeps the true base, the copy disagrees
            s = s.copy()
            j = int(rng.integers(0, len(s)))
            s[j] = (s[j] + 1) & 3
        if flip[i]:
            s = (3 - s)[::-1]
        data[int(offs[i]):int(offs[i + 1])] = lut[s]
    depths = rng.integers(depth_lo, depth_hi + 1, size=len(cuts)).astype(np.float64)
    return data, offs, depths


def make_skewed_reads(n_reads, community, read_len=150, error_rate=0.01, seed=1, repeat_frac=0.001, repeat_len=2000,
                      polya_frac=0.0002, out=None):
    """BASELINE config C5: reads of a Zipf community, plus repeat_frac of the reads from one short repeat
    (k-mers hundreds of times above the mean) and polya_frac poly-A reads with one variable base (ONE k-mer and a handful of
    neighbours seen > 65535 times: saturating counters, heavy-hitter partitions). The special reads are spread evenly through
    the set. Same output as make_reads."""
    n_rep = int(n_reads * repeat_frac) & ~1
    n_pa = int(n_reads * polya_frac)
    n_main = n_reads - n_rep - n_pa
    total = n_reads * read_len
    data = out if out is not None else np.empty(total, dtype=np.uint8)
    rows = data[:total].reshape(n_reads, read_len)
    is_special = np.zeros(n_reads, dtype=bool)
    n_sp = n_rep + n_pa
    if n_sp:
        is_special[np.linspace(0, n_reads - 1, n_sp).astype(np.int64)] = True
        n_sp = int(is_special.sum())  # linspace may repeat an index for tiny inputs
        n_pa = min(n_pa, n_sp)
        n_rep = n_sp - n_pa
        n_main = n_reads - n_sp
    main, _ = make_reads(n_main, community, read_len=read_len, error_rate=error_rate, seed=seed)
    rows[~is_special] = main.reshape(n_main, read_len)
    sp = np.empty((n_sp, read_len), dtype=np.uint8)
    if n_rep:
        rep = make_community(repeat_len, n_genomes=1, seed=seed + 101)
        r, _ = make_reads(n_rep, rep, read_len=read_len, error_rate=error_rate, seed=seed + 102)
        sp[:n_rep] = r.reshape(n_rep, read_len)
    if n_pa:
        pa = np.full((n_pa, read_len), 0 | (31 << 3), dtype=np.uint8)
        pa[:, read_len // 2] = (np.arange(n_pa) % 4) | (31 << 3)
        sp[n_rep:] = pa
    rng = np.random.default_rng(seed + 103)
    rows[is_special] = sp[rng.permutation(n_sp)]
    offsets = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(read_len)
    return data[:total], offsets

// Stage A — parse: sequences -> supermers bucketed by (owner shard, partition).
// Replaces SeqBlockInserter<MAX_K>::process_seq (reference src/kcount/kcount_cpu.cpp:73-103) together with
// Kmer::get_kmers (kmer.cpp:155-257), get_minimizer_fast (kmer.cpp:345-393) and KmerDHT::get_kmer_target_rank
// (kmer_dht.cpp:192-196) with O(1) work per k-mer. One THREAD per sequence (<= 256 bases; longer ones are tiled by the
// host), 32 sequences per warp marching in lock step, every thread working in its own interleaved slice of shared memory
// (word j of thread t at [j * blockDim + t]: conflict-free, no barriers anywhere):
//   1. 16-byte vector loads of the PackedRead bytes -> 2-bit codes (N -> G, kmer.cpp:169,187) + 1 "extension-valid" bit
//      per base (high quality AND one of ACGT: kcount.cpp:81-84, kcount_cpu.cpp:157-164,320-323), 4 bases per multiply;
//   2. phase A, one m-mer position per step: rolling forward / reverse-complement m-mer, canonical = min of the two
//      (left-aligned u64), its position packed into the 10 free low bits; sliding-window MAX over w = k-m+1 positions
//      = the minimizer of every k-mer, by the van Herk / Gil-Werman block decomposition (prefix maxima on the fly,
//      suffix maxima rewritten once per block of w) — branch-free and identical control flow in all 32 lanes;
//      each change of the window maximum appends (first k-mer, m-mer position) to a small per-thread list;
//   3. phase B, once per <= 32 list entries, again converged across lanes: recompute that m-mer, bucket =
//      (quick_hash(minimizer) % num_shards, partition from the high half of the same hash); a supermer = maximal run
//      (<= 64 k-mers) of equal buckets over k-mers 1..len-k-1 (first and last k-mer are never counted,
//      kcount_cpu.cpp:84-102); the thread appends {header, 2-bit bases, valid bits} to the bucket.
// The kernel runs twice per batch: COUNT (bucket histogram only) then FILL (exact-size buckets, no overflow possible).
#pragma once
#include "kc_device.cuh"

namespace kc {

struct BucketHist {  // per bucket, per batch; {nsup, nwords} is bumped by ONE 64-bit atomic
  u32 nsup;    // supermers
  u32 nwords;  // payload u32 words
  u32 ninst;   // k-mer instances
  u32 pad;
};

struct BucketBase {  // FILL: where the bucket's headers / payload start
  u64 hdr;
  u64 pay;
};

struct SupHdr {   // 8 bytes per supermer
  u32 woff;       // payload word offset inside the bucket's payload region
  uint16_t nb;    // bases in the supermer (= k-mers + k + 1)
  uint16_t depth; // 1 for reads; contig depth (kmer_count_t) for contig supermers (kmer_dht.hpp:78-91 Supermer::count)
};

constexpr int PARSE_TILE = 256;     // longest sequence one thread handles (positions fit 8 bits)
constexpr int PARSE_LCAP = 32;      // list entries per thread between two phase-B drains
constexpr int SUP_MAX_KMERS = 64;   // longest supermer, in k-mers (supermer boundaries are free to differ from the reference's)
constexpr int PARSE_THREADS = 128;

struct ParseArgs {
  const uint8_t *bytes;    // sequence bytes; sequence s occupies bytes[start, start+len)
  const u64 *offsets;      // direct mode: start = offsets[s], len = offsets[s+1]-start (reads, PackedReads flattened)
  const u64 *seg_start;    // segment mode (contigs / long reads): explicit start, len, depth per segment
  const u32 *seg_len;
  const uint16_t *seg_depth;
  u64 nseq;
  int ascii;               // 0: PackedRead bytes (base | qual<<3), 1: ASCII ACGTN (contigs)
  int k, m, w;             // k-mer length, minimizer length, window = k-m+1
  u32 qual_add;            // (0x80 - qual_cutoff) in every byte: SIMD "quality >= cutoff" test
  u32 num_shards, num_parts;  // partitions per shard (any number >= 1)
  int nsb, nsv;            // shared words per thread for bases / valid bits
  BucketHist *hist;        // COUNT: += ; FILL: running cursors (start at 0)
  const BucketBase *base;  // FILL
  SupHdr *hdr;
  u32 *payload;
  u64 *instances;          // += number of k-mer instances parsed
  u32 *err;                // |= 1 on a contig character outside ACGTNacgtn
};

__device__ __forceinline__ u32 bucket_of(u64 minimizer, u32 num_shards, u32 num_parts) {
  const u64 qh = quick_hash(minimizer);
  u32 shard;
  if (num_shards == 1) shard = 0;
  else if ((num_shards & (num_shards - 1)) == 0) shard = (u32)qh & (num_shards - 1);
  else shard = (u32)(qh % num_shards);                 // kmer_dht.cpp:195
  const u32 part = __umulhi((u32)(qh >> 32), num_parts);  // internal: which shared-memory-sized sub-table of that shard
  return shard * num_parts + part;
}

struct Strided {  // thread-private array interleaved across the block
  u32 *p;
  int stride;
  __device__ __forceinline__ u32 operator[](int i) const { return p[i * stride]; }
  __device__ __forceinline__ void set(int i, u32 v) const { p[i * stride] = v; }
};

// 16 PackedRead bytes (packed_reads.cpp:73-109: base 0..4 | min(qual,31) << 3) -> 16 2-bit codes MSB first + 16 valid bits
// LSB first. Four bases per multiply: the constants move each byte's field to the top byte without overlaps.
__device__ __forceinline__ void decode16_packed(const uint4 v, u32 qual_add, u32 &codes, u32 &vbits) {
  const u32 wd[4] = {v.x, v.y, v.z, v.w};
  codes = 0;
  vbits = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const u32 b3 = wd[i] & 0x07070707u;
    const u32 nm = (b3 >> 2) & 0x01010101u;                               // N (code >= 4)
    const u32 c = (b3 & 0x03030303u & ~(nm * 3u)) | (nm << 1);            // N -> G (kmer.cpp:169,187)
    const u32 q = (wd[i] >> 3) & 0x1F1F1F1Fu;
    const u32 hv = ((q + qual_add) >> 7) & 0x01010101u & ~nm;             // quality >= cutoff and not N
    codes |= ((c * 0x40100401u) >> 24) << (24 - 8 * i);
    vbits |= (((hv * 0x01020408u) >> 24) & 0xFu) << (4 * i);
  }
}

// 16 ASCII bytes (contigs: ACGTN, lower case = low quality, kcount_cpu.cpp:311-313). Bytes [lo, hi) of the chunk belong to
// the sequence and are checked for illegal characters (the reference DIEs on them, kcount_cpu.cpp:455-459).
__device__ __forceinline__ void decode16_ascii(const uint4 v, int lo, int hi, u32 &codes, u32 &vbits, bool &bad) {
  const u32 wd[4] = {v.x, v.y, v.z, v.w};
  codes = 0;
  vbits = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) {
    const u32 b = (wd[i >> 2] >> (8 * (i & 3))) & 0xFFu;
    const u32 up = b & 0xDFu;
    const u32 x = (b & 4) >> 1;  // kmer.cpp:187-188
    const u32 code = x + ((x ^ (b & 2)) >> 1);
    const bool acgt = (up == 'A') | (up == 'C') | (up == 'G') | (up == 'T');
    codes |= code << (30 - 2 * i);
    vbits |= (u32)(acgt && !(b & 0x20u)) << i;
    if (i >= lo && i < hi && !acgt && up != 'N') bad = true;
  }
}

template <bool FILL>
__global__ void __launch_bounds__(PARSE_THREADS) parse_kernel(const ParseArgs a) {
  extern __shared__ __align__(16) u32 smem_parse[];
  const int NT = blockDim.x, t = threadIdx.x;
  const Strided sb{smem_parse + t, NT};                                    // [nsb] 2-bit bases, 16 per word, MSB first
  const Strided sv{smem_parse + (size_t)a.nsb * NT + t, NT};               // [nsv] valid bits, LSB first
  u64 *vh = (u64 *)(smem_parse + (size_t)(a.nsb + a.nsv) * NT) + t;        // [w] van Herk block: element j at vh[j*NT]
  uint16_t *lst = (uint16_t *)(smem_parse + (size_t)(a.nsb + a.nsv + 2 * a.w) * NT) + t;  // [LCAP] entry e at lst[e*NT]
  const unsigned FULL = 0xffffffffu;
  const int k = a.k, m = a.m, w = a.w;
  const int msh = 64 - 2 * m;
  const u64 mmask = ~0ULL << msh;  // ZERO_MASK[m], kmer.cpp:81-87 (1 <= m <= 27)
  u32 local_inst = 0;
  bool bad = false;

  for (u64 s0 = (u64)blockIdx.x * NT; s0 < a.nseq; s0 += (u64)gridDim.x * NT) {
    const u64 s = s0 + (u64)t;
    bool act = s < a.nseq;
    u64 start = 0;
    int len = 0;
    u32 depth = 1;
    if (act) {
      if (a.seg_start) {
        start = a.seg_start[s];
        len = (int)a.seg_len[s];
        if (a.seg_depth) depth = a.seg_depth[s];
      } else {
        start = a.offsets[s];
        len = (int)(a.offsets[s + 1] - start);
      }
      if (len < k + 2) act = false;  // kcount.cpp:79 (len < k) and: fewer than k+2 bases hold no counted k-mer
    }
    int a0 = 0, nk = 0;
    // ---- 1. load + pack: aligned 16-byte chunks; base q of the sequence sits at stream position a0 + q ----
    if (act) {
      const uint8_t *src = a.bytes + start;
      a0 = (int)((uintptr_t)src & 15);
      const uint4 *p16 = (const uint4 *)(src - a0);
      const int end = a0 + len;
      const int n16 = (end + 15) >> 4;
      for (int c = 0; c < n16; c += 2) {
        const uint4 v0 = __ldg(p16 + c);
        const uint4 v1 = (c + 1 < n16) ? __ldg(p16 + c + 1) : make_uint4(0, 0, 0, 0);
        u32 c0, b0, c1, b1;
        if (a.ascii) {
          decode16_ascii(v0, a0 - 16 * c, end - 16 * c, c0, b0, bad);
          decode16_ascii(v1, a0 - 16 * (c + 1), (c + 1 < n16) ? end - 16 * (c + 1) : 0, c1, b1, bad);
        } else {
          decode16_packed(v0, a.qual_add, c0, b0);
          decode16_packed(v1, a.qual_add, c1, b1);
        }
        sb.set(c, c0);
        sb.set(c + 1, c1);
        sv.set(c >> 1, b0 | (b1 << 16));
      }
      const int nwp = (n16 + 1) & ~1;
      sb.set(nwp, 0);  // fetch64 may read 2 words past the last base
      sb.set(nwp + 1, 0);
      sb.set(nwp + 2, 0);
      sv.set(nwp >> 1, 0);
      sv.set((nwp >> 1) + 1, 0);
      for (int j = 0; j < w; j++) vh[(size_t)j * NT] = 0;
      nk = len - k - 1;  // k-mers i = 1..nk
      local_inst += (u32)nk;
    }
    const int nsteps = act ? nk + w - 1 : 0;  // m-mer positions p = 1..nsteps
    const int max_steps = __reduce_max_sync(FULL, nsteps);

    // ---- state of the two phases ----
    u64 f = 0, r = 0, pmax = 0, last = ~0ULL;
    int j = 0, cnt = 0;
    int run_i0 = 0;  // first k-mer of the open supermer (0 = none)
    u32 run_b = 0;
    if (act) {
      f = fetch64(sb, a0) & mmask;  // m-mer at position 0; every step shifts in the base that completes position p
      r = revcomp64(f) << msh;
    }

    auto emit_one = [&](int i0, int n_k, u32 b) {
      const int nb = n_k + k + 1;
      const int nwb = (nb + 15) >> 4, nwv = (nb + 31) >> 5;
      const u64 inc = 1ULL | ((u64)(nwb + nwv) << 32);
      if (!FILL) {
        atomicAdd((u64 *)&a.hist[b], inc);
        atomicAdd(&a.hist[b].ninst, (u32)n_k);
      } else {
        const u64 old = atomicAdd((u64 *)&a.hist[b], inc);
        const u32 pos = (u32)old, wpos = (u32)(old >> 32);
        const ulonglong2 bs = *(const ulonglong2 *)&a.base[b];
        SupHdr h;
        h.woff = wpos;
        h.nb = (uint16_t)nb;
        h.depth = (uint16_t)depth;
        a.hdr[bs.x + pos] = h;
        u32 *dst = a.payload + bs.y + wpos;
        const int q0 = a0 + i0 - 1;
        for (int x = 0; x < nwb; x++) dst[x] = (u32)(fetch64(sb, q0 + 16 * x) >> 32);
        for (int x = 0; x < nwv; x++) {
          const int qv = q0 + 32 * x;
          dst[nwb + x] = __funnelshift_r(sv[qv >> 5], sv[(qv >> 5) + 1], qv & 31);
        }
      }
    };
    auto emit_run = [&](int i0, int i1, u32 b) {  // k-mers i0..i1, cut into supermers of <= SUP_MAX_KMERS k-mers
      while (i0 <= i1) {
        const int n = min(SUP_MAX_KMERS, i1 - i0 + 1);
        emit_one(i0, n, b);
        i0 += n;
      }
    };
    auto drain = [&]() {  // phase B: converged across the warp, one list entry per lane and iteration
      const int maxc = __reduce_max_sync(FULL, cnt);
      for (int e = 0; e < maxc; e++) {
        if (e < cnt) {
          const u32 ent = lst[(size_t)e * NT];
          const int i = (int)(ent >> 8), pos = (int)(ent & 255u);
          const u32 b = bucket_of(canonical_mmer(fetch64(sb, a0 + pos), m), a.num_shards, a.num_parts);
          if (run_i0 == 0) {
            run_i0 = i;
            run_b = b;
          } else if (b != run_b) {
            emit_run(run_i0, i - 1, run_b);
            run_i0 = i;
            run_b = b;
          }
        }
      }
      cnt = 0;
    };

    // ---- 2. phase A ----
    for (int p = 1; p <= max_steps; p++) {
      if (p <= nsteps) {
        const u32 b = (u32)fetch_base(sb, a0 + p + m - 1);
        f = (f << 2) | ((u64)b << msh);
        r = ((r >> 2) | ((u64)(3u - b) << 62)) & mmask;
        const u64 c = (f < r ? f : r) | (u64)p;  // canonical m-mer (kmer.cpp:345-393), position in the free low bits
        pmax = (j == 0 || c > pmax) ? c : pmax;  // prefix maximum of the current block
        u64 mx = pmax;
        if (j < w - 1) {
          const u64 sfx = vh[(size_t)(j + 1) * NT];  // suffix maximum of the previous block from offset j+1
          mx = sfx > mx ? sfx : mx;
        }
        vh[(size_t)j * NT] = c;
        if (p >= w && mx != last) {  // the window [p-w+1, p] is complete: k-mer i = p-w+1 has minimizer mx
          last = mx;
          lst[(size_t)cnt * NT] = (uint16_t)(((u32)(p - w + 1) << 8) | ((u32)mx & 255u));
          cnt++;
        }
      }
      if (j == w - 1) {  // end of a block (same step in every lane): turn it into suffix maxima in place
        if (p <= nsteps) {
          u64 sfx = vh[(size_t)(w - 1) * NT];
          for (int jj = w - 2; jj >= 0; jj--) {
            const u64 v = vh[(size_t)jj * NT];
            sfx = v > sfx ? v : sfx;
            vh[(size_t)jj * NT] = sfx;
          }
        }
        j = 0;
      } else {
        j++;
      }
      if (__any_sync(FULL, cnt >= PARSE_LCAP)) drain();
    }
    drain();
    if (run_i0) emit_run(run_i0, nk, run_b);
  }
  if (bad) atomicOr(a.err, 1u);
  if (!FILL) {
    const u32 tot = __reduce_add_sync(FULL, local_inst);  // < 2^32 per warp for any realistic launch (checked on the host)
    if ((threadIdx.x & 31) == 0 && tot) atomicAdd(a.instances, (u64)tot);
  }
}

}  // namespace kc

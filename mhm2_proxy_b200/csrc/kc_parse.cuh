// Stage A — parse: sequences -> supermer records bucketed by (owner shard, partition).
// Replaces SeqBlockInserter<MAX_K>::process_seq (reference src/kcount/kcount_cpu.cpp:73-103) together with
// Kmer::get_kmers (kmer.cpp:155-257), get_minimizer_fast (kmer.cpp:345-393) and KmerDHT::get_kmer_target_rank
// (kmer_dht.cpp:192-196) with O(1) work per k-mer. One THREAD per sequence (<= 256 bases; longer ones are tiled by the
// host), 32 sequences per warp marching in lock step, every thread working in its own interleaved slice of shared memory
// (word j of thread t at [j * NT + t]: conflict-free, no barriers anywhere):
//   1. 16-byte vector loads of the PackedRead bytes -> 2-bit codes (N -> G, kmer.cpp:169,187) + 1 "extension-valid" bit
//      per base (high quality AND one of ACGT: kcount.cpp:81-84, kcount_cpu.cpp:157-164,320-323), 4 bases per multiply;
//   2. phase A, one m-mer position per step: rolling forward / reverse-complement m-mer, canonical = min of the two
//      (left-aligned u64), its position packed into the 10 free low bits; sliding-window MAX over w = k-m+1 positions
//      = the minimizer of every k-mer, by the van Herk / Gil-Werman block decomposition (prefix maxima on the fly,
//      suffix maxima rewritten once per block of w) — branch-free and identical control flow in all 32 lanes.
//      Every change of the window maximum (value or position) starts a new supermer and appends (first k-mer, m-mer
//      position) to a small per-thread list; a maximum lives at most w steps, so a supermer has <= w k-mers;
//   3. phase B, per group of 4 list entries and converged across lanes: recompute the m-mer, bucket =
//      (quick_hash(minimizer) % num_shards, partition from the high half of the same hash), reserve a record slot in
//      the bucket (ONE 32-bit atomic), build the fixed-size record in registers and write it with 16-byte stores.
// Supermer record = rw u32 words (k <= 21: 4 words = 16 bytes; else a multiple of 8 words = whole 32-byte sectors):
//      word 0: nb (bases, = k-mers + k + 1) | depth << 16     (depth: 1 for reads, contig depth for contig supermers)
//      words 1 .. nwb: 2-bit bases, 16 per word, MSB first     (nwb = ceil(nb_max / 16), nb_max = w + k + 1)
//      words 1+nwb .. : "extension-valid" bits, LSB first      (nwv = ceil(nb_max / 32))
// Only k-mers 1..len-k-1 of a sequence are counted (first and last never are, kcount_cpu.cpp:84-102); supermer
// boundaries are free to differ from the reference's.
// Two kernels per batch: parse_kernel (steps 1-2, bucket histogram, and the list of every sequence's supermers written
// to HBM as one u32 per supermer: bucket | first k-mer << 24) -> bucket scan -> fill_kernel (step 1 again, then step 3 fed
// from that list: no minimizer, no hash is computed twice) into exact-size buckets, no overflow possible.
#pragma once
#include "kc_device.cuh"

namespace kc {

struct BucketHist {  // per bucket, per batch; bumped by ONE 64-bit RED per supermer in the parse kernel
  u32 nsup;   // supermers
  u32 ninst;  // k-mer instances
};

struct RecGeom {  // geometry of a supermer record for one k
  int nb_max, nwb, nwv, rw;
};
__host__ __device__ inline RecGeom rec_geom(int k, int w) {
  RecGeom g;
  g.nb_max = w + k + 1;
  g.nwb = (g.nb_max + 15) >> 4;
  g.nwv = (g.nb_max + 31) >> 5;
  const int need = 1 + g.nwb + g.nwv;
  g.rw = need <= 4 ? 4 : (need + 7) & ~7;  // 16 bytes (k <= 21), else whole 32-byte sectors: see fill_kernel's lane-pair stores
  return g;
}

constexpr int PARSE_TILE = 256;  // longest sequence one thread handles (positions fit 8 bits)
constexpr int PARSE_LCAP = 32;   // list entries per thread between two phase-B drains

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
  int nwb, nwv, rw;        // record geometry
  BucketHist *hist;        // parse: bucket histogram (+=)
  u32 *cursor;             // fill: index of the next free record of every bucket (starts at the bucket's first record)
  int l2_hint;             // fill: record stores carry an L2 evict_last policy
  int pair_lanes;          // fill: records of whole 32-byte sectors are written sector by sector by lane pairs
  u32 *rec;                // fill: records
  u32 *elist;              // supermer lists: entry e of the sequence handled by thread t of chunk c (= NT consecutive
  int ecap;                //   sequences) at elist[((u64)c * ecap + e) * NT + t]; ecap = longest sequence - k - 1
  uint8_t *ecnt;           // [nseq] entries in the list (a sequence of <= 256 bases has <= 254 supermers)
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

template <int NT>
struct Strided {  // thread-private array interleaved across the block
  u32 *p;
  __device__ __forceinline__ u32 operator[](int i) const { return p[i * NT]; }
  __device__ __forceinline__ void set(int i, u32 v) const { p[i * NT] = v; }
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

// L2 residency hints of the fill kernel. A record store fills half (k <= 31) or all of a 32-byte sector of a bucket that
// receives its next record microseconds later: sectors evicted half-written cost a DRAM read-modify-write (ncu, v12: 5.2 GB
// read + 6.9 GB written for 2.3 + 4.6 GB of algorithmic traffic at k=21). Record stores therefore ask L2 to evict them last,
// the streaming list reads to evict them first.
__device__ __forceinline__ u64 l2_policy_evict_last() {
  u64 pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void st_record(uint4 *p, const uint4 v, int hint, u64 pol) {
  if (hint)
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
  else
    *p = v;
}
__device__ __forceinline__ u64 l2_policy_evict_first() {
  u64 pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ u32 ldg_stream(const u32 *p, u64 pol) {
  u32 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
  return v;
}

// sequence s of the launch: where it starts, how long it is, its depth; false when it holds no counted k-mer
__device__ __forceinline__ bool seq_header(const ParseArgs &a, u64 s, u64 &start, int &len, u32 &depth) {
  if (s >= a.nseq) return false;
  if (a.seg_start) {
    start = a.seg_start[s];
    len = (int)a.seg_len[s];
    if (a.seg_depth) depth = a.seg_depth[s];
  } else {
    start = a.offsets[s];
    len = (int)(a.offsets[s + 1] - start);
  }
  return len >= a.k + 2;  // kcount.cpp:79 (len < k) and: fewer than k+2 bases hold no counted k-mer
}

// ---- 1. load + pack: aligned 16-byte chunks; base q of the sequence sits at stream position a0 + q ----
template <int NT>
__device__ __forceinline__ int load_pack(const ParseArgs &a, const Strided<NT> sb, const Strided<NT> sv, u64 start, int len, bool &bad) {
  const uint8_t *src = a.bytes + start;
  const int a0 = (int)((uintptr_t)src & 15);
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
  return a0;
}

template <int NT>
__global__ void __launch_bounds__(NT, 8 * 128 / NT) parse_kernel(const ParseArgs a) {
  extern __shared__ __align__(16) u32 smem_parse[];
  const int t = threadIdx.x;
  const Strided<NT> sb{smem_parse + t};                                   // [nsb] 2-bit bases, 16 per word, MSB first
  const Strided<NT> sv{smem_parse + a.nsb * NT + t};                      // [nsv] valid bits, LSB first
  u64 *const vh = (u64 *)(smem_parse + (a.nsb + a.nsv) * NT) + t;         // [w] van Herk block: element j at vh[j*NT]
  uint16_t *const lst = (uint16_t *)(smem_parse + (a.nsb + a.nsv + 2 * a.w) * NT) + t;  // [LCAP+1] entry e at lst[e*NT]
  const unsigned FULL = 0xffffffffu;
  const int k = a.k, m = a.m, w = a.w;
  const int msh = 64 - 2 * m;
  const u64 mmask = ~0ULL << msh;  // ZERO_MASK[m], kmer.cpp:81-87 (1 <= m <= 27)
  u32 local_inst = 0;
  bool bad = false;

  for (u64 s0 = (u64)blockIdx.x * NT; s0 < a.nseq; s0 += (u64)gridDim.x * NT) {
    const u64 s = s0 + (u64)t;
    u64 start = 0;
    int len = 0;
    u32 depth = 1;
    const bool act = seq_header(a, s, start, len, depth);
    int a0 = 0, nk = 0;
    if (act) {
      a0 = load_pack<NT>(a, sb, sv, start, len, bad);
      for (int j = 0; j < w; j++) vh[j * NT] = 0;
      nk = len - k - 1;  // k-mers i = 1..nk
      local_inst += (u32)nk;
    }
    const int nsteps = act ? nk + w - 1 : 0;  // m-mer positions p = 1..nsteps
    const int max_steps = __reduce_max_sync(FULL, nsteps);

    u64 f = 0, r = 0, pmax = 0, last = ~0ULL;
    int j = 0, cnt = 0;
    if (act) {
      f = fetch64(sb, a0) & mmask;  // m-mer at position 0; every step shifts in the base that completes position p
      r = revcomp64(f) << msh;
    }
    u32 *const el = a.elist + ((u64)(s0 / NT) * (u64)a.ecap) * NT + t;  // this thread's list, entry e at el[e * NT]
    int ebase = 0;

    // Supermer e = k-mers [i_e, i_{e+1} - 1]; `fin`: the sequence is done, its last supermer ends at k-mer nk.
    // Otherwise the last entry stays pending (moved to the front of the list) until the next one arrives.
    auto drain = [&](bool fin) {
      const int ndone = fin ? cnt : max(cnt - 1, 0);
      const int maxd = __reduce_max_sync(FULL, ndone);
      for (int e = 0; e < maxd; e++) {
        if (e < ndone) {
          const u32 ent = lst[e * NT];
          const int i0 = (int)(ent >> 8), mp = (int)(ent & 255u);
          const int i1 = (e + 1 < cnt) ? (int)(lst[(e + 1) * NT] >> 8) - 1 : nk;
          const u32 b = bucket_of(canonical_mmer(fetch64(sb, a0 + mp), m), a.num_shards, a.num_parts);
          atomicAdd((u64 *)&a.hist[b], 1ULL | ((u64)(i1 - i0 + 1) << 32));
          el[(size_t)(ebase + e) * NT] = b | ((u32)i0 << 24);
        }
      }
      ebase += ndone;
      if (!fin && cnt > 0) {
        lst[0] = lst[(cnt - 1) * NT];
        cnt = 1;
      } else {
        cnt = 0;
      }
    };

    // ---- 2. phase A ----
    u64 *vh_j = vh;  // &vh[j * NT]
    for (int p = 1; p <= max_steps; p++) {
      if (p <= nsteps) {
        const u32 b = (u32)fetch_base(sb, a0 + p + m - 1);
        f = (f << 2) | ((u64)b << msh);
        r = ((r >> 2) | ((u64)(3u - b) << 62)) & mmask;
        const u64 c = (f < r ? f : r) | (u64)p;  // canonical m-mer (kmer.cpp:345-393), position in the free low bits
        pmax = (j == 0 || c > pmax) ? c : pmax;  // prefix maximum of the current block
        u64 mx = pmax;
        if (j < w - 1) {
          const u64 sfx = vh_j[NT];  // suffix maximum of the previous block from offset j+1
          mx = sfx > mx ? sfx : mx;
        }
        *vh_j = c;
        if (p >= w && mx != last) {  // the window [p-w+1, p] is complete: k-mer i = p-w+1 has minimizer mx
          last = mx;
          lst[cnt * NT] = (uint16_t)(((u32)(p - w + 1) << 8) | ((u32)mx & 255u));
          cnt++;
        }
      }
      if (j == w - 1) {  // end of a block (same step in every lane): turn it into suffix maxima in place
        if (p <= nsteps) {
          u64 sfx = *vh_j;
          for (int jj = w - 2; jj >= 0; jj--) {
            const u64 v = vh[jj * NT];
            sfx = v > sfx ? v : sfx;
            vh[jj * NT] = sfx;
          }
        }
        j = 0;
        vh_j = vh;
      } else {
        j++;
        vh_j += NT;
      }
      if (__any_sync(FULL, cnt >= PARSE_LCAP)) drain(false);
    }
    drain(true);
    if (s < a.nseq) a.ecnt[s] = (uint8_t)ebase;
  }
  if (bad) atomicOr(a.err, 1u);
  const u32 tot = __reduce_add_sync(FULL, local_inst);  // < 2^32 per warp for any realistic launch
  if ((threadIdx.x & 31) == 0 && tot) atomicAdd(a.instances, (u64)tot);
}

// Step 3 on its own: the sequence is packed again (it is the cheapest part), its supermers come from the list the parse
// kernel left. Four supermers per thread are in flight: the bucket-cursor atomics of all four are issued first, then the
// records are assembled from the 2-bit / valid words in shared memory and written with 128-bit stores.
// Must run with the grid and block size of the parse kernel that wrote the lists (chunk c = NT consecutive sequences).
template <int NT>
__global__ void __launch_bounds__(NT, 8 * 128 / NT) fill_kernel(const ParseArgs a) {
  extern __shared__ __align__(16) u32 smem_parse[];
  const int t = threadIdx.x;
  const Strided<NT> sb{smem_parse + t};
  const Strided<NT> sv{smem_parse + a.nsb * NT + t};
  const unsigned FULL = 0xffffffffu;
  const int k = a.k;
  bool bad = false;
  const u64 pol = l2_policy_evict_last(), pol_first = l2_policy_evict_first();
  for (u64 s0 = (u64)blockIdx.x * NT; s0 < a.nseq; s0 += (u64)gridDim.x * NT) {
    const u64 s = s0 + (u64)t;
    u64 start = 0;
    int len = 0;
    u32 depth = 1;
    const bool act = seq_header(a, s, start, len, depth);
    int a0 = 0, nk = 0, cnt = 0;
    if (act) {
      a0 = load_pack<NT>(a, sb, sv, start, len, bad);
      nk = len - k - 1;
      cnt = (int)a.ecnt[s];
    }
    const u32 *const el = a.elist + ((u64)(s0 / NT) * (u64)a.ecap) * NT + t;
    const int maxd = __reduce_max_sync(FULL, cnt);
    // list entries e0 .. e0+4 of the current group (the fifth closes the fourth supermer); the next group's are requested
    // before this group's atomics are issued, so that a group costs one memory round trip, not one per supermer
    u32 cur[5];
#pragma unroll
    for (int u = 0; u < 5; u++) cur[u] = (u < cnt) ? ldg_stream(el + (size_t)u * NT, pol_first) : 0u;
    // records of the previous group: written only after this group's atomics are in flight (their own atomics have had a
    // whole iteration to return). meta = first k-mer | bases << 8, 0 = no record.
    u32 ppos[4], pmeta[4] = {0u, 0u, 0u, 0u};
    // word x of the record of a supermer that starts at stream position q0 and has nb bases: 0 = header, 1..nwb = bases,
    // then valid bits; unused words are zero
    auto rec_word = [&](int x, int q0, int nb) -> u32 {
      const int nwb_s = (nb + 15) >> 4, nwv_s = (nb + 31) >> 5;  // words this supermer really fills
      if (x == 0) return (u32)nb | (depth << 16);
      if (x <= a.nwb) return (x - 1 < nwb_s) ? (u32)(fetch64(sb, q0 + 16 * (x - 1)) >> 32) : 0u;
      if (x - 1 - a.nwb < nwv_s) {
        const int qv = q0 + 32 * (x - 1 - a.nwb);
        return __funnelshift_r(sv[qv >> 5], sv[(qv >> 5) + 1], qv & 31);
      }
      return 0u;
    };
    auto write_records = [&]() {
      if (a.rw >= 8 && a.pair_lanes) {
        // Records of whole 32-byte sectors (k >= 22). Written by its own thread a sector arrives at L2 as two half-sector
        // stores, and L2 takes ~41 G store requests/s whatever their size (tools/ubench_scatter.cu: 32-byte records, one
        // thread 35 G records/s, lane pairs 67 G/s). Neighbouring lanes therefore swap halves, so that every store
        // instruction carries whole sectors: first the even lane's sector (even lane: words 0-3, odd lane: words 4-7),
        // then the odd lane's.
        const bool odd = (t & 1) != 0;
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const bool have = pmeta[u] != 0u;
          const int q0 = a0 + (int)(pmeta[u] & 255u) - 1;
          const int nb = (int)(pmeta[u] >> 8);
          const u64 mydst = (u64)(a.rec + (u64)ppos[u] * (u64)a.rw);
          const u64 odst = __shfl_xor_sync(FULL, mydst, 1);
          const bool ohave = __shfl_xor_sync(FULL, (int)have, 1) != 0;
          for (int x0 = 0; x0 < a.rw; x0 += 8) {
            u32 wd[8];
#pragma unroll
            for (int x = 0; x < 8; x++) wd[x] = have ? rec_word(x0 + x, q0, nb) : 0u;
            u32 rcv[4];
#pragma unroll
            for (int i = 0; i < 4; i++) rcv[i] = __shfl_xor_sync(FULL, odd ? wd[i] : wd[4 + i], 1);
            const u64 off = (u64)x0 * 4u;
            if (odd ? ohave : have)  // sector of the even lane's record
              st_record((uint4 *)(odd ? odst + off + 16 : mydst + off),
                        odd ? make_uint4(rcv[0], rcv[1], rcv[2], rcv[3]) : make_uint4(wd[0], wd[1], wd[2], wd[3]), a.l2_hint, pol);
            if (odd ? have : ohave)  // sector of the odd lane's record
              st_record((uint4 *)(odd ? mydst + off + 16 : odst + off),
                        odd ? make_uint4(wd[4], wd[5], wd[6], wd[7]) : make_uint4(rcv[0], rcv[1], rcv[2], rcv[3]), a.l2_hint, pol);
          }
        }
        return;
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        if (pmeta[u]) {
          uint4 *dst = (uint4 *)(a.rec + (u64)ppos[u] * (u64)a.rw);
          const int q0 = a0 + (int)(pmeta[u] & 255u) - 1;  // stream position of the supermer's first base (k-mer i0 - 1's first)
          const int nb = (int)(pmeta[u] >> 8);
          for (int x0 = 0; x0 < a.rw; x0 += 4)
            st_record(dst + (x0 >> 2), make_uint4(rec_word(x0, q0, nb), rec_word(x0 + 1, q0, nb), rec_word(x0 + 2, q0, nb), rec_word(x0 + 3, q0, nb)),
                      a.l2_hint, pol);
        }
      }
    };
    for (int e0 = 0; e0 < maxd; e0 += 4) {
      u32 nxt[4];
#pragma unroll
      for (int u = 0; u < 4; u++) nxt[u] = (e0 + 5 + u < cnt) ? ldg_stream(el + (size_t)(e0 + 5 + u) * NT, pol_first) : 0u;
      u32 npos[4], nmeta[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int e = e0 + u;
        nmeta[u] = 0u;
        npos[u] = 0u;
        if (e < cnt) {
          const u32 b = cur[u] & 0xFFFFFFu;
          const int i0 = (int)(cur[u] >> 24);
          const int i1 = (e + 1 < cnt) ? (int)(cur[u + 1] >> 24) - 1 : nk;
          nmeta[u] = (u32)i0 | ((u32)(i1 - i0 + 1 + k + 1) << 8);
          npos[u] = atomicAdd(&a.cursor[b], 1u);
        }
      }
      cur[0] = cur[4];
#pragma unroll
      for (int u = 0; u < 4; u++) cur[u + 1] = nxt[u];
      write_records();
#pragma unroll
      for (int u = 0; u < 4; u++) { ppos[u] = npos[u]; pmeta[u] = nmeta[u]; }
    }
    write_records();
  }
}

}  // namespace kc

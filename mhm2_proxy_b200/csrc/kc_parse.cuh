// Stage A — parse: sequences -> supermers bucketed by (owner shard, L2 partition).
// Replaces SeqBlockInserter<MAX_K>::process_seq (reference src/kcount/kcount_cpu.cpp:73-103) together with
// Kmer::get_kmers (kmer.cpp:155-257), get_minimizer_fast (kmer.cpp:345-393) and KmerDHT::get_kmer_target_rank
// (kmer_dht.cpp:192-196) — but as one warp per sequence with O(1)-amortised work per k-mer:
//   1. coalesced byte loads -> 2-bit codes (N -> G, kmer.cpp:169,187) + 1 "extension-valid" bit per base
//      (high quality AND one of ACGT: kcount.cpp:81-84, kcount_cpu.cpp:157-164,320-323), packed with REDUX/ballot;
//   2. canonical m-mer of every position (min(fwd, revcomp), left-aligned u64);
//   3. sliding-window MAX over k-m+1 positions by in-place doubling in shared memory = the minimizer of every k-mer;
//   4. bucket = (quick_hash(minimizer) % num_shards, partition from the high half of the same hash); supermer = maximal
//      run (<= 64 k-mers) of equal buckets over k-mers 1..len-k-1 (first and last k-mer are never counted, kcount_cpu.cpp:84-102);
//   5. one lane per supermer appends {header, 2-bit bases, valid bits} to its bucket.
// The kernel runs twice per batch: COUNT (bucket histogram only) then FILL (exact-size buckets, no overflow possible).
#pragma once
#include "kc_device.cuh"

namespace kc {

struct BucketHist {  // per bucket, per batch
  u32 nsup;    // supermers
  u32 nwords;  // payload u32 words
  u32 ninst;   // k-mer instances
};

struct SupHdr {   // 8 bytes per supermer
  u32 woff;       // payload word offset inside the bucket's payload region
  uint16_t nb;    // bases in the supermer (= k-mers + k + 1)
  uint16_t depth; // 1 for reads; contig depth (kmer_count_t) for contig supermers (kmer_dht.hpp:78-91 Supermer::count)
};

struct ParseArgs {
  const uint8_t *bytes;    // sequence bytes; sequence s occupies bytes[start, start+len)
  const u64 *offsets;      // direct mode: start = offsets[s], len = offsets[s+1]-start (reads, PackedReads flattened)
  const u64 *seg_start;    // segment mode (contigs / long reads): explicit start, len, depth per segment
  const u32 *seg_len;
  const uint16_t *seg_depth;
  u64 nseq;
  int ascii;               // 0: PackedRead bytes (base | qual<<3), 1: ASCII ACGTN (contigs)
  int k, m, w;             // k-mer length, minimizer length, window = k-m+1
  int qual_cutoff;
  u32 num_shards, num_parts;  // partitions per shard (any number >= 1)
  int max_len;             // multiple of 32, >= every sequence length in this launch
  int nsb, nsv;            // shared words for bases / valid bits per warp
  int smem_per_warp;
  BucketHist *hist;        // COUNT: += ; FILL: running cursors (start at 0)
  const u64 *hdr_base;     // FILL: first header index of each bucket
  const u64 *pay_base;     // FILL: first payload word of each bucket
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

constexpr int PARSE_THREADS = 256;
constexpr int SUP_MAX_KMERS = 64;  // longest supermer, in k-mers (supermer boundaries are free to differ from the reference's)

template <bool FILL>
__global__ void __launch_bounds__(PARSE_THREADS) parse_kernel(const ParseArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  unsigned char *base = smem_raw + (size_t)wib * a.smem_per_warp;
  u64 *c = (u64 *)base;                 // [max_len] canonical m-mers, then windowed maxima
  u32 *sb = (u32 *)(c + a.max_len);     // [nsb] 2-bit bases, 16 per word, MSB first
  u32 *sv = sb + a.nsb;                 // [nsv] valid bits, LSB first
  u32 *st = sv + a.nsv;                 // [nsv] supermer-start bits over k-mers 1..nk (bit t of word j <-> k-mer 1+32j+t)
  u32 *bk = st + a.nsv;                 // [max_len] bucket of k-mer i
  const unsigned FULL = 0xffffffffu;
  u64 local_inst = 0;
  const int k = a.k, m = a.m, w = a.w;

  for (u64 s = (u64)blockIdx.x * wpb + wib; s < a.nseq; s += (u64)gridDim.x * wpb) {
    u64 start;
    int len;
    u32 depth = 1;
    if (a.seg_start) {
      start = a.seg_start[s];
      len = (int)a.seg_len[s];
      if (a.seg_depth) depth = a.seg_depth[s];
    } else {
      start = a.offsets[s];
      len = (int)(a.offsets[s + 1] - start);
    }
    if (len < k + 2) continue;  // kcount.cpp:79 (len < k) and: fewer than k+2 bases hold no counted k-mer
    const uint8_t *src = a.bytes + start;

    // ---- 1. pack ----
    const int nchunks = (len + 31) >> 5;
    for (int j = 0; j < nchunks; j++) {
      const int i = j * 32 + lane;
      u32 code = 0, valid = 0;
      if (i < len) {
        const u32 b = src[i];
        if (a.ascii) {
          const u32 up = b & 0xDFu;
          const u32 x = (b & 4) >> 1;              // kmer.cpp:187-188
          code = x + ((x ^ (b & 2)) >> 1);
          const bool acgt = (up == 'A') | (up == 'C') | (up == 'G') | (up == 'T');
          valid = acgt && !(b & 0x20u);            // lower case = low quality (kcount_cpu.cpp:311-313)
          if (!acgt && up != 'N') atomicOr(a.err, 1u);
        } else {
          const u32 b3 = b & 7;
          code = b3 < 4 ? b3 : 2;                  // N -> G
          valid = (b3 < 4) && ((int)(b >> 3) >= a.qual_cutoff);
        }
      }
      const u32 x = code << (30 - 2 * (lane & 15));
      const u32 hm = (lane < 16) ? 0x0000FFFFu : 0xFFFF0000u;
      const u32 wv = __reduce_or_sync(hm, x);
      const u32 vm = __ballot_sync(FULL, valid);
      if ((lane & 15) == 0) sb[2 * j + (lane >> 4)] = wv;
      if (lane == 0) sv[j] = vm;
    }
    if (lane < 3) sb[2 * nchunks + lane] = 0;
    if (lane < 2) sv[nchunks + lane] = 0;
    __syncwarp();

    // ---- 2. canonical m-mers ----
    const int nm = len - m + 1;
    for (int p = lane; p < nm; p += 32) c[p] = canonical_mmer(fetch64(sb, p), m);
    __syncwarp();

    // ---- 3. windowed maximum by doubling: afterwards c[p] = max c[p .. p+span) ----
    int span = 1;
    while (span * 2 <= w) {
      for (int p0 = 0; p0 < nm; p0 += 32) {
        const int p = p0 + lane;
        u64 v = 0;
        if (p < nm) {
          v = c[p];
          if (p + span < nm) { const u64 v2 = c[p + span]; v = v2 > v ? v2 : v; }
        }
        __syncwarp();
        if (p < nm) c[p] = v;
        __syncwarp();
      }
      span *= 2;
    }

    // ---- 4. bucket of every counted k-mer, supermer starts ----
    const int nk = len - k - 1;  // k-mers i = 1..nk
    u32 carry = 0xFFFFFFFFu;
    for (int j0 = 0; j0 < nk; j0 += 32) {
      const int i = 1 + j0 + lane;
      u32 b = 0xFFFFFFFFu;
      if (i <= nk) {
        const u64 v1 = c[i], v2 = c[i + w - span];
        b = bucket_of(v1 > v2 ? v1 : v2, a.num_shards, a.num_parts);
        bk[i] = b;
      }
      u32 pb = __shfl_up_sync(FULL, b, 1);
      if (lane == 0) pb = carry;
      // a supermer also ends after SUP_MAX_KMERS k-mers (the count stage indexes k-mers inside a supermer with 6 bits)
      const u32 sm = __ballot_sync(FULL, (i <= nk) && ((b != pb) || ((i - 1) & (SUP_MAX_KMERS - 1)) == 0));
      if (lane == 0) st[j0 >> 5] = sm;
      carry = __shfl_sync(FULL, b, 31);
    }
    __syncwarp();
    if (lane == 0) local_inst += (u64)nk;

    // ---- 5. one lane per supermer ----
    const int nsw = (nk + 31) >> 5;
    int ns = 0;
    for (int j = 0; j < nsw; j++) ns += __popc(st[j]);
    for (int t0 = 0; t0 < ns; t0 += 32) {
      const int t = t0 + lane;
      if (t < ns) {
        int acc = 0, j = 0;
        u32 wd = st[0];
        for (;;) {
          const int pc = __popc(wd);
          if (acc + pc > t) break;
          acc += pc;
          wd = st[++j];
        }
        const int bit = (int)__fns(wd, 0, t - acc + 1);
        const int i0 = 1 + 32 * j + bit;
        u32 rest = (bit == 31) ? 0u : (wd & (0xFFFFFFFFu << (bit + 1)));
        int jn = j;
        while (!rest && ++jn < nsw) rest = st[jn];
        const int i1 = rest ? (32 * jn + __ffs(rest) - 1) : nk;  // (1 + 32*jn + bit') - 1
        const int n_k = i1 - i0 + 1;
        const int nb = n_k + k + 1;
        const int q0 = i0 - 1;
        const int nwb = (nb + 15) >> 4, nwv = (nb + 31) >> 5;
        const u32 b = bk[i0];
        if (!FILL) {
          atomicAdd(&a.hist[b].nsup, 1u);
          atomicAdd(&a.hist[b].nwords, (u32)(nwb + nwv));
          atomicAdd(&a.hist[b].ninst, (u32)n_k);
        } else {
          const u32 pos = atomicAdd(&a.hist[b].nsup, 1u);
          const u32 wpos = atomicAdd(&a.hist[b].nwords, (u32)(nwb + nwv));
          SupHdr h;
          h.woff = wpos;
          h.nb = (uint16_t)nb;
          h.depth = (uint16_t)depth;
          a.hdr[a.hdr_base[b] + pos] = h;
          u32 *dst = a.payload + a.pay_base[b] + wpos;
          for (int x = 0; x < nwb; x++) dst[x] = (u32)(fetch64(sb, q0 + 16 * x) >> 32);
          for (int x = 0; x < nwv; x++) {
            const int qv = q0 + 32 * x;
            dst[nwb + x] = __funnelshift_r(sv[qv >> 5], sv[(qv >> 5) + 1], qv & 31);
          }
        }
      }
    }
    __syncwarp();
  }
  if (!FILL && lane == 0 && local_inst) atomicAdd(a.instances, local_inst);
}

}  // namespace kc

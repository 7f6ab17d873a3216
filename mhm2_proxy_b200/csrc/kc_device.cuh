// Device-side primitives of the kcount path (sm_100a). Bit-level semantics follow the reference's Kmer<MAX_K>
// (src/kmer.cpp) and hash_funcs.c; citations inline. Everything here is integer work except get_ext's one fp64 multiply.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace kc {

typedef unsigned long long u64;
typedef unsigned int u32;

constexpr int EXT_NONE = 4;  // extension that counts for nothing: low quality ('0') or N (kcount_cpu.cpp:157-164,320-323)

__host__ __device__ __forceinline__ u64 rotl64(u64 x, int r) { return (x << r) | (x >> (64 - r)); }

// hash_funcs.c:65-74
__host__ __device__ __forceinline__ u64 fmix64(u64 k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdULL;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ULL;
  k ^= k >> 33;
  return k;
}

// MurmurHash3_x64_128(key, 8*KW, seed 313) low 64 bits == Kmer::hash() (hash_funcs.c:78-190, kmer.cpp:465-468),
// specialised for keys of KW whole 8-byte words (body blocks of 16 bytes, then the 8-byte tail when KW is odd).
template <int KW>
__host__ __device__ __forceinline__ u64 murmur64_words(const u64 *w) {
  const u64 c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
  u64 h1 = 313, h2 = 313;
#pragma unroll
  for (int i = 0; i < KW / 2; i++) {
    u64 k1 = w[2 * i], k2 = w[2 * i + 1];
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
    h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
    k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
    h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
  }
  if (KW & 1) {
    u64 k1 = w[KW - 1];
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
  }
  h1 ^= (u64)(KW * 8); h2 ^= (u64)(KW * 8);
  h1 += h2; h2 += h1;
  h1 = fmix64(h1); h2 = fmix64(h2);
  h1 += h2;
  return h1;
}

// Slot hash of the count stage's shared-memory table. Which slot a k-mer lands in is internal (SURVEY.md Appendix A.7:
// slot hash, probing and capacity are free to differ from KmerMapExts::insert), so the table does not pay for the 45
// instructions of MurmurHash3 per instance: two multiply / xor-fold rounds (9 instructions at KW = 1) spread the keys of
// one partition — k-mers sharing a minimizer, many one substitution apart — as evenly as fmix64 does (tools/hash_quality.py).
// murmur64_words stays the bit-exact Kmer::hash (KAT-tested through kc_prim_kmers).
template <int KW>
__host__ __device__ __forceinline__ u64 table_hash(const u64 *w) {
  const u64 c = 0xD6E8FEB86659FD93ULL;
  u64 h = w[0];
#pragma unroll
  for (int i = 1; i < KW; i++) {
    h ^= h >> 32;
    h = h * c ^ w[i];
  }
  h ^= h >> 32;
  h *= c;
  h ^= h >> 32;
  h *= c;
  h ^= h >> 32;
  return h;
}

// hash_funcs.c:332-342
__host__ __device__ __forceinline__ u64 quick_hash(u64 v) {
  v = v * 3935559000370003845ULL + 2691343689449507681ULL;
  v ^= v >> 21;
  v ^= v << 37;
  v ^= v >> 4;
  v *= 4768777513237032717ULL;
  v ^= v << 20;
  v ^= v >> 41;
  v ^= v << 5;
  return v;
}

// Reverse-complement of 32 bases held MSB-first in a u64 (what TWIN_TABLE does byte-wise, kmer.cpp:66-79,492-497):
// reverse the order of the 2-bit groups and complement (code -> 3 - code == ~code).
__device__ __forceinline__ u64 revcomp64(u64 x) {
  // per 32-bit half: BREV reverses all bits (also inside each pair), swapping neighbouring bits restores the pairs, NOT
  // complements; the halves trade places. No bit crosses a half in the neighbour swap (the masks drop bits 31 / 0), so
  // this is 2 BREV + 2 x (2 shifts + 1 LOP3) instead of 64-bit shift pairs.
  const u32 lo = __brev((u32)x), hi = __brev((u32)(x >> 32));
  const u32 a = ~(((lo >> 1) & 0x55555555u) | ((lo << 1) & 0xAAAAAAAAu));
  const u32 b = ~(((hi >> 1) & 0x55555555u) | ((hi << 1) & 0xAAAAAAAAu));
  return ((u64)a << 32) | b;
}
__host__ __forceinline__ u64 revcomp64_host(u64 x) {
  u64 r = 0;
  for (int i = 0; i < 32; i++) { r = (r << 2) | (3 - (x & 3)); x >>= 2; }
  return r;
}

// Reverse complement of a k-mer in KW words (kmer.cpp:485-505): reverse word order + per-word revcomp, then shift the
// whole thing left by 2*(32*NW - k) where NW = ceil(k/32) words actually hold bases. Words >= NW stay zero.
template <int KW>
__device__ __forceinline__ void revcomp_words(const u64 *in, u64 *out, int k) {
  const int nw = (k + 31) >> 5;  // words holding bases (kmer.cpp:488 last_long)
  u64 t[KW], rcw[KW];
#pragma unroll
  for (int i = 0; i < KW; i++) rcw[i] = revcomp64(in[i]);
  // t[nw - 1 - i] = rcw[i]. nw is KW or (k % 32 == 0) KW - 1: two compile-time candidates per word and a uniform select —
  // a run-time index would push these arrays into local memory
  const bool full = nw == KW;
#pragma unroll
  for (int i = 0; i < KW; i++) {
    const u64 a = rcw[KW - 1 - i];
    const u64 b = (KW - 2 - i >= 0) ? rcw[(KW - 2 - i >= 0) ? KW - 2 - i : 0] : 0ULL;
    t[i] = full ? a : b;
  }
  const int sh = (k & 31) ? 2 * (32 - (k & 31)) : 0;
#pragma unroll
  for (int i = 0; i < KW; i++) {
    u64 v = 0;
    if (i < nw) {
      v = t[i] << sh;
      if (sh && i + 1 < nw) v |= t[i + 1] >> (64 - sh);
    }
    out[i] = v;
  }
}

// kmer.cpp:265-272 operator<
template <int KW>
__device__ __forceinline__ bool words_less(const u64 *a, const u64 *b) {
#pragma unroll
  for (int i = 0; i < KW; i++) {
    if (a[i] < b[i]) return true;
    if (a[i] > b[i]) return false;
  }
  return false;
}

// 64 bits (32 bases) starting at base `q` of a 2-bit stream stored MSB-first in u32 words (16 bases per word).
// The caller guarantees words [q/16, q/16+2] are readable.
template <typename P>
__device__ __forceinline__ u64 fetch64(P words, int q) {
  const int idx = q >> 4, sh = (q & 15) * 2;
  u32 w0 = words[idx], w1 = words[idx + 1], w2 = words[idx + 2];
  u32 hi = __funnelshift_l(w1, w0, sh);  // (w0:w1) << sh, upper word
  u32 lo = __funnelshift_l(w2, w1, sh);
  return ((u64)hi << 32) | lo;
}
// one base code at position q
template <typename P>
__device__ __forceinline__ int fetch_base(P words, int q) {
  return (words[q >> 4] >> (30 - 2 * (q & 15))) & 3;
}
// one bit at position q of an LSB-first bit stream in u32 words
template <typename P>
__device__ __forceinline__ int fetch_bit(P bits, int q) {
  return (bits[q >> 5] >> (q & 31)) & 1;
}

// canonical m-mer at the top of x (left-aligned, 2m bits): min(fwd, revcomp) (kmer.cpp:345-393)
__device__ __forceinline__ u64 canonical_mmer(u64 x, int m) {
  const u64 mask = ~0ULL << (64 - 2 * m);  // ZERO_MASK[m], kmer.cpp:81-87 (m >= 1, m <= 31)
  u64 f = x & mask;
  u64 r = revcomp64(f) << (64 - 2 * m);
  return f < r ? f : r;
}

// ExtCounts::get_ext (kcount_cpu.cpp:135-145,173-182): order the four counters descending, ties to the LARGER letter
// (T > G > C > A); dmin_dyn = max((int)((1.0 - DYN_MIN_DEPTH) * count), dmin_thres) in IEEE fp64 — `frac` is the host's
// (1.0 - dyn_min_depth), so 0.09999999999999998 for 0.9, and there is one multiply (no contraction possible).
// dmin_dyn of a k-mer: one IEEE fp64 multiply, truncation, floor at dmin_thres — shared by its left and right side
__host__ __device__ __forceinline__ int dmin_dyn_of(u32 count, int dmin_thres, double frac) {
#ifdef __CUDA_ARCH__
  int dmin_dyn = (int)__dmul_rn(frac, (double)count);
#else
  int dmin_dyn = (int)(frac * (double)count);
#endif
  return dmin_dyn < dmin_thres ? dmin_thres : dmin_dyn;
}
// extension code of one side: 0..3 = A C G T, 4 = 'F' (fork), 5 = 'X' (none)
__host__ __device__ __forceinline__ u32 ext_code(const u32 *e, int dmin_dyn) {
  // (value, index) kept in registers: indexing e[] with a run-time `top` would put the caller's counters in local memory
  u32 top = 3;  // start from T so that ties keep the larger letter
  u32 best = e[3];
#pragma unroll
  for (int i = 2; i >= 0; i--)
    if (e[i] > best) { best = e[i]; top = (u32)i; }
  u32 runner = 0;
#pragma unroll
  for (int i = 0; i < 4; i++)
    if ((u32)i != top && e[i] > runner) runner = e[i];
  if ((int)best < dmin_dyn) return 5u;
  if ((int)runner >= dmin_dyn) return 4u;
  return top;
}
__host__ __device__ __forceinline__ char ext_char_of(u32 c) { return c < 4u ? (char)((0x54474341u >> (8u * c)) & 0xFFu) : (c == 4u ? 'F' : 'X'); }
__host__ __device__ __forceinline__ char get_ext(const u32 *e, u32 count, int dmin_thres, double frac) {
  return ext_char_of(ext_code(e, dmin_dyn_of(count, dmin_thres, frac)));
}

__device__ __forceinline__ u32 sat16(u64 v) { return v > 65535ULL ? 65535u : (u32)v; }

}  // namespace kc

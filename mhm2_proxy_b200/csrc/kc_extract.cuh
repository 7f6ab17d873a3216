// Count stage, per-instance front end: supermer record in shared memory -> (canonical k-mer, left slot, right slot).
// Replaces get_kmers_and_exts (reference src/kcount/kcount_cpu.cpp:307-335) for one k-mer instance.
//
// Two implementations with the same contract:
//   * extract_fixed<KW, K>: k is a compile-time constant (the reference's default schedule 21,33,55,77,99 and k=31, which
//     BASELINE.json's configs name): every shift count, mask, word count and the record geometry fold into immediates, the
//     loads are exactly the ceil((k+17)/16) words that can hold bases j-1 .. j+k, and nothing is predicated on k.
//   * extract_any<KW>: any k at run time (the round-1 code path, kept for every other k).
// Slots: left 0..3 = A C G T, 4 = none; right 0..3, 4 = none ("none": low quality or N, kcount_cpu.cpp:157-164,320-323).
#pragma once
#include "kc_device.cuh"
#include "kc_parse.cuh"

namespace kc {

// compile-time geometry of a k (kmer_dht.cpp:113-116 for m; kc_parse.cuh rec_geom for the record)
__host__ __device__ constexpr int kc_m_of(int k) {
  int m = k * 2 / 3 + 1;
  m = m < 15 ? 15 : m;
  m = m > 27 ? 27 : m;
  return m > k ? k : m;
}
__host__ __device__ constexpr int kc_w_of(int k) { return k - kc_m_of(k) + 1; }
// record geometry for k-mer length k and supermer window w (at most w k-mers per supermer)
__host__ __device__ constexpr int kc_nbmax_of(int k, int w) { return w + k + 1; }
__host__ __device__ constexpr int kc_nwb_of(int k, int w) { return (kc_nbmax_of(k, w) + 15) >> 4; }
__host__ __device__ constexpr int kc_nwv_of(int k, int w) { return (kc_nbmax_of(k, w) + 31) >> 5; }
__host__ __device__ constexpr int kc_rw_of(int k, int w) {
  const int need = 1 + kc_nwb_of(k, w) + kc_nwv_of(k, w);
  return need <= 4 ? 4 : (need + 7) & ~7;
}
__device__ __forceinline__ u32 lds32(u32 a) {  // a: shared-space address
  u32 v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ u32 lds16(u32 a) {
  u32 v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}

// reverse the order of the 16 bases of a u32 and complement them: BREV, swap the two bits of every pair, NOT — the swap and
// the NOT are ONE LOP3: a bitwise select between x >> 1 and x << 1 by the constant 0x55555555, inverted
__device__ __forceinline__ u32 revcomp32(u32 x) {
  const u32 v = __brev(x);
  u32 r;
  // f(a, b, c) = ~((a & c) | (b & ~c)); the lut is f evaluated on the canonical operand patterns 0xF0, 0xCC, 0xAA
  constexpr u32 LUT = ~((0xF0u & 0xAAu) | (0xCCu & ~0xAAu)) & 0xFFu;
  static_assert(LUT == 0x1B, "lop3 truth table");
  asm("lop3.b32 %0, %1, %2, 0x55555555, %3;" : "=r"(r) : "r"(v >> 1), "r"(v << 1), "n"(LUT));
  return r;
}

struct Instance {  // what the table needs to know about one k-mer instance
  u32 l, r;        // slots 0..4
};

// k and the supermer window known at compile time. rec_s: shared address of the staged record, j: 1-based k-mer index inside
// the supermer.
template <int KW, int K, int W_>
__device__ __forceinline__ Instance extract_fixed(u32 rec_s, int j, u64 (&key)[KW]) {
  constexpr int NWB = kc_nwb_of(K, W_);
  constexpr int NS = (K + 2 + 15) / 16;     // 32-bit words that hold bases j-1 .. j+k once aligned
  constexpr int NWL = (K + 17 + 15) / 16;   // words to load: the same bases at the worst alignment
  constexpr int N32 = (K + 15) / 16;        // 32-bit words of the k-mer itself
  constexpr int LASTB = 2 * (K - 16 * (N32 - 1));  // bits used in the k-mer's last 32-bit word
  constexpr u32 LASTMASK = LASTB == 32 ? 0xFFFFFFFFu : (0xFFFFFFFFu << (32 - LASTB));
  constexpr int PAD = 32 * N32 - 2 * K;     // zero bits after the k-mer in its last 32-bit word
  static_assert(2 * KW >= N32, "key words");
  const int q = j - 1;  // the base before the k-mer
  const u32 wa = rec_s + 4u + 4u * (u32)(q >> 4);
  u32 wv[NWL];
#pragma unroll
  for (int i = 0; i < NWL; i++) wv[i] = lds32(wa + 4u * i);
  const u32 sh = (u32)(q & 15) * 2u;
  u32 S[NS];  // base j-1 in the top two bits of S[0]
#pragma unroll
  for (int i = 0; i < NS; i++) S[i] = (i + 1 < NWL) ? __funnelshift_l(wv[i + 1], wv[i], sh) : (wv[i] << sh);
  const u32 lcode = S[0] >> 30;
  constexpr int RB = K + 1;  // the base after the k-mer
  const u32 rcode = (S[RB >> 4] >> (30 - 2 * (RB & 15))) & 3u;
  u32 F[2 * KW];  // the k-mer, 16 bases per word, top aligned; words >= N32 are zero
#pragma unroll
  for (int i = 0; i < 2 * KW; i++) {
    if (i < N32) F[i] = (i + 1 < NS) ? __funnelshift_l(S[i + 1], S[i], 2) : (S[i] << 2);
    else F[i] = 0u;
  }
  F[N32 - 1] &= LASTMASK;
  // reverse complement (kmer.cpp:485-505): words in reverse order, each reversed + complemented, then the PAD ones that the
  // complemented padding left at the top are shifted out
  u32 R[2 * KW];
#pragma unroll
  for (int i = 0; i < 2 * KW; i++) R[i] = (i < N32) ? revcomp32(F[N32 - 1 - i]) : 0u;
  if (PAD) {
#pragma unroll
    for (int i = 0; i < N32; i++) R[i] = (i + 1 < N32) ? __funnelshift_l(R[i + 1], R[i], PAD) : (R[i] << PAD);
  }
  u64 fwd[KW], rc[KW];
#pragma unroll
  for (int i = 0; i < KW; i++) {
    fwd[i] = ((u64)F[2 * i] << 32) | F[2 * i + 1];
    rc[i] = ((u64)R[2 * i] << 32) | R[2 * i + 1];
  }
  const bool use_rc = words_less<KW>(rc, fwd);  // strict: palindromes keep the forward orientation (kcount_cpu.cpp:326-331)
#pragma unroll
  for (int i = 0; i < KW; i++) key[i] = use_rc ? rc[i] : fwd[i];
  // extension-valid bits (LSB first) follow the base words: bit j-1 for the left neighbour, bit j+k for the right one
  const u32 va = rec_s + 4u * (1 + NWB);
  u32 lval, rval;
  if (W_ <= 32) lval = (lds32(va) >> q) & 1u;  // j-1 <= w-1 < 32: always the first word
  else lval = (lds32(va + 4u * (u32)(q >> 5)) >> (q & 31)) & 1u;
  if (((K + 1) >> 5) == ((K + W_) >> 5)) rval = (lds32(va + 4u * ((K + 1) >> 5)) >> ((j + K) & 31)) & 1u;  // one fixed word
  else rval = (lds32(va + 4u * (u32)((j + K) >> 5)) >> ((j + K) & 31)) & 1u;
  // code + 4 when the neighbour does not count; complementing (x ^ 3) keeps that flag, min(x, 4) folds it into slot 4
  const u32 lx = lcode | ((lval ^ 1u) << 2), rx = rcode | ((rval ^ 1u) << 2);
  Instance in;  // on the reverse strand the neighbours trade places and are complemented (kcount_cpu.cpp:328-330)
  in.l = min(use_rc ? rx ^ 3u : lx, 4u);
  in.r = min(use_rc ? lx ^ 3u : rx, 4u);
  return in;
}

// k at run time (any k): rec points at the staged record in shared memory
template <int KW>
__device__ __forceinline__ Instance extract_any(const u32 *rec, int j, int k, int nwb, u64 (&key)[KW]) {
  const int nw = KW == 1 ? 1 : (k + 31) >> 5;  // words of the key that hold bases: KW, or KW - 1 when k % 32 == 0
  const u64 last_mask = (k & 31) ? (~0ULL << (64 - 2 * (k & 31))) : ~0ULL;
  const int rsh = 62 - 2 * (k & 31);  // where the base after the k-mer sits in the k-mer's last word (k % 32 != 0)
  const u32 *p = rec + 1;
  const u32 *pv = p + nwb;
  // words idx-1 .. idx+2*nw of the base stream hold the base before the k-mer, the k-mer and (k % 32 != 0) the base
  // after it: one pass of funnel shifts yields all three (p[-1] is the header word, read but never selected)
  const u32 *pw = p + (j >> 4);
  const u32 sh = (u32)(j & 15) * 2u;
  u32 wv[2 * KW + 2];
#pragma unroll
  for (int i = 0; i < 2 * KW + 2; i++) wv[i] = (i <= 2 * nw + 1) ? pw[i - 1] : 0u;
  u64 fwd[KW], rc[KW];
#pragma unroll
  for (int i = 0; i < KW; i++)
    fwd[i] = ((u64)__funnelshift_l(wv[2 * i + 2], wv[2 * i + 1], sh) << 32) | __funnelshift_l(wv[2 * i + 3], wv[2 * i + 2], sh);
  const u32 lcode = __funnelshift_l(wv[1], wv[0], sh) & 3u;
  // word nw - 1 keeps only its k % 32 bases; nw is KW or KW - 1 (static indices: no local memory). Before the mask
  // the same word still holds the base that follows the k-mer.
  const u64 lastw = (nw == KW) ? fwd[KW - 1] : fwd[KW >= 2 ? KW - 2 : 0];
  const u32 rcode = (k & 31) ? ((u32)(lastw >> rsh) & 3u) : (u32)fetch_base(p, j + k);
  if (nw == KW) fwd[KW - 1] &= last_mask;
  else if (KW >= 2) { fwd[KW >= 2 ? KW - 2 : 0] &= last_mask; fwd[KW - 1] = 0ULL; }
  const u32 lval = (u32)fetch_bit(pv, j - 1), rval = (u32)fetch_bit(pv, j + k);
  revcomp_words<KW>(fwd, rc, k);
  const bool use_rc = words_less<KW>(rc, fwd);
#pragma unroll
  for (int i = 0; i < KW; i++) key[i] = use_rc ? rc[i] : fwd[i];
  const u32 lx = lcode | ((lval ^ 1u) << 2), rx = rcode | ((rval ^ 1u) << 2);
  Instance in;
  in.l = min(use_rc ? rx ^ 3u : lx, 4u);
  in.r = min(use_rc ? lx ^ 3u : rx, 4u);
  return in;
}

}  // namespace kc

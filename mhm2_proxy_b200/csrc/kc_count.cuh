// Stage B — count: for a group of partitions whose tables fit in L2 together,
//   insert_kernel : supermers -> canonical k-mers + extensions -> open-addressing table (claim by CAS, counters by RED)
//                   replaces get_kmers_and_exts + KmerMapExts::insert + insert_supermer_from_read/_from_ctg
//                   (reference src/kcount/kcount_cpu.cpp:231-267,307-406)
//   filter_kernel : per slot depth filter + extension choice + stream compaction of survivors
//                   replaces HashTableInserter::insert_into_local_hashtable (kcount_cpu.cpp:490-528) and
//                   ExtCounts::get_ext (kcount_cpu.cpp:173-182)
// Table slot = key (KW u64) + VW u32 counters: left A,C,G,T,none | right A,C,G,T,none [| ctg n, ~min depth, ext-pair mask].
// count = sum of the five left counters (every instance increments exactly one of them); u32 accumulation then
// clamping to 65535 equals the reference's saturating u16 adds because the adds are monotone.
#pragma once
#include "kc_device.cuh"
#include "kc_parse.cuh"

namespace kc {

constexpr u64 KEY_EMPTY = 0xFFFFFFFFFFFFFFFFULL;   // last key word of a real k-mer is never all ones: it has >= 2 zero
constexpr u64 KEY_LOCKED = 0xFFFFFFFFFFFFFFFEULL;  // trailing bits because N_LONGS = k/32 + 1 (main.cpp:170)
constexpr int VW_READS = 10, VW_CTG = 13;
constexpr int INS_THREADS = 256;
constexpr int CH = 256;  // supermers per chunk
constexpr int MAX_PROBE = 256;

struct SegDesc {       // one contiguous run of supermers of one partition (one per block of received/parsed data)
  const SupHdr *hdr;
  const u32 *payload;  // the partition's payload region; SupHdr::woff is relative to it
  u32 nsup;
  u32 is_ctg;
  u32 chunk0;          // index of this segment's first chunk inside the partition
  u32 pad;
};

struct PartDesc {
  u64 *keys;           // [cap * KW]
  u32 *vals;           // [cap * VW]
  u32 cap_mask;        // cap - 1 (cap is a power of two)
  u32 seg_begin, seg_end;
  u32 nchunks;
  u32 part;            // partition id
  u32 pad;
};

__device__ __forceinline__ u64 ld_acquire(const u64 *p) {
  u64 v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(u64 *p, u64 v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Find or claim the slot of `key`. Returns slot index, or 0xFFFFFFFF when MAX_PROBE slots were all foreign.
template <int KW>
__device__ __forceinline__ u32 find_or_claim(u64 *keys, u32 cap_mask, const u64 *key, u32 slot, u32 &probes) {
  u32 probe = 0;
  if (KW == 1) {
    for (;;) {
      u64 cur = __ldcg(keys + slot);
      if (cur == KEY_EMPTY) cur = atomicCAS(keys + slot, KEY_EMPTY, key[0]);
      if (cur == KEY_EMPTY || cur == key[0]) break;
      if (++probe >= (u32)MAX_PROBE) { slot = 0xFFFFFFFFu; break; }
      slot = (slot + 1) & cap_mask;
    }
  } else {
    for (;;) {
      u64 *kp = keys + (size_t)slot * KW;
      u64 last = ld_acquire(kp + KW - 1);
      if (last == KEY_EMPTY) {
        last = atomicCAS(kp + KW - 1, KEY_EMPTY, KEY_LOCKED);
        if (last == KEY_EMPTY) {  // we own the slot: write the other words, then publish the last one
#pragma unroll
          for (int i = 0; i < KW - 1; i++) __stcg(kp + i, key[i]);
          st_release(kp + KW - 1, key[KW - 1]);
          break;
        }
      }
      if (last == KEY_LOCKED) continue;  // owner is between its CAS and its release store; look again
      bool same = (last == key[KW - 1]);
#pragma unroll
      for (int i = 0; i < KW - 1; i++) same = same && (__ldcg(kp + i) == key[i]);
      if (same) break;
      if (++probe >= (u32)MAX_PROBE) { slot = 0xFFFFFFFFu; break; }
      slot = (slot + 1) & cap_mask;
    }
  }
  probes = probe;
  return slot;
}

struct InsertArgs {
  const PartDesc *parts;  // [gridDim.y]
  const SegDesc *segs;
  int k;
  u32 *overflow;          // [num_partitions] set to 1 when a table ran out of probe slots
  u32 *max_probe;         // atomicMax
};

template <int KW, bool CTG>
__global__ void __launch_bounds__(INS_THREADS) insert_kernel(const InsertArgs a) {
  constexpr int VW = CTG ? VW_CTG : VW_READS;
  const PartDesc pd = a.parts[blockIdx.y];
  __shared__ u32 pref[CH + 1];
  __shared__ u32 s_woff[CH];
  __shared__ u32 s_nbd[CH];  // nb | depth << 16
  __shared__ u32 warp_tot[INS_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int k = a.k;
  const int nw = (k + 31) >> 5;
  const u64 last_mask = (k & 31) ? (~0ULL << (64 - 2 * (k & 31))) : ~0ULL;
  u32 my_max_probe = 0;

  for (u32 chunk = blockIdx.x; chunk < pd.nchunks; chunk += gridDim.x) {
    u32 sidx = pd.seg_begin;
    while (sidx + 1 < pd.seg_end && a.segs[sidx + 1].chunk0 <= chunk) sidx++;
    const SegDesc sg = a.segs[sidx];
    const u32 first = (chunk - sg.chunk0) * CH;
    const u32 n = min((u32)CH, sg.nsup - first);
    u32 nk = 0;
    if ((u32)tid < n) {
      const SupHdr h = sg.hdr[first + tid];
      nk = (u32)h.nb - (u32)k - 1u;
      s_woff[tid] = h.woff;
      s_nbd[tid] = (u32)h.nb | ((u32)h.depth << 16);
    }
    // block exclusive scan of nk
    u32 incl = nk;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const u32 y = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += y;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    u32 woff = 0;
    for (int i = 0; i < wid; i++) woff += warp_tot[i];
    pref[tid] = woff + incl - nk;
    if (tid == INS_THREADS - 1) pref[CH] = woff + incl;
    __syncthreads();
    const u32 total = pref[CH];

    for (u32 t = tid; t < total; t += INS_THREADS) {
      // largest s with pref[s] <= t
      int lo = 0, hi = (int)n - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (pref[mid] <= t) lo = mid; else hi = mid - 1;
      }
      const int j = (int)(t - pref[lo]) + 1;  // k-mer index inside the supermer, 1-based (kcount_cpu.cpp:319)
      const u32 nb = s_nbd[lo] & 0xFFFFu, depth = s_nbd[lo] >> 16;
      const u32 *p = sg.payload + s_woff[lo];
      const u32 *pv = p + ((nb + 15) >> 4);
      u64 fwd[KW], rc[KW];
#pragma unroll
      for (int i = 0; i < KW; i++) fwd[i] = (i < nw) ? fetch64(p, j + 32 * i) : 0ULL;
#pragma unroll
      for (int i = 0; i < KW; i++)
        if (i == nw - 1) fwd[i] &= last_mask;
      const int lcode = fetch_base(p, j - 1), lval = fetch_bit(pv, j - 1);
      const int rcode = fetch_base(p, j + k), rval = fetch_bit(pv, j + k);
      revcomp_words<KW>(fwd, rc, k);
      int l, r;
      u64 key[KW];
      const bool use_rc = words_less<KW>(rc, fwd);  // strict: palindromes keep the forward orientation (kcount_cpu.cpp:326-331)
#pragma unroll
      for (int i = 0; i < KW; i++) key[i] = use_rc ? rc[i] : fwd[i];
      if (use_rc) {  // swap + complement the extensions (kcount_cpu.cpp:328-330)
        l = rval ? 3 - rcode : EXT_NONE;
        r = lval ? 3 - lcode : EXT_NONE;
      } else {
        l = lval ? lcode : EXT_NONE;
        r = rval ? rcode : EXT_NONE;
      }
      const u64 h = murmur64_words<KW>(key);  // Kmer::hash (kmer.cpp:465-468) as the slot hash
      u32 probes;
      const u32 slot = find_or_claim<KW>(pd.keys, pd.cap_mask, key, (u32)h & pd.cap_mask, probes);
      if (slot == 0xFFFFFFFFu) {
        a.overflow[pd.part] = 1u;
        continue;
      }
      my_max_probe = max(my_max_probe, probes);
      u32 *v = pd.vals + (size_t)slot * VW;
      if (!CTG || !sg.is_ctg) {
        atomicAdd(v + l, depth);       // reads: depth == 1 (kcount_cpu.cpp:75,348-352)
        atomicAdd(v + 5 + r, depth);
      } else {
        atomicAdd(v + 10, 1u);
        atomicMax(v + 11, ~depth);     // running minimum of the contig depths
        atomicOr(v + 12, 1u << (l * 5 + r));
      }
    }
    __syncthreads();
  }
  my_max_probe = __reduce_max_sync(0xffffffffu, my_max_probe);
  if (lane == 0 && my_max_probe) atomicMax(a.max_probe, my_max_probe);
}

struct FilterArgs {
  const PartDesc *parts;
  int k;
  int dmin_thres;
  double frac;            // 1.0 - DYN_MIN_DEPTH computed on the host in fp64
  const u32 *overflow;
  u64 *out_keys;          // [out_cap * KW]
  uint16_t *out_counts;
  char *out_left, *out_right;
  u64 out_cap;
  u64 *counters;          // [0] survivors cursor, [1] prefilter cursor, [2] distinct, [3] output overflow flag
  u64 *pf_keys;           // optional pre-filter dump (reads-pass counters), may be null
  uint16_t *pf_vals;
  u64 pf_cap;
  u64 *part_distinct;     // [num_partitions] distinct keys found per partition (sizing feedback)
};

template <int KW, bool CTG>
__global__ void __launch_bounds__(256) filter_kernel(const FilterArgs a) {
  constexpr int VW = CTG ? VW_CTG : VW_READS;
  const PartDesc pd = a.parts[blockIdx.y];
  if (a.overflow[pd.part]) return;  // this partition is re-run with a larger table; emit nothing now
  const u32 cap = pd.cap_mask + 1;
  const int lane = threadIdx.x & 31;
  u32 n_occ = 0;
  for (u32 s0 = blockIdx.x * blockDim.x; s0 < cap; s0 += gridDim.x * blockDim.x) {
    const u32 slot = s0 + threadIdx.x;  // cap is a multiple of blockDim.x (power of two >= 256)
    const u64 *kp = pd.keys + (size_t)slot * KW;
    const bool occ = kp[KW - 1] != KEY_EMPTY;
    bool survive = false;
    u32 count = 0;
    char le = 'X', re = 'X';
    u32 L[4] = {0, 0, 0, 0}, R[4] = {0, 0, 0, 0};
    u32 rcount = 0;
    if (occ) {
      const u32 *v = pd.vals + (size_t)slot * VW;
      u64 tot = 0;
#pragma unroll
      for (int i = 0; i < 5; i++) tot += v[i];
#pragma unroll
      for (int i = 0; i < 4; i++) { L[i] = sat16(v[i]); R[i] = sat16(v[5 + i]); }
      rcount = count = sat16(tot);
      if (CTG) {
        const u32 cn = v[10];
        if (cn) {
          bool use_ctg = (tot == 0) || (count == 1);  // new, or singleton read k-mer (kcount_cpu.cpp:365-371)
          if (!use_ctg) {
            const char l0 = get_ext(L, count, a.dmin_thres, a.frac), r0 = get_ext(R, count, a.dmin_thres, a.frac);
            use_ctg = (l0 == 'X' || l0 == 'F' || r0 == 'X' || r0 == 'F');  // non-UU read k-mer (kcount_cpu.cpp:373-376)
          }
          if (use_ctg) {
            // closed form of the sequential rules kcount_cpu.cpp:378-404 (DESIGN.md "contig pass")
            const u32 dmin = ~v[11], mask = v[12];
            const int pair = __ffs(mask) - 1;
            const int cl = pair / 5, cr = pair % 5;
#pragma unroll
            for (int i = 0; i < 4; i++) L[i] = R[i] = 0;
            if (cn == 1) {
              count = dmin;
            } else {
              const bool agree = (__popc(mask) == 1) && cl < 4 && cr < 4 && (int)dmin >= a.dmin_thres;
              count = agree ? dmin : 0;
            }
            if (cl < 4) L[cl] = count;
            if (cr < 4) R[cr] = count;
          }
        }
      }
      if (count >= 2) {
        le = get_ext(L, count, a.dmin_thres, a.frac);
        re = get_ext(R, count, a.dmin_thres, a.frac);
        survive = !(le == 'X' && re == 'X');
      }
    }
    const u32 occ_mask = __ballot_sync(0xffffffffu, occ);
    const u32 sv_mask = __ballot_sync(0xffffffffu, survive);
    n_occ += __popc(occ_mask);
    if (sv_mask) {
      u64 base = 0;
      if (lane == 0) base = atomicAdd(&a.counters[0], (u64)__popc(sv_mask));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (survive) {
        const u64 idx = base + __popc(sv_mask & ((1u << lane) - 1));
        if (idx < a.out_cap) {
#pragma unroll
          for (int i = 0; i < KW; i++) a.out_keys[idx * KW + i] = kp[i];
          a.out_counts[idx] = (uint16_t)count;
          a.out_left[idx] = le;
          a.out_right[idx] = re;
        } else {
          a.counters[3] = 1;
        }
      }
    }
    if (a.pf_keys) {
      const bool pf = occ && rcount > 0;  // k-mers seen in reads (the table as it stands after the reads pass)
      const u32 pf_mask = __ballot_sync(0xffffffffu, pf);
      if (pf_mask) {
        u64 base = 0;
        if (lane == 0) base = atomicAdd(&a.counters[1], (u64)__popc(pf_mask));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (pf) {
          const u64 idx = base + __popc(pf_mask & ((1u << lane) - 1));
          if (idx < a.pf_cap) {
            const u32 *v = pd.vals + (size_t)slot * VW;
#pragma unroll
            for (int i = 0; i < KW; i++) a.pf_keys[idx * KW + i] = kp[i];
            uint16_t *o = a.pf_vals + idx * 9;
            o[0] = (uint16_t)rcount;
#pragma unroll
            for (int i = 0; i < 4; i++) { o[1 + i] = (uint16_t)sat16(v[i]); o[5 + i] = (uint16_t)sat16(v[5 + i]); }
          } else {
            a.counters[3] = 1;
          }
        }
      }
    }
  }
  if (lane == 0 && n_occ) {
    atomicAdd(&a.counters[2], (u64)n_occ);
    atomicAdd(&a.part_distinct[pd.part], (u64)n_occ);
  }
}

}  // namespace kc

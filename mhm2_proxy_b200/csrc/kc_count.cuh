// Stage B — count: one CTA owns one partition of the shard's k-mer space at a time and keeps that partition's whole
// hash table in SHARED MEMORY (measured on B200, tools/ubench_atomics.cu: shared atomics 2.8 T/s chip-wide vs 0.19 T/s for
// L2-resident global RED, i.e. 15x). Per work item (partition, sub-pass):
//   1. stream the partition's supermers (from every parsed / received batch) through a small staging buffer,
//   2. expand them to k-mer instances, canonicalise, hash (MurmurHash3 = Kmer::hash), find-or-claim the key, bump the
//      extension counters with shared-memory atomics
//         replaces get_kmers_and_exts + KmerMapExts::insert + insert_supermer_from_read/_from_ctg
//         (reference src/kcount/kcount_cpu.cpp:231-267,307-406),
//   3. scan the dense key store: depth filter + extension choice + compaction of survivors straight to the output arrays
//         replaces HashTableInserter::insert_into_local_hashtable + ExtCounts::get_ext (kcount_cpu.cpp:173-182,490-528).
// The table never touches HBM: HBM traffic of this stage = supermers read once + survivors written once.
//
// Table layout in shared memory ("index table + dense store"):
//   tab[tab_size]  u32   open addressing, entry = fingerprint(19 bits) | (dense index + 1)(13 bits), 0 = empty
//   keys[key_cap]  KW x u64, in order of first arrival; last word all-ones = dead entry (lost a same-key race)
//   vals[key_cap]  VW x u32 counters:
//        narrow: 10 x u16 halves {left A,C,G,T,none, right A,C,G,T, spare}; a half that wraps flags the item `wide`
//        wide  : 9 x u32 (same order, no spare); used only for items that flagged (k-mers seen > 65535 times)
//        + contig fields when contigs are present: {n contig occurrences, ~min depth, (left,right) pair mask}
//   count = sum of the five left counters (every read instance increments exactly one); accumulating then clamping to
//   65535 equals the reference's saturating u16 adds because the adds are monotone.
// A partition with more distinct k-mers than key_cap is re-run as 2^n sub-passes, each taking the k-mers whose hash
// falls in its sub-range (the host grows n from the progress the failed item reports); nothing is ever dropped.
#pragma once
#include "kc_device.cuh"
#include "kc_parse.cuh"

namespace kc {

constexpr u64 KEY_EMPTY = 0xFFFFFFFFFFFFFFFFULL;  // last key word of a real k-mer is never all ones: it has >= 2 zero
                                                  // trailing bits because N_LONGS = k/32 + 1 (main.cpp:170)
constexpr int CNT_THREADS = 1024;
constexpr int EXP_CAP = 4096;      // k-mer instances per chunk
constexpr int STAGE_WORDS = 3072;  // supermer payload words per chunk
constexpr int SEG_CACHE = 64;      // segment descriptors cached per group
constexpr u32 IDX_BITS = 13, IDX_MASK = (1u << IDX_BITS) - 1u;  // dense index + 1 fits 13 bits: key_cap <= 8191
constexpr int VR_NARROW = 5, VR_WIDE = 9, V_CTG = 3;

struct SegDesc {  // one contiguous run of supermers of one partition inside one parsed / received batch
  const SupHdr *hdr;
  const u32 *payload;  // the run's payload region; SupHdr::woff is relative to it
  u32 nsup;
  u32 is_ctg;
  u32 sup0;  // index of the run's first supermer inside the partition's concatenation of runs
  u32 pad;
};

struct WorkItem {
  u32 seg_begin, seg_end;  // into CountArgs::segs
  u32 nsup;                // supermers of the partition
  u32 part;
  u32 sub, nsub;           // this item takes k-mers with ((hash >> 40) & (nsub - 1)) == sub; nsub is a power of two
  u32 pad[2];
};

struct ItemStatus {
  u32 overflow;  // ran out of key slots: emitted nothing, must be re-run with more sub-passes
  u32 progress;  // supermers consumed when it overflowed (sizes the retry)
  u32 nkeys;     // distinct k-mers found (valid when !overflow)
  u32 wide;      // a u16 counter wrapped: emitted nothing, must be re-run with wide counters
};

struct CountArgs {
  const WorkItem *items;
  const SegDesc *segs;
  ItemStatus *status;
  u32 nitems;
  u32 *next_item;  // dynamic scheduler
  int k;
  u32 tab_size, key_cap;
  int dmin_thres;
  double frac;     // 1.0 - DYN_MIN_DEPTH computed on the host in fp64
  u64 *out_keys;   // [out_cap * KW]
  uint16_t *out_counts;
  char *out_left, *out_right;
  u64 out_cap;
  u64 *counters;   // [0] survivors cursor, [1] prefilter cursor, [2] distinct, [3] output overflow flag, [4] max probe
  u64 *pf_keys;    // optional pre-filter dump (reads-pass counters), may be null
  uint16_t *pf_vals;
  u64 pf_cap;
};

__host__ __device__ inline size_t count_smem_bytes(int KW, int VW, u32 tab_size, u32 key_cap, bool ctg) {
  size_t b = (size_t)tab_size * 4 + (size_t)key_cap * (KW * 8 + VW * 4);
  b += (size_t)EXP_CAP * 2 + (size_t)(STAGE_WORDS + 4) * 4 + (size_t)CNT_THREADS * 4 * (ctg ? 2 : 1) + (size_t)SEG_CACHE * sizeof(SegDesc);
  return b;
}

template <int KW, bool WIDE, bool CTG>
__global__ void __launch_bounds__(CNT_THREADS, 1) count_kernel(const CountArgs a) {
  constexpr int VR = WIDE ? VR_WIDE : VR_NARROW;
  constexpr int VW = VR + (CTG ? V_CTG : 0);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  u32 *tab = (u32 *)smem_raw;                              // [tab_size], tab_size % 4 == 0
  u64 *keys = (u64 *)(tab + a.tab_size);                   // [key_cap * KW]
  u32 *vals = (u32 *)(keys + (size_t)a.key_cap * KW);      // [key_cap * VW]
  SegDesc *segc = (SegDesc *)(vals + (size_t)a.key_cap * VW + ((a.key_cap * VW) & 1));  // 8-byte aligned
  u32 *stage = (u32 *)(segc + SEG_CACHE);                  // [STAGE_WORDS + 4]
  u32 *meta = stage + STAGE_WORDS + 4;                     // [CNT_THREADS] staged word offset | nb << 16
  u32 *meta2 = meta + CNT_THREADS;                         // [CNT_THREADS] (CTG only) depth | is_ctg << 16
  uint16_t *expl = (uint16_t *)(meta + CNT_THREADS * (CTG ? 2 : 1));  // [EXP_CAP] supermer-in-chunk | (k-mer - 1) << 10
  __shared__ u64 s_obase[2];
  __shared__ u32 s_item, s_nkeys, s_over, s_wide, s_total, s_live;
  __shared__ u32 warp_k[32], warp_w[32], warp_c[32], warp_p[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int k = a.k;
  const int nw = (k + 31) >> 5;
  const u64 last_mask = (k & 31) ? (~0ULL << (64 - 2 * (k & 31))) : ~0ULL;
  u32 my_max_probe = 0;

  for (;;) {
    __syncthreads();
    if (tid == 0) {
      s_item = atomicAdd(a.next_item, 1u);
      s_nkeys = 0;
      s_over = 0;
      s_wide = 0;
      s_live = 0;
    }
    for (u32 i = tid; i < a.tab_size / 4; i += CNT_THREADS) ((uint4 *)tab)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const u32 item_idx = s_item;
    if (item_idx >= a.nitems) break;
    const WorkItem it = a.items[item_idx];
    const u32 nseg = it.seg_end - it.seg_begin;
    u32 consumed = 0;
    bool over = false;

    for (u32 sg0 = 0; sg0 < nseg && !over; sg0 += SEG_CACHE) {
      const int ns = (int)min((u32)SEG_CACHE, nseg - sg0);
      __syncthreads();
      if (tid < ns) segc[tid] = a.segs[it.seg_begin + sg0 + tid];
      __syncthreads();
      const u32 end_sup = segc[ns - 1].sup0 + segc[ns - 1].nsup;
      u32 c0 = segc[0].sup0;
      int seg_hint = 0;  // first segment that can contain c0
      while (c0 < end_sup) {
        // ---- a. one supermer header per thread ----
        const u32 g = c0 + (u32)tid;
        u32 nk = 0, nwd = 0, nb = 0, depth = 1, isc = 0;
        const u32 *src = nullptr;
        if (g < end_sup) {
          int si = seg_hint;
          while (si + 1 < ns && segc[si + 1].sup0 <= g) si++;
          const SupHdr h = segc[si].hdr[g - segc[si].sup0];
          nb = h.nb;
          depth = h.depth;
          isc = segc[si].is_ctg;
          nk = nb - (u32)k - 1u;
          nwd = ((nb + 15) >> 4) + ((nb + 31) >> 5);
          src = segc[si].payload + h.woff;
        }
        // ---- b. block inclusive scans of k-mer and word counts; how many supermers fit this chunk ----
        u32 ik = nk, iw = nwd;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const u32 yk = __shfl_up_sync(0xffffffffu, ik, d), yw = __shfl_up_sync(0xffffffffu, iw, d);
          if (lane >= d) { ik += yk; iw += yw; }
        }
        if (lane == 31) { warp_k[wid] = ik; warp_w[wid] = iw; }
        __syncthreads();
        u32 bk = 0, bw = 0;
        for (int i = 0; i < wid; i++) { bk += warp_k[i]; bw += warp_w[i]; }
        ik += bk;
        iw += bw;
        const bool fits = (g < end_sup) && ik <= (u32)EXP_CAP && iw <= (u32)STAGE_WORDS;
        const u32 n_take = (u32)__syncthreads_count(fits);  // prefix sums are monotone: the fitting ones are a prefix
        // ---- c. stage the payload, expand to instances ----
        if (fits) {
          const u32 woff = iw - nwd, koff = ik - nk;
          meta[tid] = woff | (nb << 16);
          if (CTG) meta2[tid] = depth | (isc << 16);
          for (u32 x = 0; x < nwd; x++) stage[woff + x] = __ldg(src + x);
          for (u32 j = 0; j < nk; j++) expl[koff + j] = (uint16_t)((u32)tid | (j << 10));
        }
        if (tid == (int)n_take - 1) s_total = ik;  // total instances of the chunk
        __syncthreads();
        const u32 total = s_total;

        // ---- d. every k-mer instance of the chunk ----
        for (u32 t = tid; t < total; t += CNT_THREADS) {
          const u32 e = expl[t];
          const u32 si = e & 1023u;
          const int j = (int)(e >> 10) + 1;  // k-mer index inside the supermer, 1-based (kcount_cpu.cpp:319)
          const u32 mt = meta[si];
          const u32 snb = mt >> 16;
          const u32 *p = stage + (mt & 0xFFFFu);
          const u32 *pv = p + ((snb + 15) >> 4);
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
          const u64 h = murmur64_words<KW>(key);  // Kmer::hash (kmer.cpp:465-468)
          if (it.nsub > 1 && ((u32)(h >> 40) & (it.nsub - 1)) != it.sub) continue;
          // ---- find or claim ----
          const u32 fp = (u32)(h >> 32) & ~IDX_MASK;
          u32 slot = __umulhi((u32)h, a.tab_size);
          u32 mine = 0xFFFFFFFFu, idx = 0xFFFFFFFFu, probes = 0;
          bool claimed = false;
          for (;;) {
            u32 en = ((volatile u32 *)tab)[slot];
            if (en == 0) {
              if (mine == 0xFFFFFFFFu) {  // reserve a dense entry (warp-aggregated), fill it, then publish it
                const unsigned am = __activemask();
                const int leader = __ffs(am) - 1;
                u32 base = 0;
                if (lane == leader) base = atomicAdd(&s_nkeys, (u32)__popc(am));
                base = __shfl_sync(am, base, leader);
                mine = base + (u32)__popc(am & ((1u << lane) - 1u));
                if (mine >= a.key_cap) {
                  s_over = 1;
                  break;
                }
#pragma unroll
                for (int i = 0; i < KW; i++) keys[(size_t)mine * KW + i] = key[i];
#pragma unroll
                for (int i = 0; i < VW; i++) vals[(size_t)mine * VW + i] = 0u;
                __threadfence_block();
              }
              en = atomicCAS(&tab[slot], 0u, fp | (mine + 1u));
              if (en == 0) {
                idx = mine;
                claimed = true;
                break;
              }
            }
            if ((en & ~IDX_MASK) == fp) {
              const u32 cand = (en & IDX_MASK) - 1u;
              bool same = true;
#pragma unroll
              for (int i = 0; i < KW; i++) same = same && (((volatile u64 *)keys)[(size_t)cand * KW + i] == key[i]);
              if (same) {
                idx = cand;
                break;
              }
            }
            probes++;
            slot = (slot + 1 == a.tab_size) ? 0u : slot + 1;
          }
          if (mine != 0xFFFFFFFFu && !claimed && mine < a.key_cap) keys[(size_t)mine * KW + KW - 1] = KEY_EMPTY;  // lost a same-key race
          if (idx == 0xFFFFFFFFu) continue;  // out of key slots; the item is re-run
          my_max_probe = max(my_max_probe, probes);
          u32 *v = vals + (size_t)idx * VW;
          const u32 mt2 = CTG ? meta2[si] : 0u;
          if (!CTG || !(mt2 >> 16)) {  // a read instance (count 1: kcount_cpu.cpp:75,348-352)
            if (WIDE) {
              atomicAdd(v + l, 1u);
              if (r < 4) atomicAdd(v + 5 + r, 1u);
            } else {
              const u32 shl = (u32)(l & 1) * 16u;
              u32 old = atomicAdd(v + (l >> 1), 1u << shl);
              if (((old >> shl) & 0xFFFFu) == 0xFFFFu) s_wide = 1;
              if (r < 4) {
                const u32 hr = 5u + (u32)r, shr = (hr & 1u) * 16u;
                old = atomicAdd(v + (hr >> 1), 1u << shr);
                if (((old >> shr) & 0xFFFFu) == 0xFFFFu) s_wide = 1;
              }
            }
          } else {  // a contig instance: remembered as (occurrences, min depth, set of extension pairs), DESIGN.md "contig pass"
            atomicAdd(v + VR, 1u);
            atomicMax(v + VR + 1, ~(mt2 & 0xFFFFu));
            atomicOr(v + VR + 2, 1u << (l * 5 + r));
          }
        }
        __syncthreads();
        c0 += n_take;
        consumed += n_take;
        if (s_over) {
          over = true;
          break;
        }
        while (seg_hint + 1 < ns && segc[seg_hint + 1].sup0 <= c0) seg_hint++;
      }
    }
    __syncthreads();
    const bool wide_needed = !WIDE && s_wide;
    const u32 n = min(s_nkeys, a.key_cap);
    if (tid == 0) {
      ItemStatus st;
      st.overflow = over ? 1u : 0u;
      st.progress = consumed;
      st.nkeys = n;
      st.wide = wide_needed ? 1u : 0u;
      a.status[item_idx] = st;
    }
    if (over || wide_needed) continue;

    // ---- e. depth filter, extension choice, compaction (kcount_cpu.cpp:490-528) ----
    for (u32 b0 = 0; b0 < n; b0 += CNT_THREADS) {
      const u32 idx = b0 + (u32)tid;
      const u64 *kp = keys + (size_t)idx * KW;
      const bool occ = idx < n && kp[KW - 1] != KEY_EMPTY;
      bool survive = false;
      u32 count = 0, rcount = 0;
      char le = 'X', re = 'X';
      u32 L[4] = {0, 0, 0, 0}, R[4] = {0, 0, 0, 0};
      if (occ) {
        const u32 *v = vals + (size_t)idx * VW;
        u64 tot;
        if (WIDE) {
          tot = (u64)v[0] + v[1] + v[2] + v[3] + v[4];
#pragma unroll
          for (int i = 0; i < 4; i++) { L[i] = sat16(v[i]); R[i] = sat16(v[5 + i]); }
        } else {
          const u32 w0 = v[0], w1 = v[1], w2 = v[2], w3 = v[3], w4 = v[4];
          L[0] = w0 & 0xFFFFu; L[1] = w0 >> 16; L[2] = w1 & 0xFFFFu; L[3] = w1 >> 16;
          const u32 lnone = w2 & 0xFFFFu;
          R[0] = w2 >> 16; R[1] = w3 & 0xFFFFu; R[2] = w3 >> 16; R[3] = w4 & 0xFFFFu;
          tot = (u64)L[0] + L[1] + L[2] + L[3] + lnone;
        }
        rcount = count = sat16(tot);
        if (CTG) {
          const u32 cn = v[VR];
          if (cn) {
            bool use_ctg = (tot == 0) || (count == 1);  // new, or singleton read k-mer (kcount_cpu.cpp:365-371)
            if (!use_ctg) {
              const char l0 = get_ext(L, count, a.dmin_thres, a.frac), r0 = get_ext(R, count, a.dmin_thres, a.frac);
              use_ctg = (l0 == 'X' || l0 == 'F' || r0 == 'X' || r0 == 'F');  // non-UU read k-mer (kcount_cpu.cpp:373-376)
            }
            if (use_ctg) {
              // closed form of the sequential rules kcount_cpu.cpp:378-404 (DESIGN.md "contig pass")
              const u32 dmin = ~v[VR + 1], mask = v[VR + 2];
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
      const bool pf = a.pf_keys && occ && rcount > 0;  // k-mers seen in reads (the table as it stands after the reads pass)
      const u32 occ_mask = __ballot_sync(0xffffffffu, occ);
      const u32 sv_mask = __ballot_sync(0xffffffffu, survive);
      const u32 pf_mask = __ballot_sync(0xffffffffu, pf);
      if (lane == 0 && occ_mask) atomicAdd(&s_live, (u32)__popc(occ_mask));
      if (lane == 0) { warp_c[wid] = __popc(sv_mask); warp_p[wid] = __popc(pf_mask); }
      __syncthreads();
      if (wid == 0) {  // one warp scans the 32 warp totals and reserves output space for the block
        const u32 c = warp_c[lane], p = warp_p[lane];
        u32 ic = c, ip = p;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const u32 yc = __shfl_up_sync(0xffffffffu, ic, d), yp = __shfl_up_sync(0xffffffffu, ip, d);
          if (lane >= d) { ic += yc; ip += yp; }
        }
        warp_c[lane] = ic - c;
        warp_p[lane] = ip - p;
        const u32 totc = __shfl_sync(0xffffffffu, ic, 31), totp = __shfl_sync(0xffffffffu, ip, 31);
        if (lane == 0) {
          s_obase[0] = totc ? atomicAdd(&a.counters[0], (u64)totc) : 0ull;
          s_obase[1] = totp ? atomicAdd(&a.counters[1], (u64)totp) : 0ull;
        }
      }
      __syncthreads();
      if (survive) {
        const u64 o = s_obase[0] + warp_c[wid] + __popc(sv_mask & ((1u << lane) - 1u));
        if (o < a.out_cap) {
#pragma unroll
          for (int i = 0; i < KW; i++) a.out_keys[o * KW + i] = kp[i];
          a.out_counts[o] = (uint16_t)count;
          a.out_left[o] = le;
          a.out_right[o] = re;
        } else {
          a.counters[3] = 1;
        }
      }
      if (pf) {
        const u64 o = s_obase[1] + warp_p[wid] + __popc(pf_mask & ((1u << lane) - 1u));
        if (o < a.pf_cap) {
          const u32 *v = vals + (size_t)idx * VW;
#pragma unroll
          for (int i = 0; i < KW; i++) a.pf_keys[o * KW + i] = kp[i];
          uint16_t *ov = a.pf_vals + o * 9;
          ov[0] = (uint16_t)rcount;
          if (WIDE) {
#pragma unroll
            for (int i = 0; i < 4; i++) { ov[1 + i] = (uint16_t)sat16(v[i]); ov[5 + i] = (uint16_t)sat16(v[5 + i]); }
          } else {
            const u32 w0 = v[0], w1 = v[1], w2 = v[2], w3 = v[3], w4 = v[4];
            ov[1] = (uint16_t)w0; ov[2] = (uint16_t)(w0 >> 16); ov[3] = (uint16_t)w1; ov[4] = (uint16_t)(w1 >> 16);
            ov[5] = (uint16_t)(w2 >> 16); ov[6] = (uint16_t)w3; ov[7] = (uint16_t)(w3 >> 16); ov[8] = (uint16_t)w4;
          }
        } else {
          a.counters[3] = 1;
        }
      }
    }
    __syncthreads();
    if (tid == 0 && s_live) atomicAdd(&a.counters[2], (u64)s_live);
  }
  my_max_probe = __reduce_max_sync(0xffffffffu, my_max_probe);
  if (lane == 0 && my_max_probe) atomicMax(&a.counters[4], (u64)my_max_probe);
}

}  // namespace kc

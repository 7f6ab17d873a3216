// Stage B — count: one CTA owns one partition of the shard's k-mer space at a time and keeps that partition's whole
// hash table in SHARED MEMORY (measured on B200, tools/ubench_atomics.cu: shared atomics 2.8 T/s chip-wide vs 0.19 T/s for
// L2-resident global RED, i.e. 15x). Per work item (partition, sub-pass):
//   1. 32 independent pipelines — every warp is one — each stream 1/32 of the partition's supermer records through their
//      own staging buffer (no barrier inside a work item, only __syncwarp) and all insert into ONE table;
//   2. records are expanded to k-mer instances, canonicalised, hashed (internal slot hash, kc_device.cuh), looked up with ONE
//      128-bit bucket load (instances whose key is not there yet wait in a per-warp miss queue and are found or claimed 32 at
//      a time), and the extension counters bumped with shared-memory atomics
//         replaces get_kmers_and_exts + KmerMapExts::insert + insert_supermer_from_read/_from_ctg
//         (reference src/kcount/kcount_cpu.cpp:231-267,307-406);
//   3. the dense key store is scanned: depth filter + extension choice + compaction of survivors straight to the output
//         replaces HashTableInserter::insert_into_local_hashtable + ExtCounts::get_ext (kcount_cpu.cpp:173-182,490-528).
// The table never touches HBM: HBM traffic of this stage = supermer records read once + survivors written once.
//
// Table layout in shared memory ("bucketised index + dense store"):
//   tab[nbuckets]  uint4  open addressing over 16-byte buckets of four entries; entry = fingerprint(19 bits) |
//                         (dense index + 1)(13 bits), 0 = empty. One 128-bit load resolves almost every lookup.
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
#include "kc_extract.cuh"

namespace kc {

constexpr u64 KEY_EMPTY = 0xFFFFFFFFFFFFFFFFULL;  // last key word of a real k-mer is never all ones: it has >= 2 zero
                                                  // trailing bits because N_LONGS = k/32 + 1 (main.cpp:170)
#ifndef KC_CNT_THREADS
#define KC_CNT_THREADS 1024
#endif
constexpr int CNT_THREADS = KC_CNT_THREADS;     // threads per CTA; CNT_CTAS_PER_SM CTAs share an SM (and its shared memory)
constexpr int CNT_CTAS_PER_SM = 1024 / CNT_THREADS;
constexpr int CNT_GROUPS = CNT_THREADS / 32;    // independent pipelines per CTA: every WARP is one (no barrier inside a work item's
constexpr int GRP_THREADS = CNT_THREADS / CNT_GROUPS;  // stream of supermers: __syncwarp only)
constexpr int GRP_STAGE_MAX = 256;     // staged record words per pipeline and chunk: min(32 records, this many words) + 4
constexpr int GRP_EXP = 160;           // k-mer instances per pipeline and chunk
static_assert(GRP_THREADS == 32, "a pipeline is one warp");
__host__ __device__ inline int stage_words(int rw) { return (32 * rw < GRP_STAGE_MAX ? 32 * rw : GRP_STAGE_MAX) + 4; }
constexpr int SEG_CACHE = 64;          // segment descriptors cached per round
// Miss queue (keys of <= 2 words): an instance whose k-mer is not where the one-load lookup expects it (new key, key in a
// neighbouring bucket, shared fingerprint) parks its KEY and counter slots in a per-warp ring and goes on; whenever 32 are
// waiting, the warp runs them through find-or-claim with every lane busy. (Until r2_v6 the instance's list entry was kept and
// the k-mer extracted again at the end of its chunk: ncu showed that pass at 12 of 32 lanes and the second extraction as 9 %
// of all instructions.)
#ifndef KC_MISS_QUEUE
#define KC_MISS_QUEUE 1
#endif
constexpr int QCAP = 64;               // ring entries per warp: < 32 waiting + up to 32 new ones
#ifndef KC_DRAIN_MIN
#define KC_DRAIN_MIN 32
#endif
constexpr unsigned DRAIN_MIN = KC_DRAIN_MIN;  // waiting instances that trigger a pass through find-or-claim (<= 32)
__host__ __device__ constexpr bool miss_queue_on(int KW) { return KC_MISS_QUEUE && KW <= 2; }
__host__ __device__ constexpr size_t miss_queue_bytes(int KW, bool ctg) {  // all warps of the CTA
  return miss_queue_on(KW) ? (size_t)CNT_GROUPS * QCAP * ((size_t)KW * 8 + (ctg ? 4 : 1)) : 0;
}
constexpr u32 IDX_BITS = 13, IDX_MASK = (1u << IDX_BITS) - 1u;  // dense index + 1 fits 13 bits: key_cap <= 8190
constexpr int VR_NARROW = 5, VR_WIDE = 9, V_CTG = 3;

struct SegDesc {  // one contiguous run of supermer records of one partition inside one parsed / received batch
  const u32 *rec;
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
  u32 heavy;               // far above the mean: scheduled through CountArgs::heavy ahead of everything else
  u32 ninst;               // k-mer instances of the partition (saturating): < 65536 means no u16 counter can wrap
};

struct BlockDesc {  // one parsed or received batch as the item builder sees it (all pointers are device memory)
  const u32 *rec;
  const BucketHist *hist;  // per bucket
  const u64 *base;         // exclusive scan of hist[].nsup
  u32 first;               // bucket of this shard's partition 0 inside hist/base
  u32 is_ctg;
};

// counters[] slots shared by the item builder, the count kernel and the host
enum { CTR_SURV = 0, CTR_PF = 1, CTR_DISTINCT = 2, CTR_ERR = 3, CTR_MAXPROBE = 4, CTR_RETRY = 5, CTR_READ_INST = 6, CTR_CTG_INST = 7,
       CTR_NHEAVY = 8, CTR_SUP = 9, CTR_WORDS = 16 };

struct ItemStatus {
  u32 overflow;  // ran out of key slots: emitted nothing, must be re-run with more sub-passes
  u32 progress;  // supermers consumed when it overflowed (sizes the retry); probe run without overflow: distinct k-mers
  u32 nkeys;     // distinct k-mers found (valid when !overflow)
  u32 wide;      // a u16 counter wrapped: emitted nothing, must be re-run with wide counters
};

struct RetryRec {  // an item that has to be run again, as the kernel saw it (the host reads only these)
  WorkItem item;
  ItemStatus st;
};

struct CountArgs {
  const WorkItem *items;
  const SegDesc *segs;
  ItemStatus *status;
  const u64 *heavy;  // [n_heavy] (nsup << 32 | item index), heaviest first
  u32 nitems, n_heavy;
  u32 measure_only;  // probe run: build the tables, report ItemStatus (distinct k-mers), emit nothing
  RetryRec *retry;   // [nitems] items of this launch that overflowed / wrapped, slot = CTR_RETRY value - retry_base
  u64 retry_base;
  int k;
  int rw, nwb;     // record geometry (kc_parse.cuh)
  int w;           // k-mers per supermer <= w (window of the minimizer), mean (w + 1) / 2
  u32 tab_size, key_cap;  // tab_size u32 entries (multiple of 4)
  int dmin_thres;
  double frac;     // 1.0 - DYN_MIN_DEPTH computed on the host in fp64
  u64 *out_keys;   // [out_cap * KW]
  uint16_t *out_counts;
  char *out_left, *out_right;
  u64 out_cap;
  u64 *counters;   // CTR_* slots
  u64 *pf_keys;    // optional pre-filter dump (reads-pass counters), may be null
  uint16_t *pf_vals;
  u64 pf_cap;
};

// dynamic shared memory of the count kernel: per-warp staging buffers (first: their addresses are cheap to derive), index
// table, dense key store, counters. The instance lists and the segment cache are static arrays (COUNT_STATIC_SMEM).
__host__ __device__ inline size_t count_smem_bytes(int KW, int VW, int rw, u32 tab_size, u32 key_cap) {
  const bool ctg = VW == VR_NARROW + V_CTG || VW == VR_WIDE + V_CTG;
  return (size_t)CNT_GROUPS * stage_words(rw) * 4 + miss_queue_bytes(KW, ctg) + (size_t)tab_size * 4 +
         (size_t)key_cap * (KW * 8 + VW * 4) + 8;
}
constexpr size_t COUNT_STATIC_SMEM = (13 * 1024) / CNT_CTAS_PER_SM + (CNT_CTAS_PER_SM > 1 ? 1536 : 0);  // >= the kernel's static arrays (ptxas -v: 12.1 KB at 1024 threads)

__device__ __forceinline__ uint4 lds128_volatile(u32 a) {  // a: shared-space address (converted once per kernel)
  uint4 v;
  asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
  return v;
}

__device__ __forceinline__ u64 lds64_volatile(u32 a) {
  u64 v;
  asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts64(u32 a, u64 v) {
  asm volatile("st.shared.u64 [%0], %1;" ::"r"(a), "l"(v) : "memory");
}
__device__ __forceinline__ void sts32(u32 a, u32 v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void sts8(u32 a, u32 v) {
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ u32 lds8_volatile(u32 a) {
  u32 v;
  asm volatile("ld.volatile.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ u32 lds32_volatile(u32 a) {
  u32 v;
  asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts16(u32 a, u32 v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ void atoms_add(u32 a, u32 v) {  // shared-space address; no result wanted (RED-style)
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ u32 atoms_add_ret(u32 a, u32 v) {
  u32 old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(a), "r"(v) : "memory");
  return old;
}

// 0..5 -> A C G T F X
__device__ __forceinline__ char ext_char(u32 c) { return c < 4u ? (char)((0x54474341u >> (8u * c)) & 0xFFu) : (c == 4u ? 'F' : 'X'); }

// KC: k as a compile-time constant (0 = any k at run time, taken from CountArgs), WC: the supermer window that goes with it
// (kc_extract.cuh)
template <int KW, bool WIDE, bool CTG, int KC, int WC>
__global__ void __launch_bounds__(CNT_THREADS, CNT_CTAS_PER_SM) count_kernel(const CountArgs a) {
  constexpr int VR = WIDE ? VR_WIDE : VR_NARROW;
  constexpr int VW = VR + (CTG ? V_CTG : 0);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int k = KC ? KC : a.k;
  const int rw = a.rw;  // run-time: the host may pad 16-byte records to whole sectors (kc_create)
  const int sw = stage_words(rw);
  constexpr bool QUEUE = miss_queue_on(KW);
  constexpr u32 QMB = CTG ? 4u : 1u;                       // bytes of a queued instance's counter slots (+ contig header bits)
  u32 *stage_all = (u32 *)smem_raw;                        // [CNT_GROUPS][sw], 16-byte aligned rows (sw % 4 == 0)
  unsigned char *mq_all = (unsigned char *)(stage_all + CNT_GROUPS * sw);  // miss queues: keys [CNT_GROUPS][QCAP][KW], then slots
  u32 *tab = (u32 *)(mq_all + miss_queue_bytes(KW, CTG));  // [tab_size], 16-byte buckets
  u64 *keys = (u64 *)(tab + a.tab_size);                   // [key_cap * KW]
  u32 *vals = (u32 *)(keys + (size_t)a.key_cap * KW);      // [key_cap * VW]
  __shared__ uint16_t expl_all[CNT_GROUPS][GRP_EXP];
  __shared__ __align__(8) SegDesc segc[SEG_CACHE];
  __shared__ u64 s_obase[2];
  __shared__ u32 s_ocur[2];
  __shared__ __align__(16) WorkItem s_work;
  __shared__ u32 s_nkeys, s_over, s_wide, s_live, s_prev, s_consumed, s_widx;
  __shared__ u32 warp_c[32], warp_p[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int grp = wid, gt = lane;  // pipeline = warp
  u32 *stage = stage_all + grp * sw;
  uint16_t *expl = expl_all[grp];
  const u32 nbuckets = a.tab_size >> 2;
  const u32 tab_s = (u32)__cvta_generic_to_shared(tab), keys_s = (u32)__cvta_generic_to_shared(keys),
            vals_s = (u32)__cvta_generic_to_shared(vals), stage_s = (u32)__cvta_generic_to_shared(stage),
            expl_s = (u32)__cvta_generic_to_shared(expl);
  const u32 mqk_s = (u32)__cvta_generic_to_shared(mq_all) + (u32)grp * (u32)(QCAP * KW * 8);
  const u32 mqm_s = (u32)__cvta_generic_to_shared(mq_all) + (u32)(CNT_GROUPS * QCAP * KW * 8) + (u32)grp * (u32)QCAP * QMB;
  const int nwb = KC ? kc_nwb_of(KC, WC) : a.nwb;
  const int w_ = KC ? WC : a.w;
  const int ncopy = (1 + nwb + ((w_ + k + 1 + 31) >> 5) + 3) >> 2;  // 16-byte pieces of a record that hold data (the rest is padding)
  // supermers staged per chunk: what the staging buffer holds, and no more than the instance list will take (a supermer has
  // (w + 1) / 2 k-mers on average; staging records that do not fit the list means staging them again)
  const int sup_per_chunk = min(min(GRP_THREADS, GRP_STAGE_MAX / rw), 2 * GRP_EXP / (w_ + 1) + 3);
  u32 my_max_probe = 0;
  // Work items are dealt round-robin to the CTAs: first the heavy list (outlier partitions, heaviest first, so that none of
  // them starts last and becomes the tail), then every item in partition order, skipping the ones the heavy list covered.
  // The next item's descriptor is fetched while the current one is being processed.
  const u32 nvirt = a.n_heavy + a.nitems;
  if (tid == 0) {
    s_prev = a.key_cap;  // shared memory starts undefined: the first item clears every counter
    if (blockIdx.x < nvirt) {
      const u32 i0 = blockIdx.x < a.n_heavy ? (u32)a.heavy[blockIdx.x] : blockIdx.x - a.n_heavy;
      s_work = a.items[i0];
      s_widx = i0;
    }
  }

  for (u32 virt = blockIdx.x; virt < nvirt; virt += gridDim.x) {
    __syncthreads();
    const u32 prev = s_prev;
    const WorkItem it = s_work;
    const u32 item_idx = s_widx;
    const bool have_nx = tid == 0 && virt + gridDim.x < nvirt;
    uint4 nx0 = make_uint4(0, 0, 0, 0), nx1 = nx0;
    u32 nx_idx = 0;
    if (have_nx) {
      const u32 nv = virt + gridDim.x;
      nx_idx = nv < a.n_heavy ? (u32)__ldg(a.heavy + nv) : nv - a.n_heavy;
      const uint4 *pn = (const uint4 *)(a.items + nx_idx);
      nx0 = __ldg(pn);
      nx1 = __ldg(pn + 1);
    }
    __syncthreads();
    if (it.nsup == 0 || (it.heavy && virt >= a.n_heavy)) {  // empty partition, or already done through the heavy list
      if (have_nx) {
        ((uint4 *)&s_work)[0] = nx0;
        ((uint4 *)&s_work)[1] = nx1;
        s_widx = nx_idx;
      }
      continue;
    }
    if (tid == 0) {
      s_nkeys = 0;
      s_over = 0;
      s_wide = 0;
      s_live = 0;
      s_consumed = 0;
    }
    for (u32 i = tid; i < nbuckets; i += CNT_THREADS) ((uint4 *)tab)[i] = make_uint4(0, 0, 0, 0);
    for (u32 i = tid; i < prev * VW; i += CNT_THREADS) vals[i] = 0;  // counters of the slots the previous item used
    __syncthreads();
    const u32 nseg = it.seg_end - it.seg_begin;
    const bool may_wrap = it.ninst >= 65536u;
    u32 spare = 0xFFFFFFFFu;  // a dense entry this thread reserved but could not publish (same key claimed concurrently)

    // counters of dense entry idx: meta = left slot | right slot << 3 | (contigs) header bits 15..31
    auto update = [&](u32 idx, u32 meta) {
      const u32 sl = meta & 7u, sr = (meta >> 3) & 7u;
      const u32 va = vals_s + idx * (u32)(VW * 4);
      if (!CTG || !(meta & 0x8000u)) {  // a read instance (count 1: kcount_cpu.cpp:75,348-352)
        if (WIDE) {
          atoms_add(va + 4u * sl, 1u);
          if (sr < 4u) atoms_add(va + 4u * (5u + sr), 1u);
        } else {
          // u16 halves: left slots 0..4 are halves 0..4, right slots 0..4 halves 5..9 (half 9 = right "none": a spare that
          // is never read, so the right add needs no predicate). A half that was 0xFFFF before the add has wrapped (into
          // its neighbour): the item is re-run with wide counters. Items of < 65536 instances cannot wrap: no check.
          const u32 hl = sl, hr = 5u + sr;
          const u32 il = 1u + (hl & 1u) * 0xFFFFu, ir = 1u + (hr & 1u) * 0xFFFFu;
          if (may_wrap) {
            u32 old = atoms_add_ret(va + ((hl << 1) & ~3u), il);
            if ((old | ((hl & 1u) ? 0x0000FFFFu : 0xFFFF0000u)) == 0xFFFFFFFFu) s_wide = 1;
            if (sr < 4u) {
              old = atoms_add_ret(va + ((hr << 1) & ~3u), ir);
              if ((old | ((hr & 1u) ? 0x0000FFFFu : 0xFFFF0000u)) == 0xFFFFFFFFu) s_wide = 1;
            }
          } else {
            atoms_add(va + ((hl << 1) & ~3u), il);
            atoms_add(va + ((hr << 1) & ~3u), ir);
          }
        }
      } else {  // a contig instance: remembered as (occurrences, min depth, set of extension pairs), DESIGN.md "contig pass"
        u32 *v = vals + (size_t)idx * VW;
        atomicAdd(v + VR, 1u);
        atomicMax(v + VR + 1, ~(meta >> 16));
        atomicOr(v + VR + 2, 1u << (sl * 5 + sr));
      }
    };
    // probe and claim: returns the key's dense entry, or 0xFFFFFFFF when the key store is full (the item is re-run)
    auto find_or_claim = [&](const u64 (&key)[KW]) -> u32 {
      const u64 h = table_hash<KW>(key);
      const u32 fp = ((u32)(h >> 32) & ~IDX_MASK) | (1u << IDX_BITS);
      u32 bkt = __umulhi((u32)h, nbuckets);
      u32 idx = 0xFFFFFFFFu, mine = 0xFFFFFFFFu, probes = 0;
      bool claimed = false;
      for (;;) {
        const uint4 bq = lds128_volatile(tab_s + 16u * bkt);
        const u32 mn = min(min(bq.x ^ fp, bq.y ^ fp), min(bq.z ^ fp, bq.w ^ fp));
        if (mn <= IDX_MASK) {  // some entry carries our fingerprint: almost always the key itself (claimed meanwhile)
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const u32 eq = q == 0 ? bq.x : q == 1 ? bq.y : q == 2 ? bq.z : bq.w;
            if ((eq ^ fp) <= IDX_MASK && idx == 0xFFFFFFFFu) {
              const u32 c2 = (eq & IDX_MASK) - 1u;
              bool s2 = true;
#pragma unroll
              for (int i = 0; i < KW; i++) s2 = s2 && (lds64_volatile(keys_s + c2 * (u32)(KW * 8) + 8u * i) == key[i]);
              if (s2) idx = c2;
            }
          }
          if (idx != 0xFFFFFFFFu) break;
        }
        // entries of a bucket are claimed in order and never removed, so the occupied ones are a prefix: the bucket has
        // room iff its last entry is empty, and the first empty entry is the number of occupied ones (a stale view can
        // only point at an entry that is already taken: the CAS fails and the bucket is read again)
        if (bq.w == 0u) {  // the key is not in the table: claim the bucket's first empty entry
          if (mine == 0xFFFFFFFFu) {  // reserve a dense entry (or reuse the one a lost race left us), fill it, publish it
            if (spare != 0xFFFFFFFFu) {
              mine = spare;
              spare = 0xFFFFFFFFu;
            } else {
              mine = atomicAdd(&s_nkeys, 1u);
              if (mine >= a.key_cap) {
                s_over = 1;
                break;
              }
            }
#pragma unroll
            for (int i = 0; i < KW; i++) keys[(size_t)mine * KW + i] = key[i];
            __threadfence_block();
          }
          const u32 q = (bq.x != 0u ? 1u : 0u) + (bq.y != 0u ? 1u : 0u) + (bq.z != 0u ? 1u : 0u);
          if (atomicCAS(&tab[4 * bkt + q], 0u, fp | (mine + 1u)) == 0u) {
            idx = mine;
            claimed = true;
            break;
          }
          // somebody else took that entry: look at this bucket again (it may now hold our key)
        } else {
          probes++;
          bkt = (bkt + 1 == nbuckets) ? 0u : bkt + 1;
        }
      }
      if (mine != 0xFFFFFFFFu && !claimed && mine < a.key_cap) {  // lost a same-key race: the entry stays ours for the next new key
        keys[(size_t)mine * KW + KW - 1] = KEY_EMPTY;
        spare = mine;
      }
      my_max_probe = max(my_max_probe, probes);
      return idx;
    };
    // miss queue of this warp (QUEUE): ring positions [qh, qh + qn) hold instances waiting for find-or-claim
    u32 qh = 0, qn = 0;
    auto drain = [&](u32 n) {  // the n <= 32 oldest waiting instances, one per lane
      __syncwarp();            // the entries were written by other lanes
      if ((u32)gt < n) {
        const u32 pos = (qh + (u32)gt) & (u32)(QCAP - 1);
        u64 key[KW];
#pragma unroll
        for (int i = 0; i < KW; i++) key[i] = lds64_volatile(mqk_s + pos * (u32)(KW * 8) + 8u * i);
        const u32 meta = CTG ? lds32_volatile(mqm_s + pos * QMB) : lds8_volatile(mqm_s + pos * QMB);
        const u32 idx = find_or_claim(key);
        if (idx != 0xFFFFFFFFu) update(idx, meta);
      }
      qh = (qh + n) & (u32)(QCAP - 1);
      qn -= n;
      __syncwarp();            // entries are free for the next pushes only once every lane has read its own
    };

    for (u32 sg0 = 0; sg0 < nseg; sg0 += SEG_CACHE) {
      const int ns = (int)min((u32)SEG_CACHE, nseg - sg0);
      __syncthreads();
      if (s_over) break;
      if (tid < ns) segc[tid] = a.segs[it.seg_begin + sg0 + tid];
      __syncthreads();
      // every warp streams its own contiguous 1/32 of this round's supermers (handing chunks out from a shared cursor was
      // measured and lost 4-8 %: v20)
      const u32 r_lo = segc[0].sup0, r_hi = segc[ns - 1].sup0 + segc[ns - 1].nsup;
      const u32 span = r_hi - r_lo;
      const u32 g_lo = r_lo + (u32)(((u64)span * grp) / CNT_GROUPS), g_hi = r_lo + (u32)(((u64)span * (grp + 1)) / CNT_GROUPS);
      u32 c0 = g_lo;
      // segment of the chunk's first supermer, and where the next segment starts (most lanes of most chunks need no search)
      int seg_hint = 0;
      u32 seg_next = 0xFFFFFFFFu;
      auto seek = [&](u32 c) {
        while (seg_hint + 1 < ns && segc[seg_hint + 1].sup0 <= c) seg_hint++;
        seg_next = seg_hint + 1 < ns ? segc[seg_hint + 1].sup0 : 0xFFFFFFFFu;
      };
      seek(c0);
      u32 consumed = 0;
      // The record a lane stages is requested one chunk ahead (its first 16 bytes: header + first bases; the rest of the
      // record shares its sector): ncu r2_v7 showed 7.5 % of all warp samples waiting for this load at the top of a chunk.
      const uint4 *nsrc = nullptr;
      uint4 nv = make_uint4(0, 0, 0, 0);
      u32 nctg = 0;
      auto fetch = [&](u32 g) {
        nsrc = nullptr;
        if (gt < sup_per_chunk && g < g_hi) {
          int si = seg_hint;
          if (g >= seg_next) {
            si++;
            while (si + 1 < ns && segc[si + 1].sup0 <= g) si++;
          }
          nsrc = (const uint4 *)(segc[si].rec + (size_t)(g - segc[si].sup0) * rw);
          nv = __ldg(nsrc);
          if (CTG) nctg = segc[si].is_ctg << 15;  // nb <= 256: bit 15 of the header is free
        }
      };
      constexpr bool PREFETCH = KW == 1;  // measured (r2_v8): k=21 count 19.3 -> 18.8 ms; k=33 (64 registers) 23.7 -> 24.2 ms
      if (PREFETCH) fetch(c0 + (u32)gt);
      while (c0 < g_hi) {
        // ---- a. one supermer record per thread: copy it to the staging buffer ----
        if (!PREFETCH) fetch(c0 + (u32)gt);
        const uint4 *src = nsrc;
        const bool have = src != nullptr;
        u32 nk = 0;
        if (have) {
          uint4 *dst = (uint4 *)(stage + gt * rw);
          uint4 v = nv;
          nk = (v.x & 0xFFFFu) - (u32)k - 1u;
          if (CTG) v.x |= nctg;
          dst[0] = v;
          for (int x = 1; x < ncopy; x++) dst[x] = __ldg(src + x);
        }
        // ---- b. scan of the k-mer counts inside the pipeline; how many supermers fit this chunk ----
        u32 ik = nk;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const u32 y = __shfl_up_sync(0xffffffffu, ik, d);
          if (lane >= d) ik += y;
        }
        const bool fits0 = have && ik <= (u32)GRP_EXP;  // prefix sums are monotone: the fitting ones are a prefix
        u32 n_take = (u32)__popc(__ballot_sync(0xffffffffu, fits0));
        // The instances are consumed 32 at a time: a chunk whose total is just above a multiple of 32 runs its last iteration
        // almost empty (ncu r2_v6: 28.2 of 32 lanes on the fast path). Unless this is the last chunk of the warp's range, give
        // up to 8 supermers back to the next chunk when that leaves a total that fills its last iteration (<= 5 idle lanes).
        {
          const u32 rem = ik & 31u;
          u32 gm = __ballot_sync(0xffffffffu, fits0 && (rem == 0u || rem >= 27u));
          if (n_take > 9u) gm &= ~((1u << (n_take - 9u)) - 1u);
          if (gm && c0 + n_take < g_hi) n_take = 32u - (u32)__clz(gm);
        }
        const bool fits = have && (u32)gt < n_take;
        const u32 total = __reduce_add_sync(0xffffffffu, fits ? nk : 0u);
        if (PREFETCH) fetch(c0 + n_take + (u32)gt);  // the next chunk's records travel while this chunk's instances are counted
        if (fits) {
          const u32 koff = ik - nk;
          if (KC && WC <= 12) {
#pragma unroll
            for (u32 j = 0; j < (u32)(KC ? WC : 1); j++)
              if (j < nk) sts16(expl_s + 2u * (koff + j), (u32)gt | (j << 8));
          } else {
            for (u32 j = 0; j < nk; j++) expl[koff + j] = (uint16_t)((u32)gt | (j << 8));
          }
        }
        __syncwarp();

        // ---- c. every k-mer instance of the chunk ----
        // The fast path (about 4 of 5 instances at 33x coverage): extract, hash, ONE 128-bit bucket load, key compare, two
        // counter adds. An instance that misses (new key, key still waiting in somebody's queue, shared fingerprint, key that
        // lives in a neighbouring bucket) needs the probe-and-claim path; handled where it occurs that path ran with 4-8 of
        // 32 lanes active in almost every iteration (a quarter of all instructions). QUEUE (keys of <= 2 words): the instance
        // parks key + counter slots in the warp's ring and find-or-claim runs whenever 32 are waiting. Otherwise (keys of 3-4
        // words, whose ring would cost a third of the table): its LIST ENTRY is compacted to the front of the list and a
        // second pass extracts those k-mers again with all lanes. Counter updates commute, so the order is free.
        if (QUEUE) {
          for (u32 t0 = 0; t0 < total; t0 += GRP_THREADS) {
            const u32 t = t0 + (u32)gt;
            bool miss = false;
            u64 key[KW];
            u32 meta = 0u;
            if (t < total) {
              const u32 e = lds16(expl_s + 2u * t);
              const u32 si = e & 255u;
              const int j = (int)(e >> 8) + 1;  // k-mer index inside the supermer, 1-based (kcount_cpu.cpp:319)
              Instance in;
              u32 hdr = 0u;
              if (KC) {
                const u32 rec_s = stage_s + si * (u32)(rw * 4);
                if (CTG) hdr = lds32(rec_s);
                in = extract_fixed<KW, KC ? KC : 21, KC ? WC : 7>(rec_s, j, key);
              } else {
                const u32 *rec = stage + si * rw;
                if (CTG) hdr = rec[0];
                in = extract_any<KW>(rec, j, k, nwb, key);
              }
              meta = in.l | (in.r << 3) | (CTG ? (hdr & 0xFFFF8000u) : 0u);
              const u64 h = table_hash<KW>(key);  // internal slot hash (kc_device.cuh); Kmer::hash is not needed for parity
              if (it.nsub <= 1 || ((u32)(h >> 40) & (it.nsub - 1)) == it.sub) {
                // find: one 128-bit load of the bucket. entry ^ fp is (dense index + 1) <= IDX_MASK exactly when the entry's upper
                // 19 bits equal the fingerprint (fp has bit 13 set: an empty entry never matches), so the minimum over the four
                // words IS the candidate's index + 1 — no selects
                const u32 fp = ((u32)(h >> 32) & ~IDX_MASK) | (1u << IDX_BITS);
                const uint4 bq = lds128_volatile(tab_s + 16u * __umulhi((u32)h, nbuckets));
                const u32 mn = min(min(bq.x ^ fp, bq.y ^ fp), min(bq.z ^ fp, bq.w ^ fp));
                miss = true;
                if (mn <= IDX_MASK) {
                  const u32 ka = keys_s + (mn - 1u) * (u32)(KW * 8);
                  bool same = true;
#pragma unroll
                  for (int i = 0; i < KW; i++) same = same && (lds64_volatile(ka + 8u * i) == key[i]);
                  if (same) {
                    miss = false;
                    update(mn - 1u, meta);
                  }
                }
              }
            }
            const u32 mm = __ballot_sync(0xffffffffu, miss);
            if (mm) {  // park key + slots at the ring's tail; 32 waiting -> one full-warp pass through find-or-claim
              if (miss) {
                const u32 pos = (qh + qn + (u32)__popc(mm & ((1u << gt) - 1u))) & (u32)(QCAP - 1);
#pragma unroll
                for (int i = 0; i < KW; i++) sts64(mqk_s + pos * (u32)(KW * 8) + 8u * i, key[i]);
                if (CTG) sts32(mqm_s + pos * QMB, meta); else sts8(mqm_s + pos * QMB, meta);
              }
              qn += (u32)__popc(mm);
              if (qn >= DRAIN_MIN) drain(min(qn, 32u));
            }
          }
        } else {
          u32 nmiss = 0;
          for (u32 t0 = 0; t0 < total; t0 += GRP_THREADS) {
            const u32 t = t0 + (u32)gt;
            bool miss = false;
            u32 e = 0;
            if (t < total) {
              e = lds16(expl_s + 2u * t);
              const u32 si = e & 255u;
              const int j = (int)(e >> 8) + 1;  // k-mer index inside the supermer, 1-based (kcount_cpu.cpp:319)
              u64 key[KW];
              Instance in;
              u32 hdr = 0u;
              if (KC) {
                const u32 rec_s = stage_s + si * (u32)(rw * 4);
                if (CTG) hdr = lds32(rec_s);
                in = extract_fixed<KW, KC ? KC : 21, KC ? WC : 7>(rec_s, j, key);
              } else {
                const u32 *rec = stage + si * rw;
                if (CTG) hdr = rec[0];
                in = extract_any<KW>(rec, j, k, nwb, key);
              }
              const u64 h = table_hash<KW>(key);  // internal slot hash (kc_device.cuh); Kmer::hash is not needed for parity
              if (it.nsub <= 1 || ((u32)(h >> 40) & (it.nsub - 1)) == it.sub) {
                // find: one 128-bit load of the bucket. entry ^ fp is (dense index + 1) <= IDX_MASK exactly when the entry's upper
                // 19 bits equal the fingerprint (fp has bit 13 set: an empty entry never matches), so the minimum over the four
                // words IS the candidate's index + 1 — no selects
                const u32 fp = ((u32)(h >> 32) & ~IDX_MASK) | (1u << IDX_BITS);
                const uint4 bq = lds128_volatile(tab_s + 16u * __umulhi((u32)h, nbuckets));
                const u32 mn = min(min(bq.x ^ fp, bq.y ^ fp), min(bq.z ^ fp, bq.w ^ fp));
                miss = true;
                if (mn <= IDX_MASK) {
                  const u32 ka = keys_s + (mn - 1u) * (u32)(KW * 8);
                  bool same = true;
#pragma unroll
                  for (int i = 0; i < KW; i++) same = same && (lds64_volatile(ka + 8u * i) == key[i]);
                  if (same) {
                    miss = false;
                    update(mn - 1u, in.l | (in.r << 3) | (CTG ? (hdr & 0xFFFF8000u) : 0u));
                  }
                }
              }
            }
            const u32 mm = __ballot_sync(0xffffffffu, miss);
            if (mm) {  // entries [0, nmiss) are behind the read position: nmiss <= t0
              if (miss) sts16(expl_s + 2u * (nmiss + (u32)__popc(mm & ((1u << gt) - 1u))), e);
              nmiss += (u32)__popc(mm);
            }
          }
          __syncwarp();
          // pass 2: the misses, all lanes busy
          for (u32 t0 = 0; t0 < nmiss; t0 += GRP_THREADS) {
            const u32 t = t0 + (u32)gt;
            if (t < nmiss) {
              const u32 e = lds16(expl_s + 2u * t);
              const u32 si = e & 255u;
              const int j = (int)(e >> 8) + 1;
              u64 key[KW];
              Instance in;
              u32 hdr = 0u;
              if (KC) {
                const u32 rec_s = stage_s + si * (u32)(rw * 4);
                if (CTG) hdr = lds32(rec_s);
                in = extract_fixed<KW, KC ? KC : 21, KC ? WC : 7>(rec_s, j, key);
              } else {
                const u32 *rec = stage + si * rw;
                if (CTG) hdr = rec[0];
                in = extract_any<KW>(rec, j, k, nwb, key);
              }
              const u32 idx = find_or_claim(key);
              if (idx != 0xFFFFFFFFu) update(idx, in.l | (in.r << 3) | (CTG ? (hdr & 0xFFFF8000u) : 0u));
            }
          }
        }
        const bool stop = __any_sync(0xffffffffu, *(volatile u32 *)&s_over != 0);  // also orders this chunk's reads of the
                                                                                   // staging buffer before the next chunk's writes
        c0 += n_take;
        consumed += n_take;
        if (stop) break;
        if (c0 >= seg_next) seek(c0);
      }
      if (gt == 0 && consumed) atomicAdd(&s_consumed, consumed);
    }
    if (QUEUE) {  // what is still waiting (fewer than 32 per warp)
      while (qn && !__any_sync(0xffffffffu, *(volatile u32 *)&s_over != 0u)) drain(min(qn, 32u));
    }
    __syncthreads();
    const bool over = s_over != 0;
    const bool wide_needed = !WIDE && s_wide;
    const u32 n = min(s_nkeys, a.key_cap);
    if (a.measure_only && !over) {  // probe run: how many of the n slots hold a key (the others were lost to same-key races)
      u32 live = 0;
      for (u32 i = tid; i < n; i += CNT_THREADS) live += keys[(size_t)i * KW + KW - 1] != KEY_EMPTY ? 1u : 0u;
      live = __reduce_add_sync(0xffffffffu, live);
      if (lane == 0 && live) atomicAdd(&s_live, live);
      __syncthreads();
    }
    if (tid == 0) {
      ItemStatus st;
      st.overflow = over ? 1u : 0u;
      st.progress = (a.measure_only && !over) ? s_live : s_consumed;  // probe: distinct k-mers; else: supermers consumed
      st.nkeys = n;
      st.wide = wide_needed ? 1u : 0u;
      a.status[item_idx] = st;
      s_prev = n;
      if (have_nx) {
        ((uint4 *)&s_work)[0] = nx0;
        ((uint4 *)&s_work)[1] = nx1;
        s_widx = nx_idx;
      }
      if ((over || wide_needed) && !a.measure_only) {
        const u64 slot = atomicAdd((unsigned long long *)&a.counters[CTR_RETRY], 1ull) - a.retry_base;
        if (a.retry) {
          a.retry[slot].item = a.items[item_idx];  // re-read: keeps the descriptor out of the registers of the hot loops
          a.retry[slot].st = st;
        }
      }
    }
    if (over || wide_needed || a.measure_only) continue;

    // ---- e. depth filter, extension choice, compaction (kcount_cpu.cpp:490-528) ----
    u32 n_occ = 0, n_sv = 0, n_pf = 0;
    for (u32 b0 = 0; b0 < n; b0 += CNT_THREADS) {
      const u32 idx = b0 + (u32)tid;
      const u64 *kp = keys + (size_t)idx * KW;
      const bool occ = idx < n && kp[KW - 1] != KEY_EMPTY;
      bool survive = false;
      u32 count = 0, rcount = 0;
      u32 lc = 5u, rc_ = 5u;  // extension codes: A C G T F X
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
              const int dm0 = dmin_dyn_of(count, a.dmin_thres, a.frac);
              use_ctg = ext_code(L, dm0) >= 4u || ext_code(R, dm0) >= 4u;  // non-UU read k-mer: an X or F side (kcount_cpu.cpp:373-376)
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
          const int dm = dmin_dyn_of(count, a.dmin_thres, a.frac);
          lc = ext_code(L, dm);
          rc_ = ext_code(R, dm);
          survive = !(lc == 5u && rc_ == 5u);
        }
      }
      const bool pf = a.pf_keys && occ && rcount > 0;  // k-mers seen in reads (the table as it stands after the reads pass)
      // verdict of this key, parked in the index table (no longer needed: insertion is over) until output space is reserved
      if (idx < n) {
        tab[idx] = (count & 0xFFFFu) | (lc << 16) | (rc_ << 19) | (survive ? 1u << 22 : 0u) | (pf ? 1u << 23 : 0u);
      }
      n_occ += occ ? 1u : 0u;
      n_sv += survive ? 1u : 0u;
      n_pf += pf ? 1u : 0u;
    }
    // one reservation of output space per item (a global atomic with result costs ~1 us: not once per 1024 keys)
#pragma unroll
    for (int d = 16; d; d >>= 1) {
      n_occ += __shfl_xor_sync(0xffffffffu, n_occ, d);
      n_sv += __shfl_xor_sync(0xffffffffu, n_sv, d);
      n_pf += __shfl_xor_sync(0xffffffffu, n_pf, d);
    }
    if (lane == 0) { warp_c[wid] = n_sv; warp_p[wid] = n_pf; if (n_occ) atomicAdd(&s_live, n_occ); }
    __syncthreads();
    if (wid == 0) {
      u32 c = lane < CNT_GROUPS ? warp_c[lane] : 0u, p = lane < CNT_GROUPS ? warp_p[lane] : 0u;
#pragma unroll
      for (int d = 16; d; d >>= 1) {
        c += __shfl_xor_sync(0xffffffffu, c, d);
        p += __shfl_xor_sync(0xffffffffu, p, d);
      }
      if (lane == 0) {
        s_obase[0] = c ? atomicAdd(&a.counters[CTR_SURV], (u64)c) : 0ull;
        s_obase[1] = p ? atomicAdd(&a.counters[CTR_PF], (u64)p) : 0ull;
        s_ocur[0] = 0;
        s_ocur[1] = 0;
      }
    }
    __syncthreads();
    for (u32 b0 = 0; b0 < n; b0 += CNT_THREADS) {
      const u32 idx = b0 + (u32)tid;
      const u32 res = idx < n ? tab[idx] : 0u;
      const bool survive = (res >> 22) & 1u, pf = (res >> 23) & 1u;
      const u64 *kp = keys + (size_t)idx * KW;
      const u32 sv_mask = __ballot_sync(0xffffffffu, survive);
      if (sv_mask) {
        u32 wbase = 0;
        if (lane == 0) wbase = atomicAdd(&s_ocur[0], (u32)__popc(sv_mask));
        wbase = __shfl_sync(0xffffffffu, wbase, 0);
        if (survive) {
          const u64 o = s_obase[0] + wbase + __popc(sv_mask & ((1u << lane) - 1u));
          if (o < a.out_cap) {
#pragma unroll
            for (int i = 0; i < KW; i++) a.out_keys[o * KW + i] = kp[i];
            a.out_counts[o] = (uint16_t)(res & 0xFFFFu);
            a.out_left[o] = ext_char((res >> 16) & 7u);
            a.out_right[o] = ext_char((res >> 19) & 7u);
          } else {
            a.counters[CTR_ERR] = 1;
          }
        }
      }
      const u32 pf_mask = __ballot_sync(0xffffffffu, pf);
      if (pf_mask) {
        u32 wbase = 0;
        if (lane == 0) wbase = atomicAdd(&s_ocur[1], (u32)__popc(pf_mask));
        wbase = __shfl_sync(0xffffffffu, wbase, 0);
        if (pf) {
          const u64 o = s_obase[1] + wbase + __popc(pf_mask & ((1u << lane) - 1u));
          if (o < a.pf_cap) {
            const u32 *v = vals + (size_t)idx * VW;
#pragma unroll
            for (int i = 0; i < KW; i++) a.pf_keys[o * KW + i] = kp[i];
            uint16_t *ov = a.pf_vals + o * 9;
            if (WIDE) {
              ov[0] = (uint16_t)sat16((u64)v[0] + v[1] + v[2] + v[3] + v[4]);
#pragma unroll
              for (int i = 0; i < 4; i++) { ov[1 + i] = (uint16_t)sat16(v[i]); ov[5 + i] = (uint16_t)sat16(v[5 + i]); }
            } else {
              const u32 w0 = v[0], w1 = v[1], w2 = v[2], w3 = v[3], w4 = v[4];
              ov[0] = (uint16_t)sat16((u64)(w0 & 0xFFFFu) + (w0 >> 16) + (w1 & 0xFFFFu) + (w1 >> 16) + (w2 & 0xFFFFu));
              ov[1] = (uint16_t)w0; ov[2] = (uint16_t)(w0 >> 16); ov[3] = (uint16_t)w1; ov[4] = (uint16_t)(w1 >> 16);
              ov[5] = (uint16_t)(w2 >> 16); ov[6] = (uint16_t)w3; ov[7] = (uint16_t)(w3 >> 16); ov[8] = (uint16_t)w4;
            }
          } else {
            a.counters[CTR_ERR] = 1;
          }
        }
      }
    }
    __syncthreads();
    if (tid == 0 && s_live) atomicAdd(&a.counters[CTR_DISTINCT], (u64)s_live);
  }
  my_max_probe = __reduce_max_sync(0xffffffffu, my_max_probe);
  if (lane == 0 && my_max_probe) atomicMax(&a.counters[CTR_MAXPROBE], (u64)my_max_probe);
}

// Work items of the count stage, built where the bucket directories already are. One thread per partition of this shard:
// its runs of supermers in every parsed / received batch become the segment list of one item (a fixed slot of nblk
// descriptors per partition, so no compaction and no host involvement); partitions far above the mean are also appended to
// the heavy list. Totals (instances from reads / contigs, supermers) are reduced into counters[].
struct BuildArgs {
  const BlockDesc *blocks;
  u32 nblk, P;   // P = buckets of this shard (the granularity the parse stage bucketed at)
  u32 nitems;    // item q covers buckets [q * step, min(q * step + g, P)): step == g merges neighbours (their records are
  u32 g, step;   //   contiguous inside every batch); step > g samples the partition space (probe run)
  int rw;
  u64 heavy_thr;  // nsup above this -> heavy
  SegDesc *segs;  // [P * nblk]
  WorkItem *items;  // [P]
  u64 *heavy;     // [P]
  u64 *counters;
};

__global__ void __launch_bounds__(256) build_items_kernel(const BuildArgs a) {
  const u32 p = blockIdx.x * blockDim.x + threadIdx.x;
  u64 nsup = 0, ri = 0, ci = 0;
  if (p < a.nitems) {
    const u32 s0 = p * a.nblk;
    const u32 b_lo = min(p * a.step, a.P), b_hi = min(b_lo + a.g, a.P);
    u32 cnt = 0;
    for (u32 j = 0; j < a.nblk; j++) {
      const BlockDesc d = a.blocks[j];
      const u64 r0 = d.base[d.first + b_lo], r1 = d.base[d.first + b_hi];  // base[] has one entry past the last bucket
      if (r1 == r0) continue;
      u64 ninst = 0;
      for (u32 b = b_lo; b < b_hi; b++) ninst += d.hist[d.first + b].ninst;
      SegDesc sd;
      sd.rec = d.rec + (size_t)r0 * a.rw;
      sd.nsup = (u32)(r1 - r0);
      sd.is_ctg = d.is_ctg;
      sd.sup0 = (u32)nsup;
      sd.pad = 0;
      a.segs[s0 + cnt++] = sd;
      nsup += r1 - r0;
      if (d.is_ctg) ci += ninst; else ri += ninst;
    }
    WorkItem it;
    it.seg_begin = s0;
    it.seg_end = s0 + cnt;
    it.nsup = (u32)nsup;
    it.part = p;
    it.sub = 0;
    it.nsub = 1;
    it.heavy = nsup > a.heavy_thr ? 1u : 0u;
    it.ninst = (u32)min(ri + ci, (u64)0xFFFFFFFFull);
    a.items[p] = it;
    if (nsup > 0xFFFFFFF0ull) a.counters[CTR_ERR] = 2;  // a partition this large cannot be indexed: more partitions needed
    if (it.heavy) {
      const u64 slot = atomicAdd((unsigned long long *)&a.counters[CTR_NHEAVY], 1ull);
      a.heavy[slot] = (nsup << 32) | p;
    }
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) {
    nsup += __shfl_xor_sync(0xffffffffu, nsup, d);
    ri += __shfl_xor_sync(0xffffffffu, ri, d);
    ci += __shfl_xor_sync(0xffffffffu, ci, d);
  }
  if ((threadIdx.x & 31) == 0) {
    if (nsup) atomicAdd((unsigned long long *)&a.counters[CTR_SUP], (unsigned long long)nsup);
    if (ri) atomicAdd((unsigned long long *)&a.counters[CTR_READ_INST], (unsigned long long)ri);
    if (ci) atomicAdd((unsigned long long *)&a.counters[CTR_CTG_INST], (unsigned long long)ci);
  }
}

}  // namespace kc

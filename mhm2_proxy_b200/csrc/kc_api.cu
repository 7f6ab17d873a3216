// Host side of the B200 kcount library and its C ABI (include/kcount_b200.h).
// One handle = one shard (one GPU). Pipeline per handle:
//   add_reads/add_contigs : [H2D copy] -> parse_kernel -> exact bucket layout -> fill_kernel   (kc_parse.cuh)
//   finish                : [NCCL all-to-all of the per-shard bucket ranges] -> per group of partitions:
//                           clear tables -> insert -> filter/compact                               (kc_count.cuh)
// There is no CPU fallback: every data-path step is a CUDA kernel; host code only sizes, allocates and launches.
#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include <dlfcn.h>

#include "../../include/kcount_b200.h"
#include "kc_count.cuh"
#include "kc_parse.cuh"

using namespace kc;

namespace {

constexpr int SEG_MAX = PARSE_TILE;  // longest sequence a parse thread handles in one piece; longer ones are tiled with k+1 overlap

struct DevBuf {
  void *p = nullptr;
  size_t bytes = 0;
};

// one parsed batch (or one received batch): supermers of every bucket laid out bucket-major
struct Block {
  u32 *rec = nullptr;            // supermer records, rw words each, bucket-major
  BucketHist *d_hist = nullptr;  // device: supermers and k-mer instances per bucket (also the directory sent to peers)
  u64 *d_base = nullptr;         // device: index of each bucket's first record (exclusive scan of d_hist[].nsup)
  // d_hist / d_base cover the PRODUCER's buckets (num_shards * P) for local blocks, P entries for received ones
  std::vector<u64> dest_base;    // host, local blocks: first record of each destination shard's range (+ total at the end)
  u64 tot_sup = 0;
  bool is_ctg = false;
  bool received = false;  // true: indexed by partition only
};

struct Nccl;  // opaque, kc_comm.inl

// a receive posted by an exchange round: becomes a received Block once its directory is back on the host (kc_comm.inl)
struct Pending {
  BucketHist *dir = nullptr;  // P entries, device
  u32 *rec = nullptr;
  u64 nsup = 0;
  bool is_ctg = false;
};

// The parse stage runs as a two-slot pipeline: batch i is parsed (ALU-bound) on one slot's stream while the records of
// batch i-1 are still being scattered (memory-bound) by the fill kernel on the other slot's stream. Each slot owns the
// buffers that live from the parse kernel to the end of the fill kernel.
struct ParseSlot {
  cudaStream_t st = nullptr;
  DevBuf d_cursor, d_tot, d_scan, d_elist, d_ecnt;
  u64 *h_tot = nullptr;  // pinned: first record of every destination shard's range (+ total)
  size_t h_tot_bytes = 0;
  cudaEvent_t e0 = nullptr, ec = nullptr, ed = nullptr, e1 = nullptr;  // parse start / parse end / fill start / fill end
  bool busy = false;   // a fill kernel was launched and its block has not been collected yet
  Block blk;
  size_t rec_bytes = 0;
  std::vector<DevBuf> temps;  // inputs that must outlive the fill kernel (segment arrays)
};

}  // namespace

struct kc_handle {
  kc_config cfg;
  int KW = 1, m = 15, w = 1;
  int smem_optin = 0, sms = 0;
  RecGeom geom;  // supermer record geometry for this k
  u32 P = 0;  // buckets per shard the parse stage uses (0 until the first add decides)
  bool auto_merge = false;  // P was chosen by the library: fine-grained, neighbours are merged into work items at finish
  bool finished = false, has_ctg = false;
  bool phase_open = true;   // multi-shard: this shard has added data (or never flushed) since its last drain of exchange rounds
  int dbg_key_cap = 0;      // kc_debug_set_key_cap
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> comm_ev;  // one pair around every exchange group on s_comm (ms_comm)
  cudaStream_t s_main = nullptr, s_copy = nullptr, s_comm = nullptr, s_meta = nullptr;
  std::vector<Pending> pending;
  DevBuf d_meta, d_meta_all, d_tot, d_scan, d_scan_comm;
  u64 *h_pin = nullptr;  // pinned scratch (PIN_BYTES): [0, PIN_META) bucket totals / counters, [PIN_META, ..) exchange metadata
  size_t pin_bytes = 0;
  cudaEvent_t ev_fill = nullptr, ev_x0 = nullptr, ev_x1 = nullptr;
  std::vector<cudaEvent_t> ev_chunk;  // kc_finish_to_host: one per slice of the count stage
  cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr, ev_d = nullptr, ev_copy[2] = {nullptr, nullptr}, ev_free[2] = {nullptr, nullptr};
  std::vector<Block> blocks;       // produced locally
  std::vector<Block> rblocks;      // received from peers (multi-shard)
  DevBuf stage_bytes[2], stage_offs[2];
  DevBuf d_misc;  // u64 instances, u32 err
  ParseSlot slot[2];
  u64 n_batches = 0;        // batches parsed so far: batch i uses slot i & 1
  cudaEvent_t ev_blk = nullptr;  // "records of the block being handed to the exchange are written" (a slot's e1)
  cudaEvent_t ev_first = nullptr;  // start of the first batch since the last drain (ms_parse = wall time of the stage)
  bool stage_open = false;
  // outputs
  DevBuf out_keys, out_counts, out_left, out_right, pf_keys, pf_vals;
  u64 n_surv = 0, n_pf = 0;
  kc_stats st;
  std::string err;
  Nccl *comm = nullptr;
};

namespace {

int exchange_round(kc_handle *h, const Block *blk, bool done, bool *all_done);  // kc_comm.inl

int fail(kc_handle *h, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (h) h->err = buf;
  return code;
}

#define CK(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess)                                                                               \
      return fail(h, e_ == cudaErrorMemoryAllocation ? KC_ERR_NOMEM : KC_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, \
                  cudaGetErrorString(e_));                                                               \
  } while (0)

#define RC(call)            \
  do {                      \
    int r_ = (call);        \
    if (r_) return r_;      \
  } while (0)

// Process-wide caching device allocator. A job runs one kcount round per k (main.cpp:169) and every round needs the same
// multi-GB buffers (supermer buckets, survivor arrays); cudaMalloc/cudaFree of those costs tens of milliseconds per round,
// so freed blocks are kept per device and handed out again. Rule kept by every caller in this file: a block is freed only
// after the stream work that used it has been synchronised. kc_release_cached_memory() gives the memory back.
struct DevCache {
  std::mutex mu;
  std::map<int, std::multimap<size_t, void *>> free_blocks;  // device -> size -> block
  std::map<void *, std::pair<int, size_t>> live;             // block -> (device, size)
};
DevCache &dev_cache() {
  static DevCache c;
  return c;
}

cudaError_t dev_alloc(void **p, size_t bytes) {
  *p = nullptr;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  bytes = (bytes + 511) & ~(size_t)511;
  DevCache &c = dev_cache();
  {
    std::lock_guard<std::mutex> lk(c.mu);
    auto &fb = c.free_blocks[dev];
    auto it = fb.lower_bound(bytes);
    if (it != fb.end() && it->first <= bytes + bytes / 2 + (1u << 20)) {
      *p = it->second;
      c.live[*p] = {dev, it->first};
      fb.erase(it);
      return cudaSuccess;
    }
  }
  e = cudaMalloc(p, bytes);
  if (e == cudaErrorMemoryAllocation) {  // give the cached blocks of this device back and try again
    cudaGetLastError();
    std::vector<void *> drop;
    {
      std::lock_guard<std::mutex> lk(c.mu);
      for (auto &kv : c.free_blocks[dev]) drop.push_back(kv.second);
      c.free_blocks[dev].clear();
    }
    for (void *q : drop) cudaFree(q);
    e = cudaMalloc(p, bytes);
  }
  if (e == cudaSuccess) {
    std::lock_guard<std::mutex> lk(c.mu);
    c.live[*p] = {dev, bytes};
  }
  return e;
}

void dev_free(void *p) {
  if (!p) return;
  DevCache &c = dev_cache();
  std::lock_guard<std::mutex> lk(c.mu);
  auto it = c.live.find(p);
  if (it == c.live.end()) {
    cudaFree(p);
    return;
  }
  c.free_blocks[it->second.first].emplace(it->second.second, p);
  c.live.erase(it);
}

int ensure(kc_handle *h, DevBuf &b, size_t bytes) {
  if (b.bytes >= bytes && b.p) return KC_OK;
  if (b.p) dev_free(b.p);
  b.p = nullptr;
  b.bytes = 0;
  size_t want = bytes + bytes / 8 + 256;
  CK(dev_alloc(&b.p, want));
  b.bytes = want;
  return KC_OK;
}

void release(DevBuf &b) {
  if (b.p) dev_free(b.p);
  b.p = nullptr;
  b.bytes = 0;
}

u32 next_pow2(u64 v) {
  u64 p = 1;
  while (p < v) p <<= 1;
  return (u32)std::min<u64>(p, 1ull << 31);
}

// host-side wall clock for KC_TRACE=1 diagnostics (stderr)
struct HostTimer {
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  bool on = getenv("KC_TRACE") != nullptr;
  void lap(const char *what) {
    if (!on) return;
    auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[kc_trace] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

float elapsed(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  return ms;
}

}  // namespace

extern "C" void kc_config_default(kc_config *cfg, int32_t kmer_len) {
  memset(cfg, 0, sizeof(*cfg));
  cfg->kmer_len = kmer_len;
  cfg->qual_cutoff = 20;     // KCOUNT_QUAL_CUTOFF, CMakeDefinitions.txt:46
  cfg->dmin_thres = 2;       // options.hpp:96
  cfg->dyn_min_depth = 0.9;  // DYN_MIN_DEPTH, CMakeDefinitions.txt:60
  cfg->num_shards = 1;
  cfg->shard_id = 0;
  cfg->device = 0;
}

extern "C" void kc_struct_sizes(uint32_t *config_bytes, uint32_t *stats_bytes) {
  if (config_bytes) *config_bytes = (uint32_t)sizeof(kc_config);
  if (stats_bytes) *stats_bytes = (uint32_t)sizeof(kc_stats);
}

extern "C" const char *kc_last_error(const kc_handle *h) { return h ? h->err.c_str() : "null handle"; }

namespace {

// pinned host scratch, pooled process-wide (cudaMallocHost / cudaFreeHost cost far more than a kcount round can afford)
constexpr size_t PIN_META = 4096;
struct PinPool {
  std::mutex mu;
  std::multimap<size_t, void *> free_blocks;
};
PinPool &pin_pool() {
  static PinPool p;
  return p;
}
cudaError_t pin_get(void **p, size_t &bytes) {
  bytes = std::max<size_t>(bytes, 64 << 10);
  PinPool &pp = pin_pool();
  {
    std::lock_guard<std::mutex> lk(pp.mu);
    auto it = pp.free_blocks.lower_bound(bytes);
    if (it != pp.free_blocks.end()) {
      *p = it->second;
      bytes = it->first;
      pp.free_blocks.erase(it);
      return cudaSuccess;
    }
  }
  return cudaMallocHost(p, bytes);
}
void pin_put(void *p, size_t bytes) {
  if (!p) return;
  PinPool &pp = pin_pool();
  std::lock_guard<std::mutex> lk(pp.mu);
  pp.free_blocks.emplace(bytes, p);
}

}  // namespace

extern "C" int kc_create(const kc_config *cfg, kc_handle **out) {
  if (!cfg || !out) return KC_ERR_INVALID;
  *out = nullptr;
  kc_handle *h = new kc_handle();
  h->cfg = *cfg;
  memset(&h->st, 0, sizeof(h->st));
  *out = h;  // returned even on failure so that kc_last_error can be read; caller destroys it
  const int k = cfg->kmer_len;
  if (k < 2 || k > 127) return fail(h, KC_ERR_INVALID, "kmer_len %d outside 2..127", k);
  if (cfg->dmin_thres < 1) return fail(h, KC_ERR_INVALID, "dmin_thres must be >= 1");
  if (cfg->num_shards < 1 || cfg->shard_id < 0 || cfg->shard_id >= cfg->num_shards)
    return fail(h, KC_ERR_INVALID, "bad shard %d of %d", cfg->shard_id, cfg->num_shards);
  if (cfg->num_partitions < 0 || cfg->num_partitions > (1 << 20)) return fail(h, KC_ERR_INVALID, "num_partitions outside 0..2^20");
  h->KW = k / 32 + 1;  // main.cpp:170, kmer.hpp:64
  int m = k * 2 / 3 + 1;  // kmer_dht.cpp:113-116
  if (m < 15) m = 15;
  if (m > 27) m = 27;
  if (m > k) m = k;  // reference asserts m <= k (kmer.cpp:346); only reachable for k < 15
  h->m = m;
  h->w = k - m + 1;
  h->geom = rec_geom(k, h->w);
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (cfg->device < 0 || cfg->device >= ndev) return fail(h, KC_ERR_INVALID, "device %d of %d", cfg->device, ndev);
  CK(cudaSetDevice(cfg->device));
  CK(cudaDeviceGetAttribute(&h->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, cfg->device));
  CK(cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, cfg->device));
  CK(cudaStreamCreateWithFlags(&h->s_main, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&h->s_copy, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&h->s_comm, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&h->s_meta, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&h->ev_fill, cudaEventDisableTiming));
  CK(cudaEventCreate(&h->ev_x0));
  CK(cudaEventCreate(&h->ev_x1));
  CK(cudaEventCreate(&h->ev_a));
  CK(cudaEventCreate(&h->ev_b));
  CK(cudaEventCreate(&h->ev_c));
  CK(cudaEventCreate(&h->ev_d));
  for (int i = 0; i < 2; i++) {
    CK(cudaEventCreateWithFlags(&h->ev_copy[i], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->ev_free[i], cudaEventDisableTiming));
  }
  CK(cudaEventCreate(&h->ev_first));
  for (ParseSlot &sl : h->slot) {
    CK(cudaStreamCreateWithFlags(&sl.st, cudaStreamNonBlocking));
    CK(cudaEventCreate(&sl.e0));
    CK(cudaEventCreate(&sl.ec));
    CK(cudaEventCreate(&sl.ed));
    CK(cudaEventCreate(&sl.e1));
    sl.h_tot_bytes = ((size_t)cfg->num_shards + 1) * 8;
    CK(pin_get((void **)&sl.h_tot, sl.h_tot_bytes));
  }
  RC(ensure(h, h->d_misc, 64));
  CK(cudaMemsetAsync(h->d_misc.p, 0, 64, h->s_main));
  CK(cudaStreamSynchronize(h->s_main));
  h->pin_bytes = PIN_META + ((size_t)cfg->num_shards + 3) * (size_t)cfg->num_shards * 8;
  if ((size_t)(cfg->num_shards + 1) * 8 > PIN_META) return fail(h, KC_ERR_INVALID, "num_shards %d too large", cfg->num_shards);
  CK(pin_get((void **)&h->h_pin, h->pin_bytes));
  return KC_OK;
}

namespace {

// shared-memory table geometry of the count stage for this handle (kc_count.cuh): how many distinct k-mers one work item holds
void table_geometry(const kc_handle *h, bool wide, bool ctg, u32 &tab_size, u32 &key_cap) {
  const int KW = h->KW, VW = (wide ? VR_WIDE : VR_NARROW) + (ctg ? V_CTG : 0);
  // the kernel's static arrays + alignment slack; CNT_CTAS_PER_SM CTAs (each with 1 KB reserved by the system) share the SM's 228 KB
  const size_t per_cta = CNT_CTAS_PER_SM == 1 ? (size_t)h->smem_optin : ((size_t)h->smem_optin + 1024) / CNT_CTAS_PER_SM - 1024;
  const size_t budget = per_cta - COUNT_STATIC_SMEM;
  const size_t fixed = count_smem_bytes(KW, VW, h->geom.rw, 0, 0) + 32;
  const size_t per_key = (size_t)KW * 8 + (size_t)VW * 4 + 8;  // + two 4-byte index-table entries (load <= 0.5)
  u64 cap = budget > fixed ? (budget - fixed) / per_key : 16;
  cap = std::min<u64>(cap, IDX_MASK - 1);
  (void)ctg;
  if (h->dbg_key_cap > 0) cap = std::min<u64>(cap, (u64)h->dbg_key_cap);
  cap = std::max<u64>(cap, 16);
  key_cap = (u32)cap;
  tab_size = (u32)((2 * cap + 15) & ~15ull);  // whole 16-byte buckets; keeps what follows 16-byte aligned
}

void decide_partitions(kc_handle *h, u64 inst_estimate) {
  if (h->P) return;
  if (h->cfg.num_partitions) {
    h->P = (u32)h->cfg.num_partitions;
  } else {
    u64 est = h->cfg.expected_instances ? h->cfg.expected_instances : inst_estimate;
    // One work item of the count stage should hold about 0.8 x key_cap distinct k-mers so that its whole table sits in one
    // CTA's shared memory. distinct / instances is not known before counting (0.13 at 33x coverage, > 0.5 at 2x), so the
    // parse stage buckets at the granularity that would suit a ratio of 1, and finish_count measures the ratio on a
    // sample of buckets and merges neighbouring buckets into items of the right size. A bucket that still turns out too
    // large is re-run in hash sub-passes, never dropped.
    u32 tab, cap;
    table_geometry(h, false, false, tab, cap);
    u64 want = (u64)((double)est / (0.8 * cap)) + 1;
    h->P = (u32)std::min<u64>(std::max<u64>(want, 1), 1u << 20);
    h->auto_merge = true;
  }
  while ((u64)h->P * (u64)h->cfg.num_shards > (1ull << 22)) h->P >>= 1;
  // NOTE multi-shard: every shard must arrive at the same P (checked at exchange time); give all shards the same
  // cfg.num_partitions or cfg.expected_instances.
  h->st.num_partitions = h->P;
}

// Exclusive prefix sum of the per-bucket supermer counts -> index of each bucket's first record, base[n] = total.
// n is 10^4..4*10^6 (num_shards * P): three small launches — per-tile sums, scan of the (<= 1024) tile sums, per-tile scan.
// totals (optional): the first record of every destination shard's range, totals[n / P] = all of them.
constexpr u32 SCAN_TILE = 4096;  // 1024 threads x 4 entries

__device__ __forceinline__ u64 block_exclusive_scan_1024(u64 v, u64 *warp_tot /* [32] shared */, u64 &block_total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  u64 inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const u64 y = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += y;
  }
  if (lane == 31) warp_tot[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    const u64 w = warp_tot[lane];
    u64 wi = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const u64 y = __shfl_up_sync(0xffffffffu, wi, d);
      if (lane >= d) wi += y;
    }
    warp_tot[lane] = wi - w;  // exclusive
    if (lane == 31) warp_tot[32] = wi;
  }
  __syncthreads();
  block_total = warp_tot[32];
  return warp_tot[wid] + inc - v;
}

__global__ void __launch_bounds__(1024) scan_tile_sums_kernel(const BucketHist *hist, u32 n, u64 *tile_sum) {
  __shared__ u64 wt[33];
  const u32 i0 = blockIdx.x * SCAN_TILE + threadIdx.x * 4;
  u64 s = 0;
#pragma unroll
  for (u32 j = 0; j < 4; j++)
    if (i0 + j < n) s += hist[i0 + j].nsup;
  u64 tot;
  block_exclusive_scan_1024(s, wt, tot);
  if (threadIdx.x == 0) tile_sum[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) scan_tile_offsets_kernel(u64 *tile_sum, u32 ntiles) {
  __shared__ u64 wt[33];
  const u64 v = threadIdx.x < ntiles ? tile_sum[threadIdx.x] : 0;
  u64 tot;
  const u64 ex = block_exclusive_scan_1024(v, wt, tot);
  if (threadIdx.x < ntiles) tile_sum[threadIdx.x] = ex;
  if (threadIdx.x == 0) tile_sum[ntiles] = tot;
}

__global__ void __launch_bounds__(1024) scan_apply_kernel(const BucketHist *hist, u32 n, const u64 *tile_off, u64 *base, u32 P,
                                                          u64 *totals, u32 *cursor) {
  __shared__ u64 wt[33];
  const u32 i0 = blockIdx.x * SCAN_TILE + threadIdx.x * 4;
  u32 c[4];
  u64 s = 0;
#pragma unroll
  for (u32 j = 0; j < 4; j++) {
    c[j] = i0 + j < n ? hist[i0 + j].nsup : 0u;
    s += c[j];
  }
  u64 tot;
  u64 run = tile_off[blockIdx.x] + block_exclusive_scan_1024(s, wt, tot);
#pragma unroll
  for (u32 j = 0; j < 4; j++) {
    const u32 i = i0 + j;
    if (i < n) {
      base[i] = run;
      if (cursor) cursor[i] = (u32)run;  // the fill kernel's bucket cursors start at the bucket's first record
      if (totals && i % P == 0) totals[i / P] = run;
    }
    run += c[j];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const u64 all = tile_off[gridDim.x];
    base[n] = all;
    if (totals) totals[n / P] = all;
  }
}

// base must hold n + 1 entries, tile_scratch (n + SCAN_TILE - 1) / SCAN_TILE + 1
int bucket_scan(kc_handle *h, const BucketHist *hist, u32 n, u64 *base, u32 P, u64 *totals, u64 *tile_scratch, cudaStream_t st,
                u32 *cursor = nullptr) {
  const u32 ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  if (ntiles > 1024) return KC_ERR_INVALID;
  scan_tile_sums_kernel<<<ntiles, 1024, 0, st>>>(hist, n, tile_scratch);
  scan_tile_offsets_kernel<<<1, 1024, 0, st>>>(tile_scratch, ntiles);
  scan_apply_kernel<<<ntiles, 1024, 0, st>>>(hist, n, tile_scratch, base, P, totals, cursor);
  h->st.gpu_launches += 3;
  return cudaGetLastError() == cudaSuccess ? KC_OK : KC_ERR_CUDA;
}

// What a slot's last batch left behind: wait for its fill kernel, book its times, keep its block for the count stage.
int collect_slot(kc_handle *h, ParseSlot &sl) {
  if (!sl.busy) return KC_OK;
  CK(cudaStreamSynchronize(sl.st));  // also surfaces kernel faults here
  h->st.ms_parse_count += elapsed(sl.e0, sl.ec);
  h->st.ms_parse_fill += elapsed(sl.ed, sl.e1);
  h->st.supermers_out += sl.blk.tot_sup;
  h->st.supermer_bytes_out += sl.rec_bytes;
  h->blocks.push_back(std::move(sl.blk));
  sl.blk = Block();
  for (DevBuf &b : sl.temps) release(b);
  sl.temps.clear();
  sl.busy = false;
  return KC_OK;
}

// End of a run of batches: both slots are collected; ms_parse gets the wall time of the run (the kernels of neighbouring
// batches overlap, so it is less than the sum of the kernel times).
int drain_parse(kc_handle *h) {
  const int first = (int)(h->n_batches & 1);  // the older batch sits in the slot the next batch would take
  cudaEvent_t last = nullptr;
  for (int i = 0; i < 2; i++) {
    ParseSlot &sl = h->slot[(first + i) & 1];
    if (sl.busy) last = sl.e1;
    RC(collect_slot(h, sl));
  }
  if (h->stage_open && last) h->st.ms_parse += elapsed(h->ev_first, last);
  h->stage_open = false;
  return KC_OK;
}

// Parse one batch that is already on the device (or will be once `in_ready` has happened). `bytes` may be a virtual base
// (bytes + start always lands inside the staged range). Either offsets (direct mode) or seg_* (segment mode) is given.
// Returns after the batch's fill kernel has been LAUNCHED: the input must stay valid until `consumed` (recorded behind the
// fill kernel when given) or the next drain_parse. `temps`: device arrays the batch reads, released when it is collected.
int parse_batch(kc_handle *h, const uint8_t *d_bytes, const u64 *d_offsets, const u64 *d_seg_start, const u32 *d_seg_len,
                const uint16_t *d_seg_depth, u64 nseq, int max_len, bool ascii, bool is_ctg, cudaEvent_t in_ready = nullptr,
                cudaEvent_t consumed = nullptr, std::vector<DevBuf> *temps = nullptr) {
  if (nseq == 0) return KC_OK;
  const u32 NP = h->P * (u32)h->cfg.num_shards;
  const int N = h->cfg.num_shards;
  ParseSlot &sl = h->slot[h->n_batches & 1];
  RC(collect_slot(h, sl));  // the batch before last: long finished, this only books it
  h->n_batches++;
  cudaStream_t st = sl.st;
  if (temps) sl.temps.swap(*temps);
  RC(ensure(h, sl.d_cursor, sizeof(u32) * NP));
  RC(ensure(h, sl.d_tot, sizeof(u64) * (size_t)(N + 1)));
  ParseArgs a;
  memset(&a, 0, sizeof(a));
  a.bytes = d_bytes;
  a.offsets = d_offsets;
  a.seg_start = d_seg_start;
  a.seg_len = d_seg_len;
  a.seg_depth = d_seg_depth;
  a.nseq = nseq;
  a.ascii = ascii ? 1 : 0;
  a.k = h->cfg.kmer_len;
  a.m = h->m;
  a.w = h->w;
  const u32 cutoff = (u32)std::min(std::max(h->cfg.qual_cutoff, 0), 64);
  a.qual_add = (0x80u - cutoff) * 0x01010101u;
  a.num_shards = (u32)h->cfg.num_shards;
  a.num_parts = h->P;
  if (max_len > PARSE_TILE) return fail(h, KC_ERR_INVALID, "internal: sequence of %d bases reached a parse thread", max_len);
  const int n16 = (15 + std::max(max_len, 16) + 15) >> 4;  // 16-byte chunks of the longest sequence at the worst alignment
  a.nsb = n16 + 4;
  a.nsv = (n16 + 1) / 2 + 2;
  a.nwb = h->geom.nwb;
  a.nwv = h->geom.nwv;
  a.rw = h->geom.rw;
  a.instances = (u64 *)h->d_misc.p;
  a.err = (u32 *)((char *)h->d_misc.p + 8);
  const int words_per_thread = a.nsb + a.nsv + 2 * a.w + PARSE_LCAP / 2;
  int threads = 128;
  while (threads > 32 && (size_t)threads * words_per_thread * 4 > 96u * 1024u) threads >>= 1;
  const size_t smem = (size_t)threads * words_per_thread * 4;
  const size_t smem_fill = (size_t)threads * (a.nsb + a.nsv) * 4;  // the fill kernel keeps only the packed sequence
  void (*k_count)(const ParseArgs) = threads == 128 ? parse_kernel<128> : threads == 64 ? parse_kernel<64> : parse_kernel<32>;
  void (*k_fill)(const ParseArgs) = threads == 128 ? fill_kernel<128> : threads == 64 ? fill_kernel<64> : fill_kernel<32>;
  CK(cudaFuncSetAttribute(k_count, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_fill, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_fill));
  int per_sm = 1, per_sm_fill = 1;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_count, threads, smem));
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_fill, k_fill, threads, smem_fill));
  per_sm = std::max(per_sm, 1);
  per_sm_fill = std::max(per_sm_fill, 1);
  const u64 need_blocks = (nseq + threads - 1) / threads;
  // The kernels of neighbouring batches overlap where one leaves SM slots free (tails, the memory-bound fill kernel's
  // stalls). Measured on C2 (r2_v2 A/B): full-size grids 79.5 ms per step, grids of half an SM's slots each
  // (KC_PARSE_SHARE=1) 95.9 ms — the ALU-bound parse kernel needs all its warps; so full-size grids are the default.
  static const int share = getenv("KC_PARSE_SHARE") ? atoi(getenv("KC_PARSE_SHARE")) : 0;
  const int res_parse = share ? std::max(per_sm / 2, 1) : per_sm * 2, res_fill = share ? std::max(per_sm_fill / 2, 1) : per_sm_fill * 2;
  const unsigned grid = (unsigned)std::min<u64>(need_blocks, (u64)h->sms * res_parse);
  const unsigned grid_fill = (unsigned)std::min<u64>(need_blocks, (u64)h->sms * res_fill);
  // supermer lists handed from the parse kernel to the fill kernel: one u32 per supermer, ecap slots per sequence
  if (NP >= (1u << 24)) return fail(h, KC_ERR_INVALID, "internal: %u buckets do not fit the 24-bit list entries", NP);
  a.ecap = std::max(1, max_len - a.k - 1);
  RC(ensure(h, sl.d_elist, need_blocks * (size_t)threads * (size_t)a.ecap * 4));
  RC(ensure(h, sl.d_ecnt, nseq));
  a.elist = (u32 *)sl.d_elist.p;
  a.ecnt = (uint8_t *)sl.d_ecnt.p;

  if (in_ready) CK(cudaStreamWaitEvent(st, in_ready, 0));
  if (!h->stage_open) {
    CK(cudaEventRecord(h->ev_first, st));
    h->stage_open = true;
  }
  CK(cudaEventRecord(sl.e0, st));
  // parse_kernel: bucket histogram + supermer lists
  Block blk;
  blk.is_ctg = is_ctg;
  CK(dev_alloc((void **)&blk.d_hist, sizeof(BucketHist) * NP));
  CK(cudaMemsetAsync(blk.d_hist, 0, sizeof(BucketHist) * NP, st));
  a.hist = blk.d_hist;
  k_count<<<grid, threads, smem, st>>>(a);
  h->st.gpu_launches++;
  h->st.n_parse_launches++;
  CK(cudaGetLastError());
  CK(cudaEventRecord(sl.ec, st));
  // bucket layout on the device: the host only needs the totals (allocation size, the range that goes to each shard)
  CK(dev_alloc((void **)&blk.d_base, sizeof(u64) * ((size_t)NP + 1)));
  RC(ensure(h, sl.d_scan, sizeof(u64) * 1032));
  RC(bucket_scan(h, blk.d_hist, NP, blk.d_base, h->P, (u64 *)sl.d_tot.p, (u64 *)sl.d_scan.p, st, (u32 *)sl.d_cursor.p));
  CK(cudaMemcpyAsync(sl.h_tot, sl.d_tot.p, sizeof(u64) * (size_t)(N + 1), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));  // the previous batch's fill kernel keeps running on the other slot's stream meanwhile
  blk.dest_base.assign(sl.h_tot, sl.h_tot + N + 1);
  const u64 hs = blk.dest_base[N];
  blk.tot_sup = hs;
  if (hs == 0) {
    dev_free(blk.d_hist);
    dev_free(blk.d_base);
    for (DevBuf &b : sl.temps) release(b);
    sl.temps.clear();
    if (consumed) CK(cudaEventRecord(consumed, st));
    return KC_OK;
  }
  const size_t rec_bytes = (size_t)hs * a.rw * 4;
  CK(dev_alloc((void **)&blk.rec, rec_bytes + 64));
  if (hs >> 32) return fail(h, KC_ERR_INVALID, "internal: %llu supermers in one batch do not fit the 32-bit bucket cursors", (unsigned long long)hs);
  // fill_kernel: records into the exact-size buckets
  a.cursor = (u32 *)sl.d_cursor.p;
  {
    static const int hint = getenv("KC_FILL_L2_HINT") ? atoi(getenv("KC_FILL_L2_HINT")) : 1;  // A/B switch (tools/prof_run.py)
    a.l2_hint = hint;
    static const int pair = getenv("KC_FILL_PAIR") ? atoi(getenv("KC_FILL_PAIR")) : 1;
    a.pair_lanes = pair;
  }
  a.rec = blk.rec;
  CK(cudaEventRecord(sl.ed, st));
  k_fill<<<grid_fill, threads, smem_fill, st>>>(a);
  h->st.gpu_launches++;
  h->st.n_parse_launches++;
  CK(cudaGetLastError());
  CK(cudaEventRecord(sl.e1, st));
  if (consumed) CK(cudaEventRecord(consumed, st));
  sl.blk = std::move(blk);
  sl.rec_bytes = rec_bytes;
  sl.busy = true;
  if (h->comm) {  // multi-shard with NCCL: this batch's supermers leave now, while the next batch is copied in and parsed
    h->phase_open = true;
    h->ev_blk = sl.e1;
    int rcx = exchange_round(h, &sl.blk, false, nullptr);
    if (rcx) return rcx;
  }
  return KC_OK;
}

}  // namespace

extern "C" int kc_add_reads_device(kc_handle *h, const uint8_t *d_bytes, const uint64_t *d_offsets, uint64_t n_reads,
                                   uint64_t total_bytes, uint32_t max_read_len) {
  if (!h) return KC_ERR_INVALID;
  if (h->finished) return fail(h, KC_ERR_STATE, "add after finish");
  if (n_reads == 0) return KC_OK;
  if (max_read_len > (uint32_t)SEG_MAX)
    return fail(h, KC_ERR_INVALID, "device-resident reads longer than %d must be tiled by the caller (use kc_add_reads)", SEG_MAX);
  CK(cudaSetDevice(h->cfg.device));
  const u64 k1 = (u64)h->cfg.kmer_len + 1;
  decide_partitions(h, total_bytes > n_reads * k1 ? total_bytes - n_reads * k1 : 0);
  // a batch is the unit of overlap: its fill kernel runs beside the next batch's parse kernel, its exchange behind both
  static const u64 def_batch = getenv("KC_BATCH_READS") ? (u64)atoll(getenv("KC_BATCH_READS")) : (2ull << 20);
  const u64 batch = h->cfg.batch_reads > 0 ? (u64)h->cfg.batch_reads : def_batch;
  for (u64 r0 = 0; r0 < n_reads; r0 += batch) {
    const u64 n = std::min<u64>(batch, n_reads - r0);
    int rc = parse_batch(h, d_bytes, (const u64 *)d_offsets + r0, nullptr, nullptr, nullptr, n, (int)max_read_len, false, false);
    if (rc) return rc;
  }
  RC(drain_parse(h));  // the caller owns the input again when this returns
  h->st.num_seqs += n_reads;
  h->st.num_bases += total_bytes;
  return KC_OK;
}

namespace {

// Tile a sequence of `len` bases starting at absolute byte `start` into segments of at most `tile` bases that overlap by
// k+1 bases, so that every counted k-mer position 1..len-k-1 belongs to exactly one segment (the reference GPU path
// refuses sequences longer than its block instead: kcount_gpu.cpp:138-140).
void tile_sequence(u64 start, u64 len, int k, int tile, uint16_t depth, std::vector<u64> &ss, std::vector<u32> &sl,
                   std::vector<uint16_t> &sd) {
  if (len < (u64)k + 2) return;
  u64 s = 0;
  for (;;) {
    u64 l = std::min<u64>((u64)tile, len - s);
    ss.push_back(start + s);
    sl.push_back((u32)l);
    sd.push_back(depth);
    if (s + l == len) break;
    s += l - (u64)k - 1;
  }
}

// copy host arrays of segments to the device and parse them (the arrays live until the batch is collected)
int parse_segments(kc_handle *h, const uint8_t *d_bytes_virtual, std::vector<u64> &ss, std::vector<u32> &sl, std::vector<uint16_t> &sd,
                   int max_len, bool ascii, bool is_ctg, cudaEvent_t in_ready = nullptr, cudaEvent_t consumed = nullptr) {
  const u64 n = ss.size();
  if (!n) return KC_OK;
  std::vector<DevBuf> t(3);
  int rc = ensure(h, t[0], n * 8);
  if (!rc) rc = ensure(h, t[1], n * 4);
  if (!rc) rc = ensure(h, t[2], n * 2);
  if (rc) {
    for (DevBuf &b : t) release(b);
    return rc;
  }
  // pageable sources: these copies return once the source has been staged; the slot's stream orders them before its kernels
  cudaStream_t st = h->slot[h->n_batches & 1].st;
  RC(collect_slot(h, h->slot[h->n_batches & 1]));  // the stream must be done with the batch before last
  cudaMemcpyAsync(t[0].p, ss.data(), n * 8, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(t[1].p, sl.data(), n * 4, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(t[2].p, sd.data(), n * 2, cudaMemcpyHostToDevice, st);
  const u64 *p0 = (const u64 *)t[0].p;
  const u32 *p1 = (const u32 *)t[1].p;
  const uint16_t *p2 = (const uint16_t *)t[2].p;
  return parse_batch(h, d_bytes_virtual, nullptr, p0, p1, p2, n, max_len, ascii, is_ctg, in_ready, consumed, &t);
}

}  // namespace

extern "C" int kc_add_reads(kc_handle *h, const uint8_t *bytes, const uint64_t *offsets, uint64_t n_reads) {
  if (!h) return KC_ERR_INVALID;
  if (h->finished) return fail(h, KC_ERR_STATE, "add after finish");
  if (n_reads == 0) return KC_OK;
  if (!bytes || !offsets) return fail(h, KC_ERR_INVALID, "null input");
  CK(cudaSetDevice(h->cfg.device));
  const int k = h->cfg.kmer_len;
  {  // instances ~ bases - reads * (k + 1); exact when every read has at least k + 1 bases (only sizes the partitions)
    const u64 bases = offsets[n_reads] - offsets[0], k1 = (u64)k + 1;
    decide_partitions(h, bases > n_reads * k1 ? bases - n_reads * k1 : 0);
  }
  u64 batch = h->cfg.batch_reads > 0 ? (u64)h->cfg.batch_reads : (2ull << 20);
  const u64 avg = std::max<u64>(1, (offsets[n_reads] - offsets[0]) / n_reads);
  batch = std::max<u64>(1, std::min<u64>(batch, (1ull << 30) / avg));  // <= ~1 GB of bases per batch
  const u64 nb = (n_reads + batch - 1) / batch;
  std::vector<cudaEvent_t> t0(nb), t1(nb);  // copy times, read once everything has drained
  for (u64 i = 0; i < nb; i++) {
    CK(cudaEventCreate(&t0[i]));
    CK(cudaEventCreate(&t1[i]));
  }
  // two staging buffers: the copy of batch i+1 overlaps the parse of batch i; a buffer is overwritten only after the FILL
  // kernel of the batch that used it (it reads the bases again), which the copy stream waits for on the device
  auto issue_copy = [&](u64 bi) -> int {
    const int sl = (int)(bi & 1);
    const u64 r0 = bi * batch, r1 = std::min<u64>(n_reads, r0 + batch);
    const u64 b0 = offsets[r0], b1 = offsets[r1];
    if (h->stage_bytes[sl].bytes < b1 - b0 + 64 || h->stage_offs[sl].bytes < (r1 - r0 + 1) * 8) RC(drain_parse(h));  // about to reallocate
    RC(ensure(h, h->stage_bytes[sl], b1 - b0 + 64));
    RC(ensure(h, h->stage_offs[sl], (r1 - r0 + 1) * 8));
    CK(cudaStreamWaitEvent(h->s_copy, h->ev_free[sl], 0));
    CK(cudaEventRecord(t0[bi], h->s_copy));
    CK(cudaMemcpyAsync(h->stage_bytes[sl].p, bytes + b0, b1 - b0, cudaMemcpyHostToDevice, h->s_copy));
    CK(cudaMemcpyAsync(h->stage_offs[sl].p, offsets + r0, (r1 - r0 + 1) * 8, cudaMemcpyHostToDevice, h->s_copy));
    CK(cudaEventRecord(t1[bi], h->s_copy));
    CK(cudaEventRecord(h->ev_copy[sl], h->s_copy));
    return KC_OK;
  };
  int rc = issue_copy(0);
  if (rc) return rc;
  for (u64 bi = 0; bi < nb; bi++) {
    const int sl = (int)(bi & 1);
    const u64 r0 = bi * batch, r1 = std::min<u64>(n_reads, r0 + batch);
    if (bi + 1 < nb) {  // the next batch's copy runs while this batch is parsed
      rc = issue_copy(bi + 1);
      if (rc) return rc;
    }
    const uint8_t *vbase = (const uint8_t *)h->stage_bytes[sl].p - offsets[r0];
    u64 max_len = 0;  // scanned while the copies are in flight
    for (u64 r = r0; r < r1; r++) max_len = std::max<u64>(max_len, offsets[r + 1] - offsets[r]);
    const bool tiled = max_len > (u64)SEG_MAX;
    if (!tiled) {
      rc = parse_batch(h, vbase, (const u64 *)h->stage_offs[sl].p, nullptr, nullptr, nullptr, r1 - r0, (int)max_len, false, false,
                       h->ev_copy[sl], h->ev_free[sl]);
    } else {
      std::vector<u64> ss;
      std::vector<u32> sl2;
      std::vector<uint16_t> sd;
      for (u64 r = r0; r < r1; r++) tile_sequence(offsets[r], offsets[r + 1] - offsets[r], k, SEG_MAX, 1, ss, sl2, sd);
      if (ss.empty()) CK(cudaEventRecord(h->ev_free[sl], h->s_copy));  // nothing reads this stage
      rc = parse_segments(h, vbase, ss, sl2, sd, SEG_MAX, false, false, h->ev_copy[sl], h->ev_free[sl]);
    }
    if (rc) return rc;
  }
  RC(drain_parse(h));
  CK(cudaStreamSynchronize(h->s_copy));
  float ms_copy = 0;
  for (u64 i = 0; i < nb; i++) {
    ms_copy += elapsed(t0[i], t1[i]);
    cudaEventDestroy(t0[i]);
    cudaEventDestroy(t1[i]);
  }
  h->st.ms_h2d += ms_copy;
  h->st.num_seqs += n_reads;
  h->st.num_bases += offsets[n_reads] - offsets[0];
  return KC_OK;
}

// ---- PackedReads resident in HBM across the k rounds of a job ----
struct kc_reads {
  int device = 0;
  uint8_t *d_bytes = nullptr;
  u64 *d_offsets = nullptr;
  u64 n_reads = 0, total_bytes = 0, batch = 0;
  cudaStream_t stream = nullptr;
  std::vector<cudaEvent_t> ready;   // per batch: its bytes and offsets have arrived
  std::vector<u32> max_len;         // per batch
  std::vector<u64> h_offsets;       // kept only when some read is longer than one parse tile (host-side tiling needs them)
  u32 *d_err = nullptr;             // kc_reads_create_fastq: set by the packing kernel on an illegal base character
  bool err_checked = false;
  void *stage[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};  // kc_reads_create_fastq: ASCII seq / qual staging, two slots
};

namespace {
// FASTQ text -> PackedRead bytes (PackedRead::PackedRead, packed_reads.cpp:73-109): low 3 bits A0 C1 G2 T3, N and the IUPAC
// codes 4; high 5 bits min(qual - qual_offset, 31) with the reference's unsigned-char wrap for negative differences.
// Anything else is the reference's DIE("Illegal char"): reported through *err. 16 bases per thread, 128-bit accesses.
__device__ __forceinline__ u32 pack_base(u32 c, u32 q, int qual_offset, u32 *err) {
  u32 code;
  switch (c) {
    case 'A': code = 0; break;
    case 'C': code = 1; break;
    case 'G': code = 2; break;
    case 'T': code = 3; break;
    case 'N': case 'U': case 'R': case 'Y': case 'K': case 'M': case 'S': case 'W': case 'B': case 'D': case 'H': case 'V': code = 4; break;
    default: code = 4; *err = 1; break;
  }
  int d = (int)(signed char)q - qual_offset;  // `char` is signed on the reference's platforms
  if (d > 31) d = 31;
  return (code | (((u32)(unsigned char)d) << 3)) & 0xFFu;
}

// seq, qual and out must be congruent modulo 16 (the staging copy is placed accordingly): a byte-wise head up to the first
// 16-byte boundary, 128-bit body, byte-wise tail
__global__ void __launch_bounds__(256) pack_reads_kernel(const uint8_t *seq, const uint8_t *qual, u64 n, int qual_offset, uint8_t *out,
                                                         u32 *err) {
  const u64 head = min(n, (u64)((16 - ((size_t)out & 15)) & 15));
  const u64 nvec = (n - head) / 16, tail0 = head + nvec * 16;
  const uint4 *s16 = (const uint4 *)(seq + head), *q16 = (const uint4 *)(qual + head);
  uint4 *o16 = (uint4 *)(out + head);
  for (u64 v = (u64)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (u64)gridDim.x * blockDim.x) {
    const uint4 s4 = s16[v], q4 = q16[v];
    const u32 sw[4] = {s4.x, s4.y, s4.z, s4.w}, qw[4] = {q4.x, q4.y, q4.z, q4.w};
    u32 ow[4];
#pragma unroll
    for (int w = 0; w < 4; w++) {
      u32 o = 0;
#pragma unroll
      for (int b = 0; b < 4; b++) o |= pack_base((sw[w] >> (8 * b)) & 0xFFu, (qw[w] >> (8 * b)) & 0xFFu, qual_offset, err) << (8 * b);
      ow[w] = o;
    }
    o16[v] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
  }
  if (blockIdx.x == 0) {
    if (threadIdx.x < head) out[threadIdx.x] = (uint8_t)pack_base(seq[threadIdx.x], qual[threadIdx.x], qual_offset, err);
    const u64 i = tail0 + threadIdx.x;
    if (threadIdx.x < 16 && i < n) out[i] = (uint8_t)pack_base(seq[i], qual[i], qual_offset, err);
  }
}
}  // namespace

extern "C" int kc_reads_create(int32_t device, const uint8_t *bytes, const uint64_t *offsets, uint64_t n_reads, int32_t batch_reads,
                               kc_reads **out) {
  if (!out) return KC_ERR_INVALID;
  *out = nullptr;
  if (!bytes || !offsets || offsets[0] != 0) return KC_ERR_INVALID;
  kc_handle *h = nullptr;  // CK reports into a handle; there is none here
  kc_reads *r = new kc_reads();
  *out = r;
  r->device = device;
  r->n_reads = n_reads;
  r->total_bytes = offsets[n_reads];
  CK(cudaSetDevice(device));
  CK(cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking));
  CK(dev_alloc((void **)&r->d_bytes, r->total_bytes + 64));
  CK(dev_alloc((void **)&r->d_offsets, (n_reads + 1) * 8));
  u64 batch = batch_reads > 0 ? (u64)batch_reads : (2ull << 20);
  const u64 avg = std::max<u64>(1, r->total_bytes / std::max<u64>(n_reads, 1));
  batch = std::max<u64>(1, std::min<u64>(batch, (1ull << 30) / avg));
  r->batch = batch;
  const u64 nb = (n_reads + batch - 1) / batch;
  bool any_long = false;
  for (u64 bi = 0; bi < nb; bi++) {  // all copies are queued first; the length scan below runs while they are in flight
    const u64 r0 = bi * batch, r1 = std::min<u64>(n_reads, r0 + batch);
    CK(cudaMemcpyAsync(r->d_bytes + offsets[r0], bytes + offsets[r0], offsets[r1] - offsets[r0], cudaMemcpyHostToDevice, r->stream));
    CK(cudaMemcpyAsync(r->d_offsets + r0, offsets + r0, (r1 - r0 + 1) * 8, cudaMemcpyHostToDevice, r->stream));
    cudaEvent_t e;
    CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CK(cudaEventRecord(e, r->stream));
    r->ready.push_back(e);
  }
  for (u64 bi = 0; bi < nb; bi++) {
    const u64 r0 = bi * batch, r1 = std::min<u64>(n_reads, r0 + batch);
    u64 ml = 0;
    for (u64 i = r0; i < r1; i++) ml = std::max<u64>(ml, offsets[i + 1] - offsets[i]);
    r->max_len.push_back((u32)std::min<u64>(ml, 0xFFFFFFFFu));
    any_long = any_long || ml > (u64)SEG_MAX;
  }
  if (any_long) r->h_offsets.assign(offsets, offsets + n_reads + 1);
  return KC_OK;
}

extern "C" int kc_reads_destroy(kc_reads *r) {
  if (!r) return KC_OK;
  cudaSetDevice(r->device);
  if (r->stream) cudaStreamSynchronize(r->stream);
  for (auto e : r->ready) cudaEventDestroy(e);
  dev_free(r->d_bytes);
  dev_free(r->d_offsets);
  dev_free(r->d_err);
  for (int sl = 0; sl < 2; sl++)
    for (int j = 0; j < 2; j++) dev_free(r->stage[sl][j]);
  if (r->stream) cudaStreamDestroy(r->stream);
  delete r;
  return KC_OK;
}

extern "C" int kc_reads_create_fastq(int32_t device, const char *seqs, const char *quals, const uint64_t *offsets, uint64_t n_reads,
                                     int32_t qual_offset, int32_t batch_reads, kc_reads **out) {
  if (!out) return KC_ERR_INVALID;
  *out = nullptr;
  if (!seqs || !quals || !offsets || offsets[0] != 0) return KC_ERR_INVALID;
  kc_handle *h = nullptr;
  kc_reads *r = new kc_reads();
  *out = r;
  r->device = device;
  r->n_reads = n_reads;
  r->total_bytes = offsets[n_reads];
  CK(cudaSetDevice(device));
  CK(cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking));
  CK(dev_alloc((void **)&r->d_bytes, r->total_bytes + 64));
  CK(dev_alloc((void **)&r->d_offsets, (n_reads + 1) * 8));
  CK(dev_alloc((void **)&r->d_err, 64));
  CK(cudaMemsetAsync(r->d_err, 0, 64, r->stream));
  u64 batch = batch_reads > 0 ? (u64)batch_reads : (2ull << 20);
  const u64 avg = std::max<u64>(1, r->total_bytes / std::max<u64>(n_reads, 1));
  batch = std::max<u64>(1, std::min<u64>(batch, (1ull << 30) / avg));
  r->batch = batch;
  const u64 nb = (n_reads + batch - 1) / batch;
  u64 slot_bytes = 0;
  for (u64 bi = 0; bi < nb; bi++) {
    const u64 r0 = bi * batch, r1 = std::min<u64>(n_reads, r0 + batch);
    slot_bytes = std::max<u64>(slot_bytes, offsets[r1] - offsets[r0]);
  }
  for (int sl = 0; sl < 2; sl++)
    for (int j = 0; j < 2; j++) CK(dev_alloc(&r->stage[sl][j], slot_bytes + 64));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  bool any_long = false;
  for (u64 bi = 0; bi < nb; bi++) {  // one stream: copy, copy, pack: slot reuse is ordered by the stream itself
    const u64 r0 = bi * batch, r1 = std::min<u64>(n_reads, r0 + batch);
    const u64 b0 = offsets[r0], nbytes = offsets[r1] - b0;
    const int sl = (int)(bi & 1);
    // the text is staged at the same offset modulo 16 as its destination, so that the kernel's 128-bit body is aligned on
    // both sides
    const u64 mis = b0 & 15;
    uint8_t *st_s = (uint8_t *)r->stage[sl][0] + mis, *st_q = (uint8_t *)r->stage[sl][1] + mis;
    CK(cudaMemcpyAsync(st_s, seqs + b0, nbytes, cudaMemcpyHostToDevice, r->stream));
    CK(cudaMemcpyAsync(st_q, quals + b0, nbytes, cudaMemcpyHostToDevice, r->stream));
    CK(cudaMemcpyAsync(r->d_offsets + r0, offsets + r0, (r1 - r0 + 1) * 8, cudaMemcpyHostToDevice, r->stream));
    if (nbytes) pack_reads_kernel<<<sms * 8, 256, 0, r->stream>>>(st_s, st_q, nbytes, qual_offset, r->d_bytes + b0, r->d_err);
    cudaEvent_t e;
    CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CK(cudaEventRecord(e, r->stream));
    r->ready.push_back(e);
  }
  CK(cudaGetLastError());
  for (u64 bi = 0; bi < nb; bi++) {
    const u64 r0 = bi * batch, r1 = std::min<u64>(n_reads, r0 + batch);
    u64 ml = 0;
    for (u64 i = r0; i < r1; i++) ml = std::max<u64>(ml, offsets[i + 1] - offsets[i]);
    r->max_len.push_back((u32)std::min<u64>(ml, 0xFFFFFFFFu));
    any_long = any_long || ml > (u64)SEG_MAX;
  }
  if (any_long) r->h_offsets.assign(offsets, offsets + n_reads + 1);
  return KC_OK;
}

namespace {
int reads_check(kc_reads *r) {  // the packing kernel's verdict (waits for the upload)
  if (!r->d_err || r->err_checked) return KC_OK;
  u32 e = 0;
  if (cudaStreamSynchronize(r->stream) != cudaSuccess) return KC_ERR_CUDA;
  if (cudaMemcpy(&e, r->d_err, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return KC_ERR_CUDA;
  r->err_checked = true;
  return e ? KC_ERR_BADCHAR : KC_OK;
}
}  // namespace

extern "C" int kc_reads_copy_packed(kc_reads *r, uint8_t *bytes_out) {
  if (!r || !bytes_out) return KC_ERR_INVALID;
  if (cudaSetDevice(r->device) != cudaSuccess) return KC_ERR_CUDA;
  if (cudaStreamSynchronize(r->stream) != cudaSuccess) return KC_ERR_CUDA;
  if (cudaMemcpy(bytes_out, r->d_bytes, r->total_bytes, cudaMemcpyDeviceToHost) != cudaSuccess) return KC_ERR_CUDA;
  return reads_check(r);
}

extern "C" int kc_add_reads_resident(kc_handle *h, kc_reads *r) {
  if (!h || !r) return KC_ERR_INVALID;
  if (h->finished) return fail(h, KC_ERR_STATE, "add after finish");
  if (r->device != h->cfg.device) return fail(h, KC_ERR_INVALID, "reads live on device %d, the handle on %d", r->device, h->cfg.device);
  if (r->n_reads == 0) return KC_OK;
  CK(cudaSetDevice(h->cfg.device));
  const int k = h->cfg.kmer_len;
  const u64 k1 = (u64)k + 1;
  decide_partitions(h, r->total_bytes > r->n_reads * k1 ? r->total_bytes - r->n_reads * k1 : 0);
  for (u64 bi = 0; bi < r->ready.size(); bi++) {
    const u64 r0 = bi * r->batch, r1 = std::min<u64>(r->n_reads, r0 + r->batch);
    int rc;  // batches wait for their upload on the device: first use right after kc_reads_create overlaps the upload
    if (r->max_len[bi] <= (u32)SEG_MAX) {
      rc = parse_batch(h, r->d_bytes, r->d_offsets + r0, nullptr, nullptr, nullptr, r1 - r0, (int)r->max_len[bi], false, false,
                       r->ready[bi]);
    } else {
      std::vector<u64> ss;
      std::vector<u32> sl;
      std::vector<uint16_t> sd;
      for (u64 i = r0; i < r1; i++) tile_sequence(r->h_offsets[i], r->h_offsets[i + 1] - r->h_offsets[i], k, SEG_MAX, 1, ss, sl, sd);
      rc = parse_segments(h, r->d_bytes, ss, sl, sd, SEG_MAX, false, false, r->ready[bi]);
    }
    if (rc) return rc;
  }
  RC(drain_parse(h));
  h->st.num_seqs += r->n_reads;
  h->st.num_bases += r->total_bytes;
  if (int rcx = reads_check(r)) return fail(h, rcx, "kc_reads_create_fastq: illegal base character in the reads (packed_reads.cpp:104)");
  return KC_OK;
}

namespace {

// ASCII sequences (ACGTN, lower case = low quality): contigs with their depths, or reads as count_kmers hands them to
// process_seq (kcount.cpp:81-86). Any length; sequences longer than one parse thread's tile are cut with k+1 overlap.
int add_ascii(kc_handle *h, const char *seqs, const uint64_t *offsets, const double *depths, uint64_t n_seqs, bool is_ctg) {
  if (!h) return KC_ERR_INVALID;
  if (h->finished) return fail(h, KC_ERR_STATE, "add after finish");
  if (n_seqs == 0) return KC_OK;
  if (!seqs || !offsets || (is_ctg && !depths)) return fail(h, KC_ERR_INVALID, "null input");
  CK(cudaSetDevice(h->cfg.device));
  const int k = h->cfg.kmer_len;
  std::vector<u64> ss;
  std::vector<u32> sl;
  std::vector<uint16_t> sd;
  u64 inst = 0;
  for (u64 c = 0; c < n_seqs; c++) {
    const u64 len = offsets[c + 1] - offsets[c];
    if (len < (u64)k + 2) continue;  // kcount.cpp:126
    uint16_t depth = 1;
    if (is_ctg) {
      const double d = depths[c];
      depth = (uint16_t)(d > 65535 ? 65535 : d);  // contigs.hpp:57-66 get_uint16_t_depth
      if (!depth) depth = 1;                       // kcount_cpu.cpp:75
    }
    inst += len - k - 1;
    tile_sequence(offsets[c], len, k, SEG_MAX, depth, ss, sl, sd);
  }
  decide_partitions(h, inst);
  if (ss.empty()) return KC_OK;
  const u64 b0 = offsets[0], b1 = offsets[n_seqs];
  DevBuf d_seq;
  int rc = ensure(h, d_seq, b1 - b0 + 64);
  if (rc) return rc;
  CK(cudaMemcpyAsync(d_seq.p, seqs + b0, b1 - b0, cudaMemcpyHostToDevice, h->s_main));
  CK(cudaStreamSynchronize(h->s_main));  // the parse slots' streams read it
  // segments are parsed in slices so that the per-batch bucket buffers stay bounded
  const size_t SLICE = 2u << 20;
  for (size_t s0 = 0; s0 < ss.size() && !rc; s0 += SLICE) {
    const size_t s1 = std::min(ss.size(), s0 + SLICE);
    std::vector<u64> a(ss.begin() + s0, ss.begin() + s1);
    std::vector<u32> b(sl.begin() + s0, sl.begin() + s1);
    std::vector<uint16_t> c(sd.begin() + s0, sd.begin() + s1);
    rc = parse_segments(h, (const uint8_t *)d_seq.p - b0, a, b, c, SEG_MAX, true, is_ctg);
  }
  {
    const int rcd = drain_parse(h);
    if (!rc) rc = rcd;
  }
  release(d_seq);
  if (rc) return rc;
  u32 err = 0;
  CK(cudaMemcpy(&err, (char *)h->d_misc.p + 8, 4, cudaMemcpyDeviceToHost));
  if (err) return fail(h, KC_ERR_BADCHAR, "sequence character outside ACGTN");
  if (is_ctg) h->has_ctg = true;
  h->st.num_seqs += n_seqs;
  h->st.num_bases += b1 - b0;
  return KC_OK;
}

}  // namespace

extern "C" int kc_add_contigs(kc_handle *h, const char *seqs, const uint64_t *offsets, const double *depths, uint64_t n_ctgs) {
  return add_ascii(h, seqs, offsets, depths, n_ctgs, true);
}

extern "C" int kc_add_reads_ascii(kc_handle *h, const char *seqs, const uint64_t *offsets, uint64_t n_reads) {
  return add_ascii(h, seqs, offsets, nullptr, n_reads, false);
}

extern "C" int kc_set_dmin_thres(kc_handle *h, int32_t dmin_thres) {
  if (!h) return KC_ERR_INVALID;
  if (h->finished) return fail(h, KC_ERR_STATE, "already finished");
  if (dmin_thres < 1) return fail(h, KC_ERR_INVALID, "dmin_thres must be >= 1");
  h->cfg.dmin_thres = dmin_thres;
  return KC_OK;
}

namespace {

template <int KW, bool WIDE, bool CTG, int KC, int WC>
int launch_count(kc_handle *h, const CountArgs &ca, size_t smem, unsigned grid) {
  CK(cudaFuncSetAttribute(count_kernel<KW, WIDE, CTG, KC, WC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  count_kernel<KW, WIDE, CTG, KC, WC><<<grid, CNT_THREADS, smem, h->s_main>>>(ca);
  CK(cudaGetLastError());
  return KC_OK;
}

// The reference's default k schedule (options.cpp:373: 21,33,55,77,99) and k=31 get kernels with k and the supermer window
// folded in at compile time (narrow counters; kc_extract.cuh); every other k, and the rare wide re-runs, take the
// run-time-k kernels.
int dispatch_count(kc_handle *h, bool wide, bool ctg, const CountArgs &ca, size_t smem, unsigned grid) {
  static const bool fixed_ok = !(getenv("KC_FIXED_K") && atoi(getenv("KC_FIXED_K")) == 0);  // A/B switch
#define GOKW(KWv, Kv, Wv)                                                                              \
  if (fixed_ok && !wide && h->cfg.kmer_len == Kv && h->w == Wv)                                        \
    return ctg ? launch_count<KWv, false, true, Kv, Wv>(h, ca, smem, grid) : launch_count<KWv, false, false, Kv, Wv>(h, ca, smem, grid)
#define GOK(KWv, Kv) GOKW(KWv, Kv, kc_w_of(Kv))
#define GO3(KWv)                                                                               \
  return wide ? (ctg ? launch_count<KWv, true, true, 0, 0>(h, ca, smem, grid) : launch_count<KWv, true, false, 0, 0>(h, ca, smem, grid)) \
              : (ctg ? launch_count<KWv, false, true, 0, 0>(h, ca, smem, grid) : launch_count<KWv, false, false, 0, 0>(h, ca, smem, grid))
  switch (h->KW) {
    case 1: GOK(1, 21); GOK(1, 31); GO3(1);
    case 2: GOK(2, 33); GOK(2, 55); GO3(2);
    case 3: GOK(3, 77); GO3(3);
    case 4: GOK(4, 99); GO3(4);
  }
#undef GO3
#undef GOK
#undef GOKW
  return fail(h, KC_ERR_INVALID, "unsupported key width %d", h->KW);
}

}  // namespace

int kc_exchange(kc_handle *h);  // kc_comm.inl
// where the surviving table goes when the caller wants it on the host as it is produced (kc_finish_to_host)
struct HostSink {
  u64 *keys = nullptr;
  uint16_t *counts = nullptr;
  char *left = nullptr, *right = nullptr;
  u64 cap = 0;     // entries the host arrays can take
  u64 copied = 0;  // entries already on their way
};

int finish_count(kc_handle *h, HostSink *sink = nullptr);

namespace {
int finish_impl(kc_handle *h, HostSink *sink);
}

extern "C" int kc_finish(kc_handle *h) { return finish_impl(h, nullptr); }

extern "C" int kc_finish_to_host(kc_handle *h, uint64_t *kmer_words, uint16_t *counts, char *left, char *right, uint64_t capacity,
                                 uint64_t *n_out) {
  if (!h || !kmer_words || !counts || !left || !right || !n_out) return KC_ERR_INVALID;
  HostSink sink;
  sink.keys = (u64 *)kmer_words;
  sink.counts = counts;
  sink.left = left;
  sink.right = right;
  sink.cap = capacity;
  *n_out = 0;
  int rc = finish_impl(h, &sink);
  if (rc) return rc;
  *n_out = h->n_surv;
  if (h->n_surv > capacity)
    return fail(h, KC_ERR_OVERFLOW, "%llu survivors do not fit the host arrays (%llu): the table is still on the device, use "
                "kc_copy_survivors", (unsigned long long)h->n_surv, (unsigned long long)capacity);
  return KC_OK;
}

namespace {
int finish_impl(kc_handle *h, HostSink *sink) {
  if (!h) return KC_ERR_INVALID;
  if (h->finished) return fail(h, KC_ERR_STATE, "finish called twice");
  CK(cudaSetDevice(h->cfg.device));
  if (h->cfg.num_shards > 1) {
    int rc = kc_exchange(h);
    if (rc) return rc;
  }
  if (!h->P) decide_partitions(h, 0);
  return finish_count(h, sink);
}
}  // namespace

namespace {
constexpr u32 SINK_SLICES = 8;
// cumulative slice boundaries in 64ths of the item range. Small slices at BOTH ends: when the copy is the longer of the two
// (k=33 on C2: 1.37 GB at 57 GB/s = 24 ms against 23 ms of counting) it should start as early as possible, when the count
// is longer (k=21) the copy nothing overlaps is the LAST slice's, so that one is 1/64 of the table
constexpr u32 SINK_BOUND64[SINK_SLICES + 1] = {0, 4, 12, 28, 44, 54, 60, 63, 64};

// survivors [sink->copied, upto) are final once `ready` has happened on the main stream: send them to the host on the copy
// stream while the main stream keeps counting
int sink_copy(kc_handle *h, HostSink *sink, u64 upto, cudaEvent_t ready) {
  const u64 n = std::min(upto, sink->cap), c = sink->copied;
  if (n <= c) return KC_OK;
  const int KW = h->KW;
  CK(cudaStreamWaitEvent(h->s_copy, ready, 0));
  CK(cudaMemcpyAsync(sink->keys + c * KW, (const u64 *)h->out_keys.p + c * KW, (n - c) * KW * 8, cudaMemcpyDeviceToHost, h->s_copy));
  CK(cudaMemcpyAsync(sink->counts + c, (const uint16_t *)h->out_counts.p + c, (n - c) * 2, cudaMemcpyDeviceToHost, h->s_copy));
  CK(cudaMemcpyAsync(sink->left + c, (const char *)h->out_left.p + c, n - c, cudaMemcpyDeviceToHost, h->s_copy));
  CK(cudaMemcpyAsync(sink->right + c, (const char *)h->out_right.p + c, n - c, cudaMemcpyDeviceToHost, h->s_copy));
  sink->copied = n;
  return KC_OK;
}
}  // namespace

// the count phase of one shard: every supermer that belongs to it is in h->blocks (own buckets) or h->rblocks (received)
int finish_count(kc_handle *h, HostSink *sink) {
  CK(cudaSetDevice(h->cfg.device));
  HostTimer ht;
  const u32 P = h->P;
  const int KW = h->KW;
  const u32 me = (u32)h->cfg.shard_id;
  // every batch that holds supermers of this shard: own buckets of the locally parsed ones, all of the received ones
  std::vector<BlockDesc> descs;
  bool any_ctg = false;
  u64 sup_expected = 0;
  for (auto &b : h->blocks) {
    const u64 mine = b.tot_sup ? b.dest_base[me + 1] - b.dest_base[me] : 0;
    if (!mine) continue;
    descs.push_back(BlockDesc{b.rec, b.d_hist, b.d_base, me * P, b.is_ctg ? 1u : 0u});
    any_ctg |= b.is_ctg;
    sup_expected += mine;
  }
  for (auto &b : h->rblocks) {
    if (!b.tot_sup) continue;
    descs.push_back(BlockDesc{b.rec, b.d_hist, b.d_base, 0u, b.is_ctg ? 1u : 0u});
    any_ctg |= b.is_ctg;
    sup_expected += b.tot_sup;
  }
  const u32 nblk = (u32)descs.size();
  if ((u64)P * std::max(nblk, 1u) > 0xFFFFFFF0ull) return fail(h, KC_ERR_INVALID, "too many batches (%u) for %u partitions", nblk, P);
  // work items and their segment lists are built on the device, where the bucket directories are (kc_count.cuh)
  DevBuf d_counters, d_segs, d_items, d_status, d_descs, d_heavy;
  RC(ensure(h, d_counters, CTR_WORDS * 8));
  CK(cudaMemsetAsync(d_counters.p, 0, CTR_WORDS * 8, h->s_main));
  RC(ensure(h, d_segs, (size_t)P * std::max(nblk, 1u) * sizeof(SegDesc) + 64));
  RC(ensure(h, d_items, (size_t)P * sizeof(WorkItem)));
  RC(ensure(h, d_status, (size_t)P * sizeof(ItemStatus)));
  RC(ensure(h, d_heavy, (size_t)P * 8));
  RC(ensure(h, d_descs, std::max<size_t>(nblk, 1) * sizeof(BlockDesc)));
  if (nblk) CK(cudaMemcpyAsync(d_descs.p, descs.data(), nblk * sizeof(BlockDesc), cudaMemcpyHostToDevice, h->s_main));
  CK(cudaMemsetAsync(d_status.p, 0, (size_t)P * sizeof(ItemStatus), h->s_main));
  u64 *ctr = h->h_pin;  // pinned
  BuildArgs ba;
  memset(&ba, 0, sizeof(ba));
  ba.blocks = (const BlockDesc *)d_descs.p;
  ba.nblk = nblk;
  ba.P = P;
  ba.rw = h->geom.rw;
  ba.segs = (SegDesc *)d_segs.p;
  ba.items = (WorkItem *)d_items.p;
  ba.heavy = (u64 *)d_heavy.p;
  ba.counters = (u64 *)d_counters.p;
  // Automatic partitioning: how many neighbouring buckets make one work item? Probe: count (without emitting anything) a
  // sample of single buckets spread over the partition space, take distinct / instances from it.
  u32 merge = 1;
  if (h->auto_merge && nblk && P > 1) {
    u32 tab_size, key_cap;
    table_geometry(h, false, any_ctg, tab_size, key_cap);
    const int VWp = VR_NARROW + (any_ctg ? V_CTG : 0);
    const size_t smem_p = count_smem_bytes(KW, VWp, h->geom.rw, tab_size, key_cap) + 8;
    struct Probe { double mean = 0, live = 0; u32 maxk = 0, over = 0; u64 inst = 0; u32 S = 0; };
    // count (emitting nothing) a sample of S candidate items of g neighbouring buckets each, spread over the partition space
    auto probe = [&](u32 g, Probe &pr) -> int {
      const u32 nq = (P + g - 1) / g;
      const u32 S = std::min<u32>(nq, 2u * (u32)h->sms);
      ba.nitems = S;
      ba.g = g;
      ba.step = (nq / S) * g;
      ba.heavy_thr = ~0ull;
      build_items_kernel<<<(S + 255) / 256, 256, 0, h->s_main>>>(ba);
      CountArgs pa;
      memset(&pa, 0, sizeof(pa));
      pa.items = (const WorkItem *)d_items.p;
      pa.segs = (const SegDesc *)d_segs.p;
      pa.status = (ItemStatus *)d_status.p;
      pa.nitems = S;
      pa.measure_only = 1;
      pa.k = h->cfg.kmer_len;
      pa.rw = h->geom.rw;
      pa.nwb = h->geom.nwb;
      pa.w = h->w;
      pa.tab_size = tab_size;
      pa.key_cap = key_cap;
      pa.dmin_thres = h->cfg.dmin_thres;
      pa.frac = 1.0 - h->cfg.dyn_min_depth;
      pa.counters = (u64 *)d_counters.p;
      RC(dispatch_count(h, false, any_ctg, pa, smem_p, std::min<unsigned>(S, (unsigned)h->sms * CNT_CTAS_PER_SM)));
      h->st.gpu_launches += 2;
      std::vector<ItemStatus> pst(S);
      CK(cudaMemcpyAsync(ctr, d_counters.p, CTR_WORDS * 8, cudaMemcpyDeviceToHost, h->s_main));
      CK(cudaMemcpyAsync(pst.data(), d_status.p, (size_t)S * sizeof(ItemStatus), cudaMemcpyDeviceToHost, h->s_main));
      CK(cudaStreamSynchronize(h->s_main));
      u64 sum = 0, live = 0;
      for (auto &x : pst) {
        sum += x.nkeys;
        if (!x.overflow) live += x.progress;
        pr.maxk = std::max(pr.maxk, x.nkeys);
        pr.over += x.overflow ? 1u : 0u;
      }
      pr.mean = (double)sum / S;
      pr.live = (double)live / S;
      pr.inst = ctr[CTR_READ_INST] + ctr[CTR_CTG_INST];
      pr.S = S;
      CK(cudaMemsetAsync(d_counters.p, 0, CTR_WORDS * 8, h->s_main));
      CK(cudaMemsetAsync(d_status.p, 0, (size_t)S * sizeof(ItemStatus), h->s_main));
      return KC_OK;
    };
    // Key slots an item needs = its distinct k-mers + the slots same-key races leave unused. The second term does not grow
    // with the item (a race needs two copies of a new key in flight at once: in a single bucket, counted in a handful of
    // warp iterations, that is a third of the keys; in a merged item it is the same few hundred slots), so single buckets
    // overstate what a merged item needs. Probe 1 (single buckets) reports both terms; the merge factor is sized from
    // distinct k-mers per bucket + that constant, then probe 2 runs real candidates of that size and the factor is accepted
    // only if every sampled item stays clear of the key store's capacity (else: one bucket fewer per item, re-probed once,
    // and finally the conservative factor that scales both terms).
    static const double fill = getenv("KC_TARGET_FILL") ? atof(getenv("KC_TARGET_FILL")) : 0.9;
    Probe p1, p2;
    RC(probe(1, p1));
    u32 g_safe = 1;
    if (!p1.over && p1.inst && p1.mean > 0) g_safe = (u32)std::min(64.0, std::max(1.0, std::floor(fill * key_cap / (1.1 * p1.mean))));
    merge = g_safe;
    u32 g_try = g_safe;
    if (!p1.over && p1.live > 0) {
      const double dead = std::max(0.0, p1.mean - p1.live);
      g_try = (u32)std::min(64.0, std::max((double)g_safe, std::floor((fill * key_cap / 1.1 - dead) / p1.live)));
      g_try = std::min<u32>(g_try, P);
    }
    for (int attempt = 0; attempt < 2 && g_try > g_safe; attempt++) {
      p2 = Probe();
      RC(probe(g_try, p2));
      if (!p2.over && p2.maxk <= 0.95 * key_cap && 1.1 * p2.mean <= fill * key_cap) {
        merge = g_try;
        break;
      }
      g_try--;
    }
    if (getenv("KC_TRACE"))
      fprintf(stderr, "[kc_trace] probe: P %u key_cap %u | g=1: %.0f slots/item, %.0f distinct (%u items, %llu inst) -> safe %u | g=%u: mean %.0f max %u"
              " over %u -> merge %u\n", P, key_cap, p1.mean, p1.live, p1.S, (unsigned long long)p1.inst, g_safe, g_try, p2.mean, p2.maxk, p2.over, merge);
    ht.lap("finish: partition probe");
  }
  const u32 NQ = (P + merge - 1) / merge;  // work items
  h->st.num_partitions = NQ;
  ba.nitems = NQ;
  ba.g = merge;
  ba.step = merge;
  // Hash partitions are statistically alike; only outliers (skewed data: heavy-hitter partitions) are scheduled ahead
  ba.heavy_thr = 2 * (sup_expected / NQ) + 1;
  build_items_kernel<<<(NQ + 255) / 256, 256, 0, h->s_main>>>(ba);
  h->st.gpu_launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(ctr, d_counters.p, CTR_WORDS * 8, cudaMemcpyDeviceToHost, h->s_main));
  CK(cudaStreamSynchronize(h->s_main));
  if (ctr[CTR_ERR]) return fail(h, KC_ERR_INVALID, "a partition holds too many supermers: use more partitions");
  const u64 read_inst = ctr[CTR_READ_INST], ctg_inst = ctr[CTR_CTG_INST];
  u32 n_heavy = (u32)ctr[CTR_NHEAVY];
  if (ctr[CTR_SUP] != sup_expected) return fail(h, KC_ERR_CUDA, "internal: bucket directories do not add up");
  if (n_heavy) {  // heaviest first
    std::vector<u64> hv(n_heavy);
    CK(cudaMemcpy(hv.data(), d_heavy.p, (size_t)n_heavy * 8, cudaMemcpyDeviceToHost));
    std::sort(hv.begin(), hv.end(), [](u64 x, u64 y) { return x > y; });
    CK(cudaMemcpyAsync(d_heavy.p, hv.data(), (size_t)n_heavy * 8, cudaMemcpyHostToDevice, h->s_main));
    CK(cudaStreamSynchronize(h->s_main));
  }
  ht.lap("finish: build items/segs (device)");
  h->st.instances_inserted = read_inst + ctg_inst;
  h->st.ctg_instances_inserted = ctg_inst;
  h->st.supermers_in = sup_expected;

  // outputs: a k-mer needs count >= 2 to survive, so survivors <= read instances / 2 + contig instances
  const u64 out_cap = read_inst / 2 + ctg_inst + 1024;
  RC(ensure(h, h->out_keys, out_cap * KW * 8));
  RC(ensure(h, h->out_counts, out_cap * 2));
  RC(ensure(h, h->out_left, out_cap));
  RC(ensure(h, h->out_right, out_cap));
  u64 pf_cap = 0;
  if (h->cfg.keep_prefilter) {
    pf_cap = read_inst + 1024;
    RC(ensure(h, h->pf_keys, pf_cap * KW * 8));
    RC(ensure(h, h->pf_vals, pf_cap * 18));
  }

  CountArgs ca;
  memset(&ca, 0, sizeof(ca));
  ca.segs = (const SegDesc *)d_segs.p;
  ca.k = h->cfg.kmer_len;
  ca.rw = h->geom.rw;
  ca.nwb = h->geom.nwb;
  ca.w = h->w;
  ca.dmin_thres = h->cfg.dmin_thres;
  ca.frac = 1.0 - h->cfg.dyn_min_depth;  // kcount_cpu.cpp:178, evaluated in fp64 exactly as the reference does
  ca.out_keys = (u64 *)h->out_keys.p;
  ca.out_counts = (uint16_t *)h->out_counts.p;
  ca.out_left = (char *)h->out_left.p;
  ca.out_right = (char *)h->out_right.p;
  ca.out_cap = out_cap;
  ca.counters = (u64 *)d_counters.p;
  ca.pf_keys = h->cfg.keep_prefilter ? (u64 *)h->pf_keys.p : nullptr;
  ca.pf_vals = (uint16_t *)h->pf_vals.p;
  ca.pf_cap = pf_cap;

  ht.lap("finish: output buffers");
  cudaEvent_t e0 = h->ev_a, e1 = h->ev_b;
  CK(cudaEventRecord(e0, h->s_main));
  // Round 0 runs the device-built items. Items that report trouble come back to the host: a full key store is re-run as
  // 2^n hash sub-passes, a wrapped u16 counter is re-run with u32 counters.
  std::vector<WorkItem> items, wide_items;  // host copies exist only for retry rounds
  std::vector<ItemStatus> status;
  bool wide = false, first = true;
  u64 retries_seen = 0;
  int round = 0;
  DevBuf d_ritems, d_rstatus, d_retry;
  while (first || !items.empty() || !wide_items.empty()) {
    if (!first && items.empty()) {  // narrow rounds are done: re-run the items whose u16 counters wrapped with u32 counters
      items.swap(wide_items);
      wide = true;
    }
    if (++round > 64) return fail(h, KC_ERR_CUDA, "count stage does not converge");
    u32 tab_size, key_cap;
    table_geometry(h, wide, any_ctg, tab_size, key_cap);
    const int VW = (wide ? VR_WIDE : VR_NARROW) + (any_ctg ? V_CTG : 0);
    const size_t smem = count_smem_bytes(KW, VW, h->geom.rw, tab_size, key_cap) + 8;
    size_t nitems;
    if (first) {
      ca.items = (const WorkItem *)d_items.p;
      ca.status = (ItemStatus *)d_status.p;
      ca.heavy = (const u64 *)d_heavy.p;
      ca.n_heavy = n_heavy;
      nitems = nblk ? NQ : 0;
    } else {
      std::stable_sort(items.begin(), items.end(), [](const WorkItem &x, const WorkItem &y) { return x.nsup > y.nsup; });
      nitems = items.size();
      RC(ensure(h, d_ritems, nitems * sizeof(WorkItem)));
      RC(ensure(h, d_rstatus, nitems * sizeof(ItemStatus)));
      CK(cudaMemcpyAsync(d_ritems.p, items.data(), nitems * sizeof(WorkItem), cudaMemcpyHostToDevice, h->s_main));
      ca.items = (const WorkItem *)d_ritems.p;
      ca.status = (ItemStatus *)d_rstatus.p;
      ca.heavy = nullptr;
      ca.n_heavy = 0;
    }
    ca.nitems = (u32)nitems;
    ca.tab_size = tab_size;
    ca.key_cap = key_cap;
    RC(ensure(h, d_retry, std::max<size_t>(nitems, 1) * sizeof(RetryRec)));
    ca.retry = (RetryRec *)d_retry.p;
    ca.retry_base = retries_seen;
    if (nitems) {
      // With a host sink the device-built round runs as a few slices of the partition range: the survivors of slice g travel
      // to the host while slice g+1 is being counted (the heavy list belongs to slice 0).
      const u32 G = (first && sink && nitems >= 16u * (size_t)h->sms) ? SINK_SLICES : 1u;
      while (h->ev_chunk.size() < G) {
        cudaEvent_t e;
        CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->ev_chunk.push_back(e);
      }
      CK(cudaEventRecord(h->ev_c, h->s_main));
      for (u32 g = 0; g < G; g++) {
        CountArgs cg = ca;
        const size_t lo = G == 1 ? 0 : nitems * SINK_BOUND64[g] / 64, hi = G == 1 ? nitems : nitems * SINK_BOUND64[g + 1] / 64;
        cg.items = ca.items + lo;
        cg.status = ca.status + lo;
        cg.nitems = (u32)(hi - lo);
        if (g) {
          cg.heavy = nullptr;
          cg.n_heavy = 0;
        }
        const unsigned grid = (unsigned)std::min<size_t>((size_t)cg.nitems + cg.n_heavy, (size_t)h->sms * CNT_CTAS_PER_SM);
        RC(dispatch_count(h, wide, any_ctg, cg, smem, grid));
        h->st.gpu_launches++;
        h->st.n_insert_launches++;
        if (G > 1) {
          CK(cudaMemcpyAsync(ctr + CTR_WORDS * (g + 1), d_counters.p, CTR_WORDS * 8, cudaMemcpyDeviceToHost, h->s_main));
          CK(cudaEventRecord(h->ev_chunk[g], h->s_main));
        }
      }
      CK(cudaEventRecord(h->ev_d, h->s_main));
      for (u32 g = 0; G > 1 && g < G; g++) {
        CK(cudaEventSynchronize(h->ev_chunk[g]));
        RC(sink_copy(h, sink, ctr[CTR_WORDS * (g + 1) + CTR_SURV], h->ev_chunk[g]));
      }
    }
    CK(cudaMemcpyAsync(ctr, d_counters.p, CTR_WORDS * 8, cudaMemcpyDeviceToHost, h->s_main));
    CK(cudaStreamSynchronize(h->s_main));
    ht.lap("finish: count kernel");
    if (nitems) h->st.ms_insert += elapsed(h->ev_c, h->ev_d);
    h->st.table_bytes_max = std::max<u64>(h->st.table_bytes_max, smem);
    std::vector<WorkItem> again;
    if (ctr[CTR_RETRY] != retries_seen) {
      // the kernel listed the items that have to run again (descriptor + status): the host reads those few records only
      const size_t n_new = (size_t)(ctr[CTR_RETRY] - retries_seen);
      retries_seen = ctr[CTR_RETRY];
      std::vector<RetryRec> recs(n_new);
      CK(cudaMemcpy(recs.data(), d_retry.p, n_new * sizeof(RetryRec), cudaMemcpyDeviceToHost));
      for (const RetryRec &r : recs) {
        WorkItem it = r.item;
        it.heavy = 0;
        if (r.st.overflow) {
          // it filled its key store after `progress` of `nsup` supermers: split its hash range so that each child should
          // end up around 60 % full
          const double frac_done = std::max(1.0, (double)r.st.progress) / (double)it.nsup;
          u32 f = 2;
          while (f < 256 && (double)f * frac_done < 1.6) f <<= 1;
          if ((u64)it.nsub * f > (1u << 20)) return fail(h, KC_ERR_CUDA, "partition %u cannot be split further", it.part);
          for (u32 c = 0; c < f; c++) {
            WorkItem ch = it;
            ch.sub = it.sub + c * it.nsub;
            ch.nsub = it.nsub * f;
            again.push_back(ch);
          }
          h->st.table_retries++;
        } else if (r.st.wide) {
          wide_items.push_back(it);
          h->st.table_retries++;
        }
      }
    }
    items.swap(again);
    first = false;
  }
  CK(cudaEventRecord(e1, h->s_main));
  CK(cudaStreamSynchronize(h->s_main));
  if (sink && !ctr[CTR_ERR]) {  // whatever the slices have not sent yet (retry rounds, unsliced runs)
    RC(sink_copy(h, sink, ctr[CTR_SURV], e1));
    CK(cudaStreamSynchronize(h->s_copy));
  }
  h->st.ms_finish += elapsed(e0, e1);
  for (DevBuf *b : {&d_counters, &d_segs, &d_items, &d_status, &d_descs, &d_heavy, &d_ritems, &d_rstatus, &d_retry}) release(*b);
  if (ctr[CTR_ERR]) return fail(h, KC_ERR_OVERFLOW, "survivor output capacity exceeded");
  h->n_surv = ctr[CTR_SURV];
  h->n_pf = ctr[CTR_PF];
  h->st.distinct = ctr[CTR_DISTINCT];
  h->st.survivors = ctr[CTR_SURV];
  h->st.purged = ctr[CTR_DISTINCT] - ctr[CTR_SURV];
  h->st.max_probe = ctr[CTR_MAXPROBE];
  u64 inst_parsed = 0;
  CK(cudaMemcpy(&inst_parsed, h->d_misc.p, 8, cudaMemcpyDeviceToHost));
  h->st.instances_parsed = inst_parsed;
  // supermer buffers are no longer needed (KmerDHT::clear_stores, kmer_dht.cpp:156-159)
  for (auto &b : h->blocks) { dev_free(b.rec); dev_free(b.d_hist); dev_free(b.d_base); }
  for (auto &b : h->rblocks) { dev_free(b.rec); dev_free(b.d_hist); dev_free(b.d_base); }
  h->blocks.clear();
  h->rblocks.clear();
  h->finished = true;
  ht.lap("finish: epilogue");
  return KC_OK;
}

extern "C" int kc_num_survivors(kc_handle *h, uint64_t *n) {
  if (!h || !n) return KC_ERR_INVALID;
  if (!h->finished) return fail(h, KC_ERR_STATE, "not finished");
  *n = h->n_surv;
  return KC_OK;
}

extern "C" int kc_copy_survivors(kc_handle *h, uint64_t *kmer_words, uint16_t *counts, char *left, char *right) {
  if (!h) return KC_ERR_INVALID;
  if (!h->finished) return fail(h, KC_ERR_STATE, "not finished");
  CK(cudaSetDevice(h->cfg.device));
  const u64 n = h->n_surv;
  if (!n) return KC_OK;
  if (kmer_words) CK(cudaMemcpyAsync(kmer_words, h->out_keys.p, n * h->KW * 8, cudaMemcpyDeviceToHost, h->s_main));
  if (counts) CK(cudaMemcpyAsync(counts, h->out_counts.p, n * 2, cudaMemcpyDeviceToHost, h->s_main));
  if (left) CK(cudaMemcpyAsync(left, h->out_left.p, n, cudaMemcpyDeviceToHost, h->s_main));
  if (right) CK(cudaMemcpyAsync(right, h->out_right.p, n, cudaMemcpyDeviceToHost, h->s_main));
  CK(cudaStreamSynchronize(h->s_main));
  return KC_OK;
}

// KmerDHT::dump_kmers (kmer_dht.cpp:238-266): one line per surviving k-mer, "KMER count left right". The reference writes
// kmers-<k>.txt.gz per rank through zstr; here the caller names the file, ".gz" selects gzip (libz is loaded on demand), and
// the order of lines is unspecified, as it is in the reference (hash-map iteration order).
extern "C" int kc_dump_kmers(kc_handle *h, const char *path) {
  if (!h || !path) return KC_ERR_INVALID;
  if (!h->finished) return fail(h, KC_ERR_STATE, "not finished");
  CK(cudaSetDevice(h->cfg.device));
  const u64 n = h->n_surv;
  const int KW = h->KW, k = h->cfg.kmer_len;
  const size_t plen = strlen(path);
  const bool gz = plen > 3 && strcmp(path + plen - 3, ".gz") == 0;
  void *zlib = nullptr;
  void *(*zopen)(const char *, const char *) = nullptr;
  int (*zwrite)(void *, const void *, unsigned) = nullptr;
  int (*zclose)(void *) = nullptr;
  void *zf = nullptr;
  FILE *f = nullptr;
  if (gz) {
    zlib = dlopen("libz.so.1", RTLD_NOW);
    if (!zlib) zlib = dlopen("libz.so", RTLD_NOW);
    if (!zlib) return fail(h, KC_ERR_INVALID, "%s: libz is not available, use a name without .gz", path);
    *(void **)(&zopen) = dlsym(zlib, "gzopen");
    *(void **)(&zwrite) = dlsym(zlib, "gzwrite");
    *(void **)(&zclose) = dlsym(zlib, "gzclose");
    if (!zopen || !zwrite || !zclose) return fail(h, KC_ERR_INVALID, "libz lacks gzopen/gzwrite/gzclose");
    zf = zopen(path, "wb1");
    if (!zf) return fail(h, KC_ERR_INVALID, "cannot open %s", path);
  } else {
    f = fopen(path, "wb");
    if (!f) return fail(h, KC_ERR_INVALID, "cannot open %s", path);
  }
  const u64 CH = 1u << 20;  // entries per round trip
  std::vector<u64> keys(std::min(n, CH) * KW);
  std::vector<uint16_t> counts(std::min(n, CH));
  std::vector<char> left(std::min(n, CH)), right(std::min(n, CH));
  std::string buf;
  int rc = KC_OK;
  for (u64 c0 = 0; c0 < n && rc == KC_OK; c0 += CH) {
    const u64 m = std::min(CH, n - c0);
    CK(cudaMemcpy(keys.data(), (const u64 *)h->out_keys.p + c0 * KW, m * KW * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(counts.data(), (const uint16_t *)h->out_counts.p + c0, m * 2, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(left.data(), (const char *)h->out_left.p + c0, m, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(right.data(), (const char *)h->out_right.p + c0, m, cudaMemcpyDeviceToHost));
    buf.clear();
    buf.reserve(m * (k + 12));
    char num[16];
    for (u64 i = 0; i < m; i++) {
      for (int b = 0; b < k; b++) buf.push_back("ACGT"[(keys[i * KW + (b >> 5)] >> (2 * (31 - (b & 31)))) & 3]);  // kmer.cpp:322-343
      const int nl = snprintf(num, sizeof(num), " %u ", (unsigned)counts[i]);
      buf.append(num, nl);
      buf.push_back(left[i]);
      buf.push_back(' ');
      buf.push_back(right[i]);
      buf.push_back('\n');
    }
    if (gz) {
      for (size_t o = 0; o < buf.size() && rc == KC_OK; o += (1u << 30)) {
        const unsigned len = (unsigned)std::min<size_t>(buf.size() - o, 1u << 30);
        if (zwrite(zf, buf.data() + o, len) != (int)len) rc = fail(h, KC_ERR_INVALID, "write to %s failed", path);
      }
    } else if (fwrite(buf.data(), 1, buf.size(), f) != buf.size()) {
      rc = fail(h, KC_ERR_INVALID, "write to %s failed", path);
    }
  }
  if (gz) {
    if (zclose(zf) != 0 && rc == KC_OK) rc = fail(h, KC_ERR_INVALID, "closing %s failed", path);
  } else if (fclose(f) != 0 && rc == KC_OK) {
    rc = fail(h, KC_ERR_INVALID, "closing %s failed", path);
  }
  return rc;
}

namespace {
// kc_table_digest: order-independent digest of the surviving table (include/kcount_b200.h)
__global__ void __launch_bounds__(256) digest_kernel(const u64 *keys, const uint16_t *counts, const char *left, const char *right, u64 n,
                                                     int KW, u64 *out4) {
  u64 sc = 0, sm = 0, xm = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    u64 x = 0x9E3779B97F4A7C15ULL;
    for (int j = 0; j < KW; j++) x = fmix64(x ^ keys[i * KW + j]);
    const u64 c = counts[i];
    x = fmix64(x ^ (c | ((u64)(unsigned char)left[i] << 16) | ((u64)(unsigned char)right[i] << 24)));
    sc += c;
    sm += x;
    xm ^= x;
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) {
    sc += __shfl_xor_sync(0xffffffffu, sc, d);
    sm += __shfl_xor_sync(0xffffffffu, sm, d);
    xm ^= __shfl_xor_sync(0xffffffffu, xm, d);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd((unsigned long long *)&out4[1], sc);
    atomicAdd((unsigned long long *)&out4[2], sm);
    atomicXor((unsigned long long *)&out4[3], xm);
  }
}
}  // namespace

extern "C" int kc_table_digest(kc_handle *h, uint64_t out4[4]) {
  if (!h || !out4) return KC_ERR_INVALID;
  if (!h->finished) return fail(h, KC_ERR_STATE, "not finished");
  CK(cudaSetDevice(h->cfg.device));
  DevBuf d;
  RC(ensure(h, d, 32));
  CK(cudaMemsetAsync(d.p, 0, 32, h->s_main));
  const u64 n = h->n_surv;
  if (n) {
    const unsigned grid = (unsigned)std::min<u64>((n + 255) / 256, (u64)h->sms * 8);
    digest_kernel<<<grid, 256, 0, h->s_main>>>((const u64 *)h->out_keys.p, (const uint16_t *)h->out_counts.p, (const char *)h->out_left.p,
                                               (const char *)h->out_right.p, n, h->KW, (u64 *)d.p);
    h->st.gpu_launches++;
    CK(cudaGetLastError());
  }
  CK(cudaMemcpyAsync(out4, d.p, 32, cudaMemcpyDeviceToHost, h->s_main));
  CK(cudaStreamSynchronize(h->s_main));
  out4[0] = n;
  release(d);
  return KC_OK;
}

extern "C" int kc_device_count(int32_t *n) {
  if (!n) return KC_ERR_INVALID;
  int c = 0;
  if (cudaGetDeviceCount(&c) != cudaSuccess) return KC_ERR_CUDA;
  *n = c;
  return KC_OK;
}

extern "C" int kc_debug_set_key_cap(kc_handle *h, int32_t key_cap) {
  if (!h || key_cap < 0) return KC_ERR_INVALID;
  if (h->P) return fail(h, KC_ERR_STATE, "kc_debug_set_key_cap after the first add");
  h->dbg_key_cap = key_cap;
  return KC_OK;
}

extern "C" int kc_survivors_device(kc_handle *h, const uint64_t **d_kmer_words, const uint16_t **d_counts, const char **d_left,
                                   const char **d_right) {
  if (!h) return KC_ERR_INVALID;
  if (!h->finished) return fail(h, KC_ERR_STATE, "not finished");
  if (d_kmer_words) *d_kmer_words = (const uint64_t *)h->out_keys.p;
  if (d_counts) *d_counts = (const uint16_t *)h->out_counts.p;
  if (d_left) *d_left = (const char *)h->out_left.p;
  if (d_right) *d_right = (const char *)h->out_right.p;
  return KC_OK;
}

extern "C" int kc_num_prefilter(kc_handle *h, uint64_t *n) {
  if (!h || !n) return KC_ERR_INVALID;
  if (!h->finished) return fail(h, KC_ERR_STATE, "not finished");
  if (!h->cfg.keep_prefilter) return fail(h, KC_ERR_STATE, "keep_prefilter was not set");
  *n = h->n_pf;
  return KC_OK;
}

extern "C" int kc_copy_prefilter(kc_handle *h, uint64_t *kmer_words, uint16_t *vals9) {
  if (!h) return KC_ERR_INVALID;
  if (!h->finished) return fail(h, KC_ERR_STATE, "not finished");
  if (!h->cfg.keep_prefilter) return fail(h, KC_ERR_STATE, "keep_prefilter was not set");
  CK(cudaSetDevice(h->cfg.device));
  const u64 n = h->n_pf;
  if (!n) return KC_OK;
  CK(cudaMemcpy(kmer_words, h->pf_keys.p, n * h->KW * 8, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(vals9, h->pf_vals.p, n * 18, cudaMemcpyDeviceToHost));
  return KC_OK;
}

extern "C" int kc_get_stats(kc_handle *h, kc_stats *st) {
  if (!h || !st) return KC_ERR_INVALID;
  h->st.ms_total = h->st.ms_parse + h->st.ms_exchange + h->st.ms_finish;
  *st = h->st;
  return KC_OK;
}

extern "C" int kc_release_cached_memory(void) {
  DevCache &c = dev_cache();
  std::vector<std::pair<int, void *>> drop;
  {
    std::lock_guard<std::mutex> lk(c.mu);
    for (auto &d : c.free_blocks) {
      for (auto &kv : d.second) drop.push_back({d.first, kv.second});
      d.second.clear();
    }
  }
  int cur = 0;
  cudaGetDevice(&cur);
  for (auto &x : drop) {
    cudaSetDevice(x.first);
    cudaFree(x.second);
  }
  cudaSetDevice(cur);
  return KC_OK;
}

extern "C" int kc_destroy(kc_handle *h) {
  if (!h) return KC_OK;
  cudaSetDevice(h->cfg.device);
  if (h->s_main) cudaStreamSynchronize(h->s_main);
  if (h->s_copy) cudaStreamSynchronize(h->s_copy);
  if (h->s_comm) cudaStreamSynchronize(h->s_comm);
  if (h->s_meta) cudaStreamSynchronize(h->s_meta);
  for (auto &b : h->blocks) { dev_free(b.rec); dev_free(b.d_hist); dev_free(b.d_base); }
  for (auto &b : h->rblocks) { dev_free(b.rec); dev_free(b.d_hist); dev_free(b.d_base); }
  for (ParseSlot &sl : h->slot) {
    if (sl.st) cudaStreamSynchronize(sl.st);
    if (sl.busy) { dev_free(sl.blk.rec); dev_free(sl.blk.d_hist); dev_free(sl.blk.d_base); }
    for (DevBuf &b : sl.temps) release(b);
    for (DevBuf *b : {&sl.d_cursor, &sl.d_tot, &sl.d_scan, &sl.d_elist, &sl.d_ecnt}) release(*b);
    for (cudaEvent_t e : {sl.e0, sl.ec, sl.ed, sl.e1})
      if (e) cudaEventDestroy(e);
    if (sl.st) cudaStreamDestroy(sl.st);
    pin_put(sl.h_tot, sl.h_tot_bytes);
  }
  if (h->ev_first) cudaEventDestroy(h->ev_first);
  DevBuf *bufs[] = {&h->stage_bytes[0], &h->stage_bytes[1], &h->stage_offs[0], &h->stage_offs[1], &h->d_meta, &h->d_meta_all,
                    &h->d_tot, &h->d_scan, &h->d_scan_comm, &h->d_misc, &h->out_keys, &h->out_counts, &h->out_left, &h->out_right,
                    &h->pf_keys, &h->pf_vals};
  for (DevBuf *b : bufs) release(*b);
  if (h->ev_a) cudaEventDestroy(h->ev_a);
  if (h->ev_b) cudaEventDestroy(h->ev_b);
  if (h->ev_c) cudaEventDestroy(h->ev_c);
  if (h->ev_d) cudaEventDestroy(h->ev_d);
  for (int i = 0; i < 2; i++) {
    if (h->ev_copy[i]) cudaEventDestroy(h->ev_copy[i]);
    if (h->ev_free[i]) cudaEventDestroy(h->ev_free[i]);
  }
  if (h->s_main) cudaStreamDestroy(h->s_main);
  if (h->s_copy) cudaStreamDestroy(h->s_copy);
  if (h->s_comm) cudaStreamDestroy(h->s_comm);
  if (h->s_meta) cudaStreamDestroy(h->s_meta);
  pin_put(h->h_pin, h->pin_bytes);
  for (auto &p : h->pending) { dev_free(p.dir); dev_free(p.rec); }
  if (h->ev_fill) cudaEventDestroy(h->ev_fill);
  for (auto e : h->ev_chunk) cudaEventDestroy(e);
  if (h->ev_x0) cudaEventDestroy(h->ev_x0);
  if (h->ev_x1) cudaEventDestroy(h->ev_x1);
  for (auto &pr : h->comm_ev) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  void kc_comm_destroy(kc_handle *);
  kc_comm_destroy(h);
  delete h;
  return KC_OK;
}

// ---------------------------------------------------------------- primitives for known-answer tests ----------------

namespace {

template <int KW>
__global__ void prim_kmers_kernel(const char *kmers, u64 n, int k, int m, u32 num_shards, u64 *words, u64 *rcw, u64 *minim, u64 *hash,
                                  int32_t *owner) {
  const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const char *s = kmers + i * k;
  // pack into the u32 stream layout the kernels use (16 bases per word, MSB first), then fetch like they do
  u32 sb[KW * 2 + 3];
  for (int j = 0; j < KW * 2 + 3; j++) sb[j] = 0;
  for (int j = 0; j < k; j++) {
    const u32 b = (u32)s[j];
    const u32 x = (b & 4) >> 1;
    const u32 code = x + ((x ^ (b & 2)) >> 1);
    sb[j >> 4] |= code << (30 - 2 * (j & 15));
  }
  const int nw = (k + 31) >> 5;
  const u64 last_mask = (k & 31) ? (~0ULL << (64 - 2 * (k & 31))) : ~0ULL;
  u64 fwd[KW], rc[KW];
  for (int j = 0; j < KW; j++) fwd[j] = (j < nw) ? fetch64(sb, 32 * j) : 0ULL;
  for (int j = 0; j < KW; j++)
    if (j == nw - 1) fwd[j] &= last_mask;
  revcomp_words<KW>(fwd, rc, k);
  u64 best = 0;
  for (int p = 0; p + m <= k; p++) {
    const u64 c = canonical_mmer(fetch64(sb, p), m);
    best = c > best ? c : best;
  }
  for (int j = 0; j < KW; j++) {
    words[i * KW + j] = fwd[j];
    rcw[i * KW + j] = rc[j];
  }
  minim[i] = best;
  hash[i] = murmur64_words<KW>(fwd);
  owner[i] = (int32_t)(quick_hash(best) % num_shards);
}

__global__ void prim_get_ext_kernel(const uint16_t *rows, u64 n, int dmin_thres, double frac, char *out) {
  const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  u32 e[4] = {rows[i * 5], rows[i * 5 + 1], rows[i * 5 + 2], rows[i * 5 + 3]};
  out[i] = get_ext(e, rows[i * 5 + 4], dmin_thres, frac);
}

}  // namespace

#define CKP(call)                                   \
  do {                                              \
    cudaError_t e_ = (call);                        \
    if (e_ != cudaSuccess) { rc = KC_ERR_CUDA; goto done; } \
  } while (0)

extern "C" int kc_prim_kmers(int32_t device, int32_t k, const char *kmers, uint64_t n, int32_t num_shards, uint64_t *words,
                             uint64_t *rc_words, uint64_t *minimizer, uint64_t *hash, int32_t *owner) {
  if (k < 2 || k > 127 || !n) return KC_ERR_INVALID;
  const int KW = k / 32 + 1;
  int m = k * 2 / 3 + 1;
  if (m < 15) m = 15;
  if (m > 27) m = 27;
  if (m > k) m = k;
  int rc = KC_OK;
  char *d_k = nullptr;
  u64 *d_w = nullptr, *d_r = nullptr, *d_m = nullptr, *d_h = nullptr;
  int32_t *d_o = nullptr;
  const unsigned grid = (unsigned)((n + 127) / 128);
  CKP(cudaSetDevice(device));
  CKP(cudaMalloc((void **)&d_k, n * k));
  CKP(cudaMalloc((void **)&d_w, n * KW * 8));
  CKP(cudaMalloc((void **)&d_r, n * KW * 8));
  CKP(cudaMalloc((void **)&d_m, n * 8));
  CKP(cudaMalloc((void **)&d_h, n * 8));
  CKP(cudaMalloc((void **)&d_o, n * 4));
  CKP(cudaMemcpy(d_k, kmers, n * k, cudaMemcpyHostToDevice));
  switch (KW) {
    case 1: prim_kmers_kernel<1><<<grid, 128>>>(d_k, n, k, m, (u32)num_shards, d_w, d_r, d_m, d_h, d_o); break;
    case 2: prim_kmers_kernel<2><<<grid, 128>>>(d_k, n, k, m, (u32)num_shards, d_w, d_r, d_m, d_h, d_o); break;
    case 3: prim_kmers_kernel<3><<<grid, 128>>>(d_k, n, k, m, (u32)num_shards, d_w, d_r, d_m, d_h, d_o); break;
    default: prim_kmers_kernel<4><<<grid, 128>>>(d_k, n, k, m, (u32)num_shards, d_w, d_r, d_m, d_h, d_o); break;
  }
  CKP(cudaGetLastError());
  CKP(cudaMemcpy(words, d_w, n * KW * 8, cudaMemcpyDeviceToHost));
  CKP(cudaMemcpy(rc_words, d_r, n * KW * 8, cudaMemcpyDeviceToHost));
  CKP(cudaMemcpy(minimizer, d_m, n * 8, cudaMemcpyDeviceToHost));
  CKP(cudaMemcpy(hash, d_h, n * 8, cudaMemcpyDeviceToHost));
  CKP(cudaMemcpy(owner, d_o, n * 4, cudaMemcpyDeviceToHost));
done:
  cudaFree(d_k); cudaFree(d_w); cudaFree(d_r); cudaFree(d_m); cudaFree(d_h); cudaFree(d_o);
  return rc;
}

extern "C" int kc_prim_get_ext(int32_t device, const uint16_t *acgt_count5, uint64_t n, int32_t dmin_thres, double dyn_min_depth,
                               char *out) {
  if (!n) return KC_ERR_INVALID;
  int rc = KC_OK;
  uint16_t *d_in = nullptr;
  char *d_out = nullptr;
  CKP(cudaSetDevice(device));
  CKP(cudaMalloc((void **)&d_in, n * 10));
  CKP(cudaMalloc((void **)&d_out, n));
  CKP(cudaMemcpy(d_in, acgt_count5, n * 10, cudaMemcpyHostToDevice));
  prim_get_ext_kernel<<<(unsigned)((n + 127) / 128), 128>>>(d_in, n, dmin_thres, 1.0 - dyn_min_depth, d_out);
  CKP(cudaGetLastError());
  CKP(cudaMemcpy(out, d_out, n, cudaMemcpyDeviceToHost));
done:
  cudaFree(d_in); cudaFree(d_out);
  return rc;
}

// ---------------------------------------------------------------- allocation-free digest ----------
// kc_table_digest_static: the digest of kc_table_digest for callers that cannot allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc). No counterpart in the reference. Its scratch — one partial result per CTA and a
// ticket — lives in a file-scope __device__ array that every GPU loading this library holds until exit
// (KC_DIGEST_STATIC_BLOCKS * 32 + 256 bytes = 128.25 KB). In host code the array's NAME is only a host-side stand-in: the
// device address comes from cudaGetSymbolAddress, looked up once per device with that device current.
namespace {
constexpr u32 KC_DIGEST_STATIC_BLOCKS = 4096;        // partial results the scratch array holds
constexpr u32 KC_DIGEST_STATIC_ROWS_PER_BLOCK = 65536;
static_assert((u64)KC_DIGEST_STATIC_BLOCKS * KC_DIGEST_STATIC_ROWS_PER_BLOCK == KC_DIGEST_STATIC_MAX_ROWS, "header and scratch size agree");
// [0, 32): ticket (must be zero when a launch starts) + padding; then 4 words per CTA
static __device__ __align__(256) unsigned long long kc_digest_scratch[32 + 4 * KC_DIGEST_STATIC_BLOCKS];

__global__ void __launch_bounds__(256) digest_static_kernel(const u64 *keys, const uint16_t *counts, const char *left, const char *right,
                                                            u64 n, int KW, unsigned long long *scratch, u64 *out4) {
  __shared__ u64 sh[3][8];
  __shared__ bool s_last;
  const u64 r0 = (u64)blockIdx.x * KC_DIGEST_STATIC_ROWS_PER_BLOCK;
  const u64 r1 = r0 + KC_DIGEST_STATIC_ROWS_PER_BLOCK < n ? r0 + KC_DIGEST_STATIC_ROWS_PER_BLOCK : n;
  u64 sc = 0, sm = 0, xm = 0;
  for (u64 i = r0 + threadIdx.x; i < r1; i += blockDim.x) {
    u64 x = 0x9E3779B97F4A7C15ULL;
    for (int j = 0; j < KW; j++) x = fmix64(x ^ keys[i * KW + j]);
    const u64 c = counts[i];
    x = fmix64(x ^ (c | ((u64)(unsigned char)left[i] << 16) | ((u64)(unsigned char)right[i] << 24)));
    sc += c;
    sm += x;
    xm ^= x;
  }
  auto block_reduce = [&](u64 &a, u64 &b, u64 &c) {
#pragma unroll
    for (int d = 16; d; d >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, d);
      b += __shfl_xor_sync(0xffffffffu, b, d);
      c ^= __shfl_xor_sync(0xffffffffu, c, d);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) { sh[0][wid] = a; sh[1][wid] = b; sh[2][wid] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < 8; w++) { a += sh[0][w]; b += sh[1][w]; c ^= sh[2][w]; }
    }
  };
  block_reduce(sc, sm, xm);
  unsigned long long *part = scratch + 32;
  if (threadIdx.x == 0) {
    part[4 * blockIdx.x + 1] = sc;
    part[4 * blockIdx.x + 2] = sm;
    part[4 * blockIdx.x + 3] = xm;
    __threadfence();
    s_last = atomicAdd(&scratch[0], 1ull) == (unsigned long long)gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  sc = sm = xm = 0;
  for (u32 b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
    sc += ((volatile unsigned long long *)part)[4 * b + 1];
    sm += ((volatile unsigned long long *)part)[4 * b + 2];
    xm ^= ((volatile unsigned long long *)part)[4 * b + 3];
  }
  block_reduce(sc, sm, xm);
  if (threadIdx.x == 0) {
    out4[0] = n;
    out4[1] = sc;
    out4[2] = sm;
    out4[3] = xm;
    scratch[0] = 0;  // leave the ticket as the next call needs it (the call clears it as well)
  }
}

constexpr int KC_MAX_DEVICES = 64;
std::mutex g_digest_static_mu;
unsigned long long *g_digest_static_ptr[KC_MAX_DEVICES];  // device address of kc_digest_scratch on each device
}  // namespace

extern "C" int kc_table_digest_static(int32_t device, const uint64_t *d_kmer_words, const uint16_t *d_counts, const char *d_left,
                                      const char *d_right, uint64_t n, int32_t n_longs, uint64_t *d_out4, void *cuda_stream,
                                      char *errbuf, uint64_t errbuf_len) {
  auto say = [&](int code, const char *fmt, ...) {
    if (errbuf && errbuf_len) {
      va_list ap;
      va_start(ap, fmt);
      vsnprintf(errbuf, (size_t)errbuf_len, fmt, ap);
      va_end(ap);
    }
    return code;
  };
  if (errbuf && errbuf_len) errbuf[0] = 0;
  if (device < 0 || device >= KC_MAX_DEVICES) return say(KC_ERR_INVALID, "device %d outside 0..%d", device, KC_MAX_DEVICES - 1);
  if (n_longs < 1 || n_longs > 4) return say(KC_ERR_INVALID, "n_longs %d outside 1..4", n_longs);
  if (!d_out4 || (n && (!d_kmer_words || !d_counts || !d_left || !d_right))) return say(KC_ERR_INVALID, "null pointer");
  if (n > KC_DIGEST_STATIC_MAX_ROWS)
    return say(KC_ERR_INVALID, "%llu rows exceed the static scratch's limit of %llu rows (KC_DIGEST_STATIC_MAX_ROWS): use kc_table_digest",
               (unsigned long long)n, (unsigned long long)KC_DIGEST_STATIC_MAX_ROWS);
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return say(KC_ERR_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  unsigned long long *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_digest_static_mu);
    if (!g_digest_static_ptr[device]) {  // set-up, once per device, with that device current
      void *p = nullptr;
      e = cudaGetSymbolAddress(&p, kc_digest_scratch);
      if (e != cudaSuccess || !p) return say(KC_ERR_CUDA, "cudaGetSymbolAddress(kc_digest_scratch): %s", cudaGetErrorString(e));
      g_digest_static_ptr[device] = (unsigned long long *)p;
    }
    scratch = g_digest_static_ptr[device];
  }
  cudaStream_t st = (cudaStream_t)cuda_stream;
  // the array keeps the last call's contents: the ticket must start at zero (partials are overwritten before they are read)
  e = cudaMemsetAsync(scratch, 0, 256, st);
  if (e != cudaSuccess) return say(KC_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
  const unsigned grid = (unsigned)std::max<u64>(1, (n + KC_DIGEST_STATIC_ROWS_PER_BLOCK - 1) / KC_DIGEST_STATIC_ROWS_PER_BLOCK);
  digest_static_kernel<<<grid, 256, 0, st>>>((const u64 *)d_kmer_words, d_counts, d_left, d_right, n, n_longs, scratch, (u64 *)d_out4);
  e = cudaGetLastError();
  if (e != cudaSuccess) return say(KC_ERR_CUDA, "digest_static_kernel launch: %s", cudaGetErrorString(e));
  return KC_OK;
}

#include "kc_comm.inl"

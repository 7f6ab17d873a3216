// Multi-shard supermer exchange. Included at the end of kc_api.cu.
//
// Replaces the reference's ThreeTierAggrStore<Supermer> / rpc_ff data plane (kmer_dht.cpp:133,146-149,222-231;
// upcxx-utils three_tier_aggr_store.hpp:914-960): every shard has parsed its reads into buckets ordered
// (owner shard, partition), so the data for one destination is ONE contiguous header range and ONE contiguous payload
// range per parsed batch. The exchange is a personalised all-to-all of those ranges plus a P-entry directory each:
//   * NCCL transport (one process per GPU): ncclSend/ncclRecv inside one group over NVLink/NVSwitch, sizes agreed by a
//     small ncclAllGather first. libnccl is dlopen'ed so that single-GPU use has no NCCL dependency.
//   * local transport (several shards inside one process, e.g. tests on one GPU or one process driving a box):
//     cudaMemcpyPeerAsync between the handles' devices.
// Both end in the same place: h->rblocks, which the count phase treats exactly like locally parsed blocks.
#if __has_include(<nccl.h>)
#include <nccl.h>
#define KC_HAVE_NCCL_H 1
#else
#define KC_HAVE_NCCL_H 0
#endif
#include <map>
#include <mutex>

namespace {

// a received run: directory (P entries) already on the host, hdr/payload device buffers owned by the block
int make_rblock(kc_handle *h, const BucketHist *dir, u32 *rec, bool is_ctg) {
  const u32 P = h->P;
  Block b;
  b.received = true;
  b.is_ctg = is_ctg;
  b.rec = rec;
  b.hist.assign(dir, dir + P);
  b.rec_base.resize(P + 1);
  u64 hs = 0;
  for (u32 p = 0; p < P; p++) {
    b.rec_base[p] = hs;
    hs += dir[p].nsup;
  }
  b.rec_base[P] = hs;
  b.tot_sup = hs;
  h->rblocks.push_back(std::move(b));
  return KC_OK;
}

#if KC_HAVE_NCCL_H
struct NcclApi {
  void *lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

NcclApi &nccl_api() {
  static NcclApi a;
  static std::once_flag once;
  std::call_once(once, [] {
    // if the process already carries an NCCL (e.g. torch's bundled one) dlopen returns that same library (SONAME match)
    a.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!a.lib) a.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!a.lib) return;
#define SYM(field, name) *(void **)(&a.field) = dlsym(a.lib, name)
    SYM(GetUniqueId, "ncclGetUniqueId");
    SYM(CommInitRank, "ncclCommInitRank");
    SYM(CommDestroy, "ncclCommDestroy");
    SYM(Send, "ncclSend");
    SYM(Recv, "ncclRecv");
    SYM(AllGather, "ncclAllGather");
    SYM(GroupStart, "ncclGroupStart");
    SYM(GroupEnd, "ncclGroupEnd");
    SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    a.ok = a.GetUniqueId && a.CommInitRank && a.Send && a.Recv && a.AllGather && a.GroupStart && a.GroupEnd;
  });
  return a;
}

struct Nccl {
  ncclComm_t comm = nullptr;
  int nranks = 0, rank = 0;
};

// communicators are cached per (unique id, rank): creating one costs hundreds of ms and a job runs one kcount round per k
std::map<std::string, Nccl *> &comm_cache() {
  static std::map<std::string, Nccl *> m;
  return m;
}

#define NC(call)                                                                                              \
  do {                                                                                                        \
    ncclResult_t r_ = (call);                                                                                 \
    if (r_ != ncclSuccess)                                                                                    \
      return fail(h, KC_ERR_COMM, "%s:%d %s: %s", __FILE__, __LINE__, #call,                                  \
                  nccl_api().GetErrorString ? nccl_api().GetErrorString(r_) : "nccl error");                  \
  } while (0)
#else
struct Nccl {
  int nranks = 0, rank = 0;
};
#endif

}  // namespace

extern "C" int kc_comm_unique_id(uint8_t id128[128]) {
#if KC_HAVE_NCCL_H
  NcclApi &a = nccl_api();
  if (!a.ok) return KC_ERR_COMM;
  ncclUniqueId id;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  if (a.GetUniqueId(&id) != ncclSuccess) return KC_ERR_COMM;
  memcpy(id128, &id, 128);
  return KC_OK;
#else
  (void)id128;
  return KC_ERR_COMM;
#endif
}

extern "C" int kc_comm_init(kc_handle *h, const uint8_t id128[128]) {
  if (!h || !id128) return KC_ERR_INVALID;
#if KC_HAVE_NCCL_H
  NcclApi &a = nccl_api();
  if (!a.ok) return fail(h, KC_ERR_COMM, "libnccl.so.2 could not be loaded");
  if (h->cfg.num_shards < 2) return KC_OK;
  CK(cudaSetDevice(h->cfg.device));
  std::string key((const char *)id128, 128);
  key += ":" + std::to_string(h->cfg.shard_id) + "/" + std::to_string(h->cfg.num_shards);
  auto it = comm_cache().find(key);
  if (it == comm_cache().end()) {
    Nccl *c = new Nccl();
    c->nranks = h->cfg.num_shards;
    c->rank = h->cfg.shard_id;
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    NC(a.CommInitRank(&c->comm, c->nranks, id, c->rank));
    it = comm_cache().emplace(key, c).first;
  }
  h->comm = it->second;
  return KC_OK;
#else
  return fail(h, KC_ERR_COMM, "built without nccl.h");
#endif
}

void kc_comm_destroy(kc_handle *h) { h->comm = nullptr; /* cached communicators live until process exit */ }

// NCCL transport. Collective: every shard calls it once, from kc_finish.
int kc_exchange(kc_handle *h) {
#if KC_HAVE_NCCL_H
  if (!h->comm) return fail(h, KC_ERR_COMM, "multi-shard handle without a communicator: call kc_comm_init first");
  NcclApi &a = nccl_api();
  const int N = h->cfg.num_shards, me = h->cfg.shard_id;
  const u32 P = h->P;
  cudaStream_t st = h->s_main;
  cudaEvent_t e0 = h->ev_a, e1 = h->ev_b;
  CK(cudaEventRecord(e0, st));
  // 1. agree on (P, number of blocks) ...
  DevBuf d_meta, d_all;
  u64 mine[2] = {P, (u64)h->blocks.size()};
  RC(ensure(h, d_meta, 16));
  RC(ensure(h, d_all, 16 * (size_t)N));
  CK(cudaMemcpyAsync(d_meta.p, mine, 16, cudaMemcpyHostToDevice, st));
  NC(a.AllGather(d_meta.p, d_all.p, 2, ncclUint64, h->comm->comm, st));
  std::vector<u64> all(2 * (size_t)N);
  CK(cudaMemcpyAsync(all.data(), d_all.p, 16 * (size_t)N, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  u64 maxb = 0;
  for (int r = 0; r < N; r++) {
    if (all[2 * r] != P)
      return fail(h, KC_ERR_INVALID, "shards disagree on num_partitions (%llu vs %u): set cfg.num_partitions or expected_instances",
                  (unsigned long long)all[2 * r], P);
    maxb = std::max(maxb, all[2 * r + 1]);
  }
  if (maxb == 0) return KC_OK;
  // 2. ... then on every (block, destination) size: per block [is_ctg, nsup x N]
  const size_t rec = 1 + (size_t)N, per_rank = maxb * rec;
  const size_t RB = (size_t)h->geom.rw * 4;  // bytes per supermer record
  std::vector<u64> sz(per_rank, 0), szall(per_rank * N);
  for (size_t b = 0; b < h->blocks.size(); b++) {
    const Block &blk = h->blocks[b];
    sz[b * rec] = blk.is_ctg ? 1 : 0;
    for (int d = 0; d < N; d++) {
      sz[b * rec + 1 + d] = blk.tot_sup ? blk.rec_base[(size_t)(d + 1) * P] - blk.rec_base[(size_t)d * P] : 0;
    }
  }
  RC(ensure(h, d_meta, per_rank * 8));
  RC(ensure(h, d_all, per_rank * 8 * N));
  CK(cudaMemcpyAsync(d_meta.p, sz.data(), per_rank * 8, cudaMemcpyHostToDevice, st));
  NC(a.AllGather(d_meta.p, d_all.p, per_rank, ncclUint64, h->comm->comm, st));
  CK(cudaMemcpyAsync(szall.data(), d_all.p, per_rank * 8 * N, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  // 3. directories of my blocks on the device (P entries per destination are sent along)
  std::vector<BucketHist *> d_dirs(h->blocks.size(), nullptr);
  for (size_t b = 0; b < h->blocks.size(); b++) {
    if (!h->blocks[b].tot_sup) continue;
    CK(dev_alloc((void **)&d_dirs[b], sizeof(BucketHist) * (size_t)N * P));
    CK(cudaMemcpyAsync(d_dirs[b], h->blocks[b].hist.data(), sizeof(BucketHist) * (size_t)N * P, cudaMemcpyHostToDevice, st));
  }
  // 4. receive buffers
  struct Incoming { BucketHist *dir; u32 *rec; u64 nsup; bool ctg; int src; };
  std::vector<Incoming> inc;
  for (int s = 0; s < N; s++) {
    if (s == me) continue;
    for (u64 b = 0; b < all[2 * s + 1]; b++) {
      const u64 *r = &szall[(size_t)s * per_rank + b * rec];
      const u64 nsup = r[1 + me];
      if (!nsup) continue;
      Incoming in{nullptr, nullptr, nsup, r[0] != 0, s};
      CK(dev_alloc((void **)&in.dir, sizeof(BucketHist) * P));
      CK(dev_alloc((void **)&in.rec, RB * nsup + 64));
      inc.push_back(in);
    }
  }
  // 5. one group: all sends and receives; between a pair of ranks the order is (block asc; dir, hdr, payload)
  NC(a.GroupStart());
  u64 sent = 0;
  for (int d = 0; d < N; d++) {
    if (d == me) continue;
    for (size_t b = 0; b < h->blocks.size(); b++) {
      const Block &blk = h->blocks[b];
      const u64 nsup = sz[b * rec + 1 + d];
      if (!nsup) continue;
      NC(a.Send(d_dirs[b] + (size_t)d * P, sizeof(BucketHist) * P, ncclUint8, d, h->comm->comm, st));
      NC(a.Send(blk.rec + blk.rec_base[(size_t)d * P] * h->geom.rw, RB * nsup, ncclUint8, d, h->comm->comm, st));
      sent += sizeof(BucketHist) * P + RB * nsup;
    }
  }
  for (auto &in : inc) {
    NC(a.Recv(in.dir, sizeof(BucketHist) * P, ncclUint8, in.src, h->comm->comm, st));
    NC(a.Recv(in.rec, RB * in.nsup, ncclUint8, in.src, h->comm->comm, st));
  }
  NC(a.GroupEnd());
  // 6. directories back to the host -> rblocks
  std::vector<BucketHist> dirs(inc.size() * (size_t)P);
  for (size_t i = 0; i < inc.size(); i++)
    CK(cudaMemcpyAsync(&dirs[i * P], inc[i].dir, sizeof(BucketHist) * P, cudaMemcpyDeviceToHost, st));
  CK(cudaEventRecord(e1, st));
  CK(cudaStreamSynchronize(st));
  for (size_t i = 0; i < inc.size(); i++) {
    make_rblock(h, &dirs[i * P], inc[i].rec, inc[i].ctg);
    dev_free(inc[i].dir);
  }
  for (auto p : d_dirs)
    if (p) dev_free(p);
  release(d_meta);
  release(d_all);
  h->st.ms_exchange += elapsed(e0, e1);
  h->st.supermer_bytes_sent += sent;
  return KC_OK;
#else
  return fail(h, KC_ERR_COMM, "built without nccl.h");
#endif
}

int finish_count(kc_handle *h);  // kc_api.cu

// Local transport: all shards live in this process (handles[i] must be shard i of n). Exchange by peer copies, then count.
extern "C" int kc_finish_local_group(kc_handle **hs, int32_t n) {
  if (!hs || n < 1) return KC_ERR_INVALID;
  for (int i = 0; i < n; i++) {
    kc_handle *h = hs[i];
    if (!h) return KC_ERR_INVALID;
    if (h->finished) return fail(h, KC_ERR_STATE, "finish called twice");
    if (h->cfg.num_shards != n || h->cfg.shard_id != i) return fail(h, KC_ERR_INVALID, "handle %d is not shard %d of %d", i, i, n);
    if (!h->P) decide_partitions(h, 0);
    if (h->P != hs[0]->P) return fail(h, KC_ERR_INVALID, "shards disagree on num_partitions: set cfg.num_partitions");
  }
  const u32 P = hs[0]->P;
  for (int s = 0; s < n; s++) {
    kc_handle *h = hs[s];  // `h` is what CK reports into
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->s_main));
    for (auto &blk : h->blocks) {
      if (!blk.tot_sup) continue;
      for (int d = 0; d < n; d++) {
        if (d == s) continue;
        const u64 nsup = blk.rec_base[(size_t)(d + 1) * P] - blk.rec_base[(size_t)d * P];
        if (!nsup) continue;
        kc_handle *hd = hs[d];
        const size_t RB = (size_t)h->geom.rw * 4;
        u32 *rec = nullptr;
        CK(cudaSetDevice(hd->cfg.device));
        CK(dev_alloc((void **)&rec, RB * nsup + 64));
        CK(cudaMemcpyPeer(rec, hd->cfg.device, blk.rec + blk.rec_base[(size_t)d * P] * h->geom.rw, h->cfg.device, RB * nsup));
        make_rblock(hd, &blk.hist[(size_t)d * P], rec, blk.is_ctg);
        h->st.supermer_bytes_sent += RB * nsup;
      }
    }
  }
  for (int i = 0; i < n; i++) {
    int rc = finish_count(hs[i]);
    if (rc) return rc;
  }
  return KC_OK;
}

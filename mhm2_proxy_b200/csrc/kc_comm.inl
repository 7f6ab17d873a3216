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

// a received run: directory (P entries) and records are device buffers now owned by the block; its bucket layout is
// computed on `st` (the stream the data arrived on)
int make_rblock(kc_handle *h, BucketHist *dir, u32 *rec, u64 nsup, bool is_ctg, cudaStream_t st) {
  Block b;
  b.received = true;
  b.is_ctg = is_ctg;
  b.rec = rec;
  b.d_hist = dir;
  b.tot_sup = nsup;
  CK(dev_alloc((void **)&b.d_base, sizeof(u64) * ((size_t)h->P + 1)));
  // scratch per stream: received blocks are scanned on the communication stream (or, local transport, on the main stream)
  DevBuf &scratch = st == h->s_main ? h->d_scan : h->d_scan_comm;
  RC(ensure(h, scratch, sizeof(u64) * 1032));
  RC(bucket_scan(h, dir, h->P, b.d_base, h->P, nullptr, (u64 *)scratch.p, st));
  h->rblocks.push_back(std::move(b));
  return KC_OK;
}

#if KC_HAVE_NCCL_H
struct NcclApi {
  void *lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t *, void *) = nullptr;  // NCCL >= 2.18; optional
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
    SYM(CommSplit, "ncclCommSplit");
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
  ncclComm_t comm = nullptr;   // bulk data: the per-round groups of sends / receives
  ncclComm_t meta = nullptr;   // a duplicate for the per-round size all-gather (== comm when ncclCommSplit is unavailable)
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
    // The size all-gather of round i must not queue behind the bulk transfer of round i-1 (operations on one communicator
    // run in issue order): it gets its own communicator and its own stream, so that the host only ever waits for the peers
    // to REACH round i, not for data to move (N=2, C2: the parse stage went from 25.3 to the single-GPU 19 ms per k round).
    c->meta = c->comm;
    static const bool split = !(getenv("KC_COMM_SPLIT") && atoi(getenv("KC_COMM_SPLIT")) == 0);
    if (a.CommSplit && split) {
      ncclComm_t m = nullptr;
      if (a.CommSplit(c->comm, 0, c->rank, &m, nullptr) == ncclSuccess && m) c->meta = m;
    }
    it = comm_cache().emplace(key, c).first;
  }
  h->comm = it->second;
  // Shards must bucket with the same number of partitions. Agree once, here: the largest instance hint of any shard sizes
  // the partitions of all of them (an explicit num_partitions must already be equal everywhere). Shards that give no hint
  // at all still work when their first adds are of similar size; otherwise the first exchange round reports the mismatch.
  {
    const int N = h->cfg.num_shards;
    RC(ensure(h, h->d_meta, 16));
    RC(ensure(h, h->d_meta_all, 16 * (size_t)N));
    u64 mine[2] = {(u64)h->cfg.num_partitions, h->cfg.expected_instances};
    std::vector<u64> all(2 * (size_t)N);
    CK(cudaMemcpyAsync(h->d_meta.p, mine, 16, cudaMemcpyHostToDevice, h->s_meta));
    NC(a.AllGather(h->d_meta.p, h->d_meta_all.p, 2, ncclUint64, h->comm->meta, h->s_meta));
    CK(cudaMemcpyAsync(all.data(), h->d_meta_all.p, 16 * (size_t)N, cudaMemcpyDeviceToHost, h->s_meta));
    CK(cudaStreamSynchronize(h->s_meta));
    u64 hint = 0;
    for (int r = 0; r < N; r++) {
      if (all[2 * r] != mine[0])
        return fail(h, KC_ERR_INVALID, "shards disagree on cfg.num_partitions (%llu here, %llu on shard %d)", (unsigned long long)mine[0],
                    (unsigned long long)all[2 * r], r);
      hint = std::max(hint, all[2 * r + 1]);
    }
    h->cfg.expected_instances = hint;
  }
  return KC_OK;
#else
  return fail(h, KC_ERR_COMM, "built without nccl.h");
#endif
}

void kc_comm_destroy(kc_handle *h) { h->comm = nullptr; /* cached communicators live until process exit */ }

namespace {

// round metadata travels to the device as kernel parameters: a host-to-device copy would queue behind the bulk copy of
// the next batch of reads on the copy engine
constexpr int META_INLINE = 96;
struct MetaWords {
  u64 v[META_INLINE];
};
__global__ void meta_store_kernel(const MetaWords m, int n, u64 *dst) {
  if ((int)threadIdx.x < n) dst[threadIdx.x] = m.v[threadIdx.x];
}

// One exchange round over NCCL. Collective, but NOT tied to a batch number: a shard calls it whenever it has a parsed
// block to hand out (blk != null, from parse_batch right after the fill kernel was launched) and, from kc_finish, with blk == null and
// done == true until every shard has said "done" in the same round. Round r of one shard pairs with round r of every
// other shard, whatever each of them brought. Per round:
//   1. all-gather of [done, is_ctg, P, nsup for each destination] (N+3 words per shard),
//   2. one NCCL group on the comm stream: my directory slice (P entries) + my record range to every destination that gets
//      something, and the matching receives. Between a pair of ranks the order is (directory, records).
// The group is not waited for: it runs on s_comm behind the "fill done" event while the main stream parses the next batch;
// receives are turned into blocks in finish_exchange.
int exchange_round(kc_handle *h, const Block *blk, bool done, bool *all_done) {
#if KC_HAVE_NCCL_H
  NcclApi &a = nccl_api();
  const int N = h->cfg.num_shards, me = h->cfg.shard_id;
  const size_t MW = (size_t)N + 3;
  const size_t RB = (size_t)h->geom.rw * 4;  // bytes per supermer record
  cudaStream_t st = h->s_comm, sm = h->comm->meta != h->comm->comm ? h->s_meta : h->s_comm;
  RC(ensure(h, h->d_meta, MW * 8));
  RC(ensure(h, h->d_meta_all, MW * 8 * N));
  std::vector<u64> mine(MW, 0);
  mine[0] = done ? 1 : 0;
  mine[1] = blk && blk->is_ctg ? 1 : 0;
  mine[2] = (u64)h->P | ((u64)(h->auto_merge ? 1 : 0) << 32);
  if (blk && blk->tot_sup)
    for (int d = 0; d < N; d++)
      if (d != me) mine[3 + d] = blk->dest_base[d + 1] - blk->dest_base[d];
  if (MW <= (size_t)META_INLINE) {
    MetaWords mw;
    memcpy(mw.v, mine.data(), MW * 8);
    meta_store_kernel<<<1, META_INLINE, 0, sm>>>(mw, (int)MW, (u64 *)h->d_meta.p);
    h->st.gpu_launches++;
    CK(cudaGetLastError());
  } else {
    CK(cudaMemcpyAsync(h->d_meta.p, mine.data(), MW * 8, cudaMemcpyHostToDevice, sm));
  }
  NC(a.AllGather(h->d_meta.p, h->d_meta_all.p, MW, ncclUint64, h->comm->meta, sm));
  u64 *h_all = h->h_pin + PIN_META / 8;
  CK(cudaMemcpyAsync(h_all, h->d_meta_all.p, MW * 8 * N, cudaMemcpyDeviceToHost, sm));
  CK(cudaStreamSynchronize(sm));
  const u64 *all = h_all;
  bool everyone = true;
  u64 P = h->P;
  h->st.exchange_rounds++;
  for (int r = 0; r < N; r++) {
    everyone = everyone && all[r * MW] != 0;
    const u64 pr = all[r * MW + 2] & 0xFFFFFFFFull;
    if (pr && !h->P && !P) h->auto_merge = (all[r * MW + 2] >> 32) != 0;  // adopted geometry: merge neighbours like the peers do
    if (pr && !P) P = pr;
    if (pr && pr != P)
      return fail(h, KC_ERR_INVALID, "shards disagree on num_partitions (%llu vs %llu): set cfg.num_partitions or expected_instances",
                  (unsigned long long)pr, (unsigned long long)P);
  }
  if (all_done) *all_done = everyone;
  if (!h->P && P) {  // this shard had nothing to parse so far: take the peers' geometry
    h->P = (u32)P;
    h->st.num_partitions = h->P;
  }
  // receive buffers
  struct In { Pending p; int src; };
  std::vector<In> inc;
  for (int s = 0; s < N; s++) {
    if (s == me) continue;
    const u64 nsup = all[s * MW + 3 + me];
    if (!nsup) continue;
    In in;
    in.src = s;
    in.p.nsup = nsup;
    in.p.is_ctg = all[s * MW + 1] != 0;
    CK(dev_alloc((void **)&in.p.dir, sizeof(BucketHist) * P));
    CK(dev_alloc((void **)&in.p.rec, RB * nsup + 64));
    inc.push_back(in);
  }
  bool any_send = false;
  for (int d = 0; d < N; d++) any_send = any_send || mine[3 + d] != 0;
  if (inc.empty() && !any_send) return KC_OK;
  if (any_send) CK(cudaStreamWaitEvent(st, h->ev_blk, 0));  // the records are being written by the fill kernel on a parse slot's stream
  std::pair<cudaEvent_t, cudaEvent_t> ev;  // what the group occupies on the communication stream: ms_comm
  CK(cudaEventCreate(&ev.first));
  CK(cudaEventCreate(&ev.second));
  CK(cudaEventRecord(ev.first, st));
  NC(a.GroupStart());
  u64 sent = 0, recvd = 0;
  for (int d = 0; d < N; d++) {
    const u64 nsup = mine[3 + d];
    if (!nsup) continue;
    NC(a.Send(blk->d_hist + (size_t)d * P, sizeof(BucketHist) * P, ncclUint8, d, h->comm->comm, st));
    NC(a.Send(blk->rec + blk->dest_base[d] * h->geom.rw, RB * nsup, ncclUint8, d, h->comm->comm, st));
    sent += sizeof(BucketHist) * P + RB * nsup;
  }
  for (auto &in : inc) {
    NC(a.Recv(in.p.dir, sizeof(BucketHist) * P, ncclUint8, in.src, h->comm->comm, st));
    NC(a.Recv(in.p.rec, RB * in.p.nsup, ncclUint8, in.src, h->comm->comm, st));
    recvd += sizeof(BucketHist) * P + RB * in.p.nsup;
  }
  NC(a.GroupEnd());
  CK(cudaEventRecord(ev.second, st));
  h->comm_ev.push_back(ev);
  for (auto &in : inc) h->pending.push_back(in.p);
  h->st.supermer_bytes_sent += sent;
  h->st.supermer_bytes_recv += recvd;
  return KC_OK;
#else
  (void)blk; (void)done; (void)all_done;
  return fail(h, KC_ERR_COMM, "built without nccl.h");
#endif
}

}  // namespace

// NCCL transport, the part that runs in kc_finish: empty rounds until every shard is done, then the receives become blocks.
// ms_exchange is the exchange time the count phase actually waits for (what the per-batch rounds did not hide).
namespace {
// drain: empty "done" rounds until every shard has said done in the same round. Every shard leaves at the same round index,
// so the next phase (contigs after reads, kcount.cpp:148-150) starts with the rounds paired up again.
int comm_drain(kc_handle *h) {
  for (bool all_done = false; !all_done;) RC(exchange_round(h, nullptr, true, &all_done));
  h->phase_open = false;
  return KC_OK;
}
}  // namespace

extern "C" int kc_flush(kc_handle *h) {
  if (!h) return KC_ERR_INVALID;
  if (h->finished) return fail(h, KC_ERR_STATE, "flush after finish");
  if (h->cfg.num_shards < 2) return KC_OK;
#if KC_HAVE_NCCL_H
  if (!h->comm) return fail(h, KC_ERR_COMM, "multi-shard handle without a communicator: call kc_comm_init first");
  CK(cudaSetDevice(h->cfg.device));
  return comm_drain(h);
#else
  return fail(h, KC_ERR_COMM, "built without nccl.h");
#endif
}

int kc_exchange(kc_handle *h) {
#if KC_HAVE_NCCL_H
  if (!h->comm) return fail(h, KC_ERR_COMM, "multi-shard handle without a communicator: call kc_comm_init first");
  CK(cudaEventRecord(h->ev_x0, h->s_main));  // the main stream is idle here: x0..x1 is what kc_finish waits for
  if (h->phase_open) RC(comm_drain(h));  // a shard that flushed and added nothing since has already taken its last round
  if (!h->P) decide_partitions(h, 0);
  for (auto &pd : h->pending) RC(make_rblock(h, pd.dir, pd.rec, pd.nsup, pd.is_ctg, h->s_comm));
  h->pending.clear();
  CK(cudaEventRecord(h->ev_fill, h->s_comm));
  CK(cudaStreamWaitEvent(h->s_main, h->ev_fill, 0));
  CK(cudaEventRecord(h->ev_x1, h->s_main));
  CK(cudaStreamSynchronize(h->s_main));
  CK(cudaStreamSynchronize(h->s_comm));
  h->st.ms_exchange += elapsed(h->ev_x0, h->ev_x1);
  for (auto &pr : h->comm_ev) {
    h->st.ms_comm += elapsed(pr.first, pr.second);
    cudaEventDestroy(pr.first);
    cudaEventDestroy(pr.second);
  }
  h->comm_ev.clear();
  return KC_OK;
#else
  return fail(h, KC_ERR_COMM, "built without nccl.h");
#endif
}


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
        const u64 nsup = blk.dest_base[d + 1] - blk.dest_base[d];
        if (!nsup) continue;
        kc_handle *hd = hs[d];
        const size_t RB = (size_t)h->geom.rw * 4;
        u32 *rec = nullptr;
        BucketHist *dir = nullptr;
        CK(cudaSetDevice(hd->cfg.device));
        CK(dev_alloc((void **)&rec, RB * nsup + 64));
        CK(dev_alloc((void **)&dir, sizeof(BucketHist) * P));
        // the source was synchronised above; everything on the destination side is ordered on its main stream
        CK(cudaMemcpyPeerAsync(rec, hd->cfg.device, blk.rec + blk.dest_base[d] * h->geom.rw, h->cfg.device, RB * nsup, hd->s_main));
        CK(cudaMemcpyPeerAsync(dir, hd->cfg.device, blk.d_hist + (size_t)d * P, h->cfg.device, sizeof(BucketHist) * P, hd->s_main));
        RC(make_rblock(hd, dir, rec, nsup, blk.is_ctg, hd->s_main));
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

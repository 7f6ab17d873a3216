// kcount_b200.cpp — the reference-side binding of libkcount_b200.so (include/kcount_b200.h).
//
// A maintainer adds this ONE file to the reference tree as src/kcount/kcount_b200.cpp and selects it in
// src/kcount/CMakeLists.txt:24-44 in place of kcount_cpu.cpp / kcount_gpu.cpp (INTEGRATION.md has the CMake lines). It
// implements the reference's own backend plug point, the two pimpl classes
//     SeqBlockInserter<MAX_K>   (kcount.hpp:57-69)
//     HashTableInserter<MAX_K>  (kmer_dht.hpp:95-118)
// on top of the C ABI, so that kcount.cpp (count_kmers / add_ctg_kmers / analyze_kmers), kmer_dht.cpp, contigging.cpp
// and the consumer dbjg_traversal.cpp compile and run UNCHANGED: analyze_kmers<MAX_K> keeps its signature and its
// post-condition (kmer_dht->local_kmers holds this rank's surviving k-mers with count and chosen extensions).
//
// Differences from kcount_cpu.cpp in what happens behind the interface:
//   * process_seq only stages the sequence; done_processing hands the staged block to the GPU (parse -> supermers);
//   * nothing goes through kmer_dht->add_supermer / ThreeTierAggrStore: the supermer exchange between ranks is the
//     NCCL all-to-all inside the library (kc_comm_init + kc_finish), so HashTableInserter::insert_supermer is never
//     called (it DIEs if it is);
//   * insert_into_local_hashtable runs exchange + count + filter on the GPU and drains the survivors into local_kmers.
// This file contains no CUDA and includes no torch header: plain C++17 + the C ABI.
#include <array>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "upcxx_utils.hpp"
#include "kcount.hpp"
#include "kmer_dht.hpp"

#include "kcount_b200.h"

using namespace std;
using namespace upcxx;
using namespace upcxx_utils;

namespace {

// One kcount round is live per MAX_K at a time (contigging.cpp:115-118 builds a KmerDHT, runs analyze_kmers, traverses,
// destroys it). SeqBlockInserter has no access to KmerDHT's private ht_inserter, so the two pimpl classes meet here.
template <int MAX_K>
struct B200Round {
  kc_handle *h = nullptr;
  bool using_ctg_kmers = false;
  uint64_t expected_instances = 0;
  static B200Round &get() {
    static B200Round r;
    return r;
  }
};

// where the wall time of a round goes on this side of the C ABI (seconds, accumulated per process; read by benchmarks through
// kcount_b200_seam_seconds): [0] staging sequences in process_seq, [1] inside the library (kc_add_* / kc_flush / kc_finish /
// kc_copy_survivors), [2] filling the reference's local_kmers hash map from the survivor arrays
double g_seam_seconds[3] = {0, 0, 0};
struct SeamTimer {
  int slot;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  explicit SeamTimer(int s) : slot(s) {}
  ~SeamTimer() { g_seam_seconds[slot] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
};

void die_on(kc_handle *h, int rc, const char *what) {
  if (rc != KC_OK) DIE("kcount_b200: ", what, " failed (", rc, "): ", kc_last_error(h));
}

// KCOUNT_SEQ_BLOCK_SIZE (CMakeDefinitions.txt:51) is 3 MB for the reference GPU path; a B200 batch is far larger
constexpr size_t STAGE_BYTES = 256u << 20;

}  // namespace

// ------------------------------------------------------------------------------------------ SeqBlockInserter ----

template <int MAX_K>
struct SeqBlockInserter<MAX_K>::SeqBlockInserterState {
  int qual_offset = 0;
  vector<char> seqs;            // ASCII, lower case = low quality, exactly what count_kmers produces (kcount.cpp:81-84)
  vector<uint64_t> offsets{0};
  vector<double> depths;
  int64_t num_kmers = 0;
};

template <int MAX_K>
SeqBlockInserter<MAX_K>::SeqBlockInserter(int qual_offset, int minimizer_len) {
  state = new SeqBlockInserterState();
  state->qual_offset = qual_offset;
  // the library derives the minimizer length from k exactly as KmerDHT's constructor does (kmer_dht.cpp:113-116)
  (void)minimizer_len;
}

template <int MAX_K>
SeqBlockInserter<MAX_K>::~SeqBlockInserter() {
  if (state) delete state;
}

template <int MAX_K>
static void flush_block(typename SeqBlockInserter<MAX_K>::SeqBlockInserterState *st) {
  auto &round = B200Round<MAX_K>::get();
  const uint64_t n = st->offsets.size() - 1;
  if (!n) return;
  if (!round.h) DIE("kcount_b200: process_seq before HashTableInserter::init");
  SeamTimer tm(1);
  int rc;
  if (round.using_ctg_kmers)  // kcount_cpu.cpp:459-462: after init_ctg_kmers every supermer is a contig supermer
    rc = kc_add_contigs(round.h, st->seqs.data(), st->offsets.data(), st->depths.data(), n);
  else
    rc = kc_add_reads_ascii(round.h, st->seqs.data(), st->offsets.data(), n);
  die_on(round.h, rc, round.using_ctg_kmers ? "kc_add_contigs" : "kc_add_reads_ascii");
  st->seqs.clear();
  st->offsets.assign(1, 0);
  st->depths.clear();
}

template <int MAX_K>
void SeqBlockInserter<MAX_K>::process_seq(string &seq, kmer_count_t depth, dist_object<KmerDHT<MAX_K>> &kmer_dht) {
  (void)kmer_dht;
  const unsigned k = Kmer<MAX_K>::get_k();
  if (seq.length() < k + 2) return;  // holds no counted k-mer (kcount_cpu.cpp:84-102 starts at 1, stops before the last)
  SeamTimer tm(0);
  state->num_kmers += seq.length() - k - 1;
  state->seqs.insert(state->seqs.end(), seq.begin(), seq.end());
  state->offsets.push_back(state->seqs.size());
  state->depths.push_back((double)depth);  // 0 -> 1 happens in the library (kcount_cpu.cpp:75)
  if (state->seqs.size() >= STAGE_BYTES) {
    flush_block<MAX_K>(state);
  }
}

template <int MAX_K>
void SeqBlockInserter<MAX_K>::done_processing(dist_object<KmerDHT<MAX_K>> &kmer_dht) {
  (void)kmer_dht;
  flush_block<MAX_K>(state);
  // ranks stage different numbers of batches; every batch is one exchange round inside the library. kc_flush keeps this
  // rank taking (empty) rounds until every rank is done adding, so no rank waits in a round while a peer sits in the
  // reductions and barriers that follow (kcount.cpp:93-96, kmer_dht.cpp:227-231)
  if (auto &round = B200Round<MAX_K>::get(); round.h) {
    SeamTimer tm(1);
    die_on(round.h, kc_flush(round.h), "kc_flush");
  }
  auto tot_kmers = reduce_one(state->num_kmers, op_fast_add, 0).wait();
  SLOG_VERBOSE("Processed a total of ", tot_kmers, " kmers\n");  // kcount_cpu.cpp:107-112
}

// ----------------------------------------------------------------------------------------- HashTableInserter ----

template <int MAX_K>
struct HashTableInserter<MAX_K>::HashTableInserterState {
  kc_handle *h = nullptr;
};

template <int MAX_K>
HashTableInserter<MAX_K>::HashTableInserter() {}

template <int MAX_K>
HashTableInserter<MAX_K>::~HashTableInserter() {
  if (state) {
    if (state->h) kc_destroy(state->h);
    auto &round = B200Round<MAX_K>::get();
    if (round.h == state->h) round = B200Round<MAX_K>();
    delete state;
  }
}

template <int MAX_K>
void HashTableInserter<MAX_K>::init(int num_elems, bool use_qf) {
  this->use_qf = use_qf;
  state = new HashTableInserterState();
  kc_config cfg;
  kc_config_default(&cfg, (int32_t)Kmer<MAX_K>::get_k());
  cfg.dmin_thres = _dmin_thres;  // set by analyze_kmers before count_kmers runs (kcount.cpp:145); re-read at finish
  cfg.num_shards = rank_n();     // the table is sharded exactly as the UPC++ ranks shard it (kmer_dht.cpp:192-196)
  cfg.shard_id = rank_me();
  // one rank per GPU of the node, as the reference does (gpu_utils.cpp:68-75: rank % device count);
  // KCOUNT_B200_DEVICES restricts the ranks to the first n devices
  int32_t ndev = 1;
  if (kc_device_count(&ndev) != KC_OK || ndev < 1) DIE("kcount_b200: no CUDA device");
  if (const char *e = getenv("KCOUNT_B200_DEVICES")) ndev = max(1, min((int)ndev, atoi(e)));
  cfg.device = local_team().rank_me() % ndev;
  // num_elems = KmerDHT's my_num_kmers = estimate_num_kmers's instances per rank (contigging.cpp:61-91, kmer_dht.cpp:152);
  // every rank passes the same reduce_all'ed value, so every shard derives the same partition count
  cfg.expected_instances = (uint64_t)max(num_elems, 1);
  die_on(nullptr, kc_create(&cfg, &state->h), "kc_create");
  if (rank_n() > 1) {
    array<uint8_t, 128> id{};
    if (!rank_me() && kc_comm_unique_id(id.data()) != KC_OK) DIE("kcount_b200: kc_comm_unique_id failed");
    id = broadcast(id, 0).wait();
    die_on(state->h, kc_comm_init(state->h, id.data()), "kc_comm_init");
  }
  auto &round = B200Round<MAX_K>::get();
  round.h = state->h;
  round.using_ctg_kmers = false;
}

template <int MAX_K>
void HashTableInserter<MAX_K>::init_ctg_kmers(int max_elems) {
  (void)max_elems;  // contig k-mers share the read k-mers' table; nothing to pre-size
  B200Round<MAX_K>::get().using_ctg_kmers = true;
}

template <int MAX_K>
void HashTableInserter<MAX_K>::insert_supermer(const std::string &supermer_seq, kmer_count_t supermer_count) {
  (void)supermer_seq;
  (void)supermer_count;
  DIE("kcount_b200: insert_supermer reached — supermers travel over NCCL inside the library, not through kmer_store");
}

template <int MAX_K>
void HashTableInserter<MAX_K>::flush_inserts() {
  // KmerDHT::flush_updates calls this after each of the two passes (kmer_dht.cpp:227-231); the exchange and the count
  // run once, in insert_into_local_hashtable, when reads AND contigs are known
  barrier();
}

template <int MAX_K>
void HashTableInserter<MAX_K>::insert_into_local_hashtable(dist_object<KmerMap<MAX_K>> &local_kmers) {
  kc_handle *h = state->h;
  uint64_t n = 0;
  const int N_LONGS = (int)Kmer<MAX_K>::get_N_LONGS();
  vector<uint64_t> words;
  vector<uint16_t> counts;
  vector<char> left, right;
  {
    SeamTimer tm(1);
    die_on(h, kc_set_dmin_thres(h, _dmin_thres), "kc_set_dmin_thres");
    die_on(h, kc_finish(h), "kc_finish");  // collective: NCCL all-to-all of the supermers, then count + filter per shard
    die_on(h, kc_num_survivors(h, &n), "kc_num_survivors");
    words.resize(n * N_LONGS);
    counts.resize(n);
    left.resize(n);
    right.resize(n);
    die_on(h, kc_copy_survivors(h, words.data(), counts.data(), left.data(), right.data()), "kc_copy_survivors");
  }
  {
    SeamTimer tm(2);
    local_kmers->reserve(n);
    for (uint64_t i = 0; i < n; i++) {
      Kmer<MAX_K> kmer(&words[i * N_LONGS]);  // kmer.hpp:75
      KmerCounts kmer_counts = {.uutig_frag = nullptr, .count = counts[i], .left = left[i], .right = right[i]};
      local_kmers->insert({kmer, kmer_counts});  // kcount_cpu.cpp:513-521
    }
  }
  kc_stats st;
  die_on(h, kc_get_stats(h, &st), "kc_get_stats");
  barrier();
  auto tot_distinct = reduce_one(st.distinct, op_fast_add, 0).wait();
  auto tot_purged = reduce_one(st.purged, op_fast_add, 0).wait();
  auto tot_surv = reduce_one(st.survivors, op_fast_add, 0).wait();
  auto max_surv = reduce_one(st.survivors, op_fast_max, 0).wait();
  SLOG_VERBOSE("B200 kcount: ", tot_distinct, " distinct kmers, purged ", perc_str(tot_purged, tot_distinct), ", kept ", tot_surv,
               " (avg per rank ", tot_surv / rank_n(), ", max ", max_surv, "); nothing dropped (", st.table_retries,
               " partitions re-run in sub-passes)\n");  // kcount_cpu.cpp:467-487,525-527
}

// benchmarks: where a round's wall time went on this side of the ABI (see g_seam_seconds); resets the counters
extern "C" void kcount_b200_seam_seconds(double out3[3]) {
  for (int i = 0; i < 3; i++) {
    out3[i] = g_seam_seconds[i];
    g_seam_seconds[i] = 0;
  }
}

// explicit instantiations for every MAX_K the tree is built for, as kcount_cpu.cpp:533-555 does
#define SEQ_BLOCK_INSERTER_K(KMER_LEN) template struct SeqBlockInserter<KMER_LEN>;
#define HASH_TABLE_INSERTER_K(KMER_LEN) template class HashTableInserter<KMER_LEN>;

SEQ_BLOCK_INSERTER_K(32);
HASH_TABLE_INSERTER_K(32);
#if MAX_BUILD_KMER >= 64
SEQ_BLOCK_INSERTER_K(64);
HASH_TABLE_INSERTER_K(64);
#endif
#if MAX_BUILD_KMER >= 96
SEQ_BLOCK_INSERTER_K(96);
HASH_TABLE_INSERTER_K(96);
#endif
#if MAX_BUILD_KMER >= 128
SEQ_BLOCK_INSERTER_K(128);
HASH_TABLE_INSERTER_K(128);
#endif
#if MAX_BUILD_KMER >= 160
#error "kcount_b200 supports k <= 127 (MAX_BUILD_KMER <= 128): libkcount_b200.so rejects longer k-mers"
#endif
#undef SEQ_BLOCK_INSERTER_K
#undef HASH_TABLE_INSERTER_K

// TEST INFRASTRUCTURE ONLY — the drop-in proof (SURVEY.md §7.2 "boundary proof").
//
// Builds the reference's OWN analyze_kmers (kcount/kcount.cpp: count_kmers, add_ctg_kmers, analyze_kmers), its KmerDHT
// (kcount/kmer_dht.cpp) and Kmer (kmer.cpp), #included UNMODIFIED from /root/reference, together with
// integration/kcount_b200.cpp — this repository's implementation of the reference's backend plug point
// (SeqBlockInserter / HashTableInserter) over the C ABI — in place of kcount_cpu.cpp. The test then calls the reference's
// analyze_kmers<MAX_K> on PackedReads + Contigs and reads kmer_dht->local_kmers, exactly what dbjg_traversal.cpp:301
// consumes. Output: oracle/_ref/libkcount_dropin.so (git-ignored, links ../../mhm2_proxy_b200/libkcount_b200.so).
//
// PackedRead / PackedReads are the reference's own packed_reads.cpp (compiled unmodified into _ref/packed_reads.o). The
// three Contigs members that kcount.cpp calls are restated below from contigs.cpp:81-90 (that translation unit needs
// upcxx-utils' dist_ofstream and the file layer, which are out of scope).
#include <array>
#include <algorithm>
#include <bitset>
#include <cassert>
#include <cctype>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <iomanip>
#include <iostream>
#include <iterator>
#include <limits>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <string_view>
#include <unordered_map>
#include <utility>
#include <vector>

#include "upcxx_utils.hpp"

#include "kmer.cpp"
#include "kcount/kmer_dht.cpp"
#include "kcount/kcount.cpp"
#include "../integration/kcount_b200.cpp"

// what cmake generates from the extern-template inputs (src/CMakeLists.txt, kcount/CMakeLists.txt:3-12)
__MACRO_KMER__(32, template);
__MACRO_KMER__(64, template);
__MACRO_KMER__(96, template);
__MACRO_KMER__(128, template);
__MACRO_KCOUNT__(32, template);
__MACRO_KCOUNT__(64, template);
__MACRO_KCOUNT__(96, template);
__MACRO_KCOUNT__(128, template);

// utils.cpp:71-75 (declared in utils.hpp, used by KmerDHT's constructor)
size_t estimate_hashtable_memory(size_t num_elements, size_t element_size) {
  HASH_TABLE<char, char> tmp;
  double max_load_factor = tmp.max_load_factor();
  size_t expanded_num_elements = num_elements / max_load_factor + 1;
  return expanded_num_elements * element_size;
}

// PackedRead / PackedReads: the reference's own packed_reads.cpp, compiled where it lies into _ref/packed_reads.o (Makefile)

// ---- contigs.cpp:81-90 restated ----
void Contigs::clear() {
  contigs.clear();
  vector<Contig>().swap(contigs);
}
void Contigs::add_contig(Contig contig) { contigs.push_back(contig); }
size_t Contigs::size() { return contigs.size(); }

namespace {

struct Result {
  int n_longs = 0;
  uint64_t n_local = 0;
  double t_analyze = 0;  // wall seconds inside the reference's analyze_kmers (reads in PackedReads -> local_kmers filled)
  std::vector<uint64_t> keys;
  std::vector<uint16_t> counts;
  std::vector<char> left, right;
};

template <int MAX_K>
int run(int k, int qual_offset, int dmin_thres, const uint8_t *bytes, const uint64_t *offs, uint64_t n_reads, const char *cseqs,
        const uint64_t *coffs, const double *cdepths, uint64_t n_ctgs, Result *res, bool want_rows = true) {
  Kmer<MAX_K>::set_k(k);
  upcxx::shim_state::nranks() = 1;
  upcxx::shim_state::me() = 0;
  upcxx_utils::shim_mem::free_mem() = 64e9;
  res->n_longs = Kmer<MAX_K>::get_N_LONGS();
  // PackedReads as merge_reads leaves them: one PackedRead per read, bytes = base | qual << 3 (packed_reads.cpp:73-109)
  vector<PackedRead> prs(n_reads);
  uint64_t inst = 0;
  for (uint64_t r = 0; r < n_reads; r++) {
    const uint64_t len = offs[r + 1] - offs[r];
    prs[r].read_id = (r & 1) ? (int64_t)(r / 2 + 1) : -(int64_t)(r / 2 + 1);
    prs[r].read_len = (uint16_t)len;
    prs[r].bytes = new unsigned char[len ? len : 1];
    memcpy(prs[r].bytes, bytes + offs[r], len);
    if (len >= (uint64_t)k + 2) inst += len - k - 1;
  }
  PackedReads packed_reads(qual_offset, prs);
  vector<PackedReads *> packed_reads_list = {&packed_reads};
  Contigs ctgs;
  for (uint64_t c = 0; c < n_ctgs; c++) {
    Contig ctg{.id = (int64_t)c, .seq = std::string(cseqs + coffs[c], cseqs + coffs[c + 1]), .depth = cdepths[c]};
    ctgs.add_contig(ctg);
    if (ctg.seq.size() >= (size_t)k + 2) inst += ctg.seq.size() - k - 1;
  }
  {
    // contigging.cpp:115-118: the KmerDHT is built for the estimated number of k-mers, then analyze_kmers runs
    dist_object<KmerDHT<MAX_K>> kmer_dht((uint64_t)std::max<uint64_t>(inst, 1000), (int)0, (int)100, false, false);
    const auto t0 = std::chrono::steady_clock::now();
    analyze_kmers<MAX_K>((unsigned)k, 0u, qual_offset, packed_reads_list, dmin_thres, ctgs, kmer_dht, false);
    res->t_analyze = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (want_rows) {
    struct Row { std::array<uint64_t, 4> w; uint16_t c; char l, r; };
    std::vector<Row> rows;
    rows.reserve(kmer_dht->get_local_num_kmers());
    for (auto it = kmer_dht->local_kmers_begin(); it != kmer_dht->local_kmers_end(); ++it) {
      Row row{};
      for (int l = 0; l < res->n_longs; l++) row.w[l] = it->first.get_longs()[l];
      row.c = it->second.count; row.l = it->second.left; row.r = it->second.right;
      rows.push_back(row);
    }
    std::sort(rows.begin(), rows.end(), [](const Row &a, const Row &b) { return a.w < b.w; });
    for (auto &row : rows) {
      for (int l = 0; l < res->n_longs; l++) res->keys.push_back(row.w[l]);
      res->counts.push_back(row.c); res->left.push_back(row.l); res->right.push_back(row.r);
    }
    }
    res->n_local = kmer_dht->get_local_num_kmers();
  }
  return 0;
}

}  // namespace

extern "C" {
void *odrop_run(int k, int qual_offset, int dmin_thres, const uint8_t *bytes, const uint64_t *offs, uint64_t n_reads, const char *cseqs,
                const uint64_t *coffs, const double *cdepths, uint64_t n_ctgs) {
  Result *res = new Result();
  int rc = -1;
  switch ((k / 32 + 1) * 32) {  // main.cpp:170
    case 32: rc = run<32>(k, qual_offset, dmin_thres, bytes, offs, n_reads, cseqs, coffs, cdepths, n_ctgs, res); break;
    case 64: rc = run<64>(k, qual_offset, dmin_thres, bytes, offs, n_reads, cseqs, coffs, cdepths, n_ctgs, res); break;
    case 96: rc = run<96>(k, qual_offset, dmin_thres, bytes, offs, n_reads, cseqs, coffs, cdepths, n_ctgs, res); break;
    case 128: rc = run<128>(k, qual_offset, dmin_thres, bytes, offs, n_reads, cseqs, coffs, cdepths, n_ctgs, res); break;
  }
  if (rc != 0) { delete res; return nullptr; }
  return res;
}
// timing of the seam at bench size: analyze_kmers only, the table stays in local_kmers (no sorted copy); returns seconds
extern "C" void kcount_b200_seam_seconds(double out3[3]);
// split of the last odrop_time: seconds staging in process_seq / inside libkcount_b200 / filling local_kmers
void odrop_seam_seconds(double out3[3]) { kcount_b200_seam_seconds(out3); }
double odrop_time(int k, int qual_offset, int dmin_thres, const uint8_t *bytes, const uint64_t *offs, uint64_t n_reads,
                  uint64_t *n_local_kmers) {
  Result res;
  double scrap[3];
  kcount_b200_seam_seconds(scrap);  // reset
  int rc = -1;
  switch ((k / 32 + 1) * 32) {
    case 32: rc = run<32>(k, qual_offset, dmin_thres, bytes, offs, n_reads, nullptr, nullptr, nullptr, 0, &res, false); break;
    case 64: rc = run<64>(k, qual_offset, dmin_thres, bytes, offs, n_reads, nullptr, nullptr, nullptr, 0, &res, false); break;
    case 96: rc = run<96>(k, qual_offset, dmin_thres, bytes, offs, n_reads, nullptr, nullptr, nullptr, 0, &res, false); break;
    case 128: rc = run<128>(k, qual_offset, dmin_thres, bytes, offs, n_reads, nullptr, nullptr, nullptr, 0, &res, false); break;
  }
  if (n_local_kmers) *n_local_kmers = res.n_local;
  return rc == 0 ? res.t_analyze : -1.0;
}
int odrop_n_longs(void *h) { return ((Result *)h)->n_longs; }
uint64_t odrop_num_survivors(void *h) { return ((Result *)h)->counts.size(); }
void odrop_copy_survivors(void *h, uint64_t *keys, uint16_t *counts, char *left, char *right) {
  Result *r = (Result *)h;
  memcpy(keys, r->keys.data(), r->keys.size() * 8);
  memcpy(counts, r->counts.data(), r->counts.size() * 2);
  memcpy(left, r->left.data(), r->left.size());
  memcpy(right, r->right.data(), r->right.size());
}
void odrop_free(void *h) { delete (Result *)h; }
}

// TEST INFRASTRUCTURE ONLY — the drop-in proof (SURVEY.md §7.2 "boundary proof").
//
// Builds the reference's OWN analyze_kmers (kcount/kcount.cpp: count_kmers, add_ctg_kmers, analyze_kmers), its KmerDHT
// (kcount/kmer_dht.cpp) and Kmer (kmer.cpp), #included UNMODIFIED from /root/reference, together with
// integration/kcount_b200.cpp — this repository's implementation of the reference's backend plug point
// (SeqBlockInserter / HashTableInserter) over the C ABI — in place of kcount_cpu.cpp. The test then calls the reference's
// analyze_kmers<MAX_K> on PackedReads + Contigs and reads kmer_dht->local_kmers, exactly what dbjg_traversal.cpp:301
// consumes. Output: oracle/_ref/libkcount_dropin.so (git-ignored, links ../../mhm2_proxy_b200/libkcount_b200.so).
//
// The few PackedRead / PackedReads / Contigs members that kcount.cpp calls are restated below from packed_reads.cpp and
// contigs.cpp (those translation units drag in the FASTQ reader and file layer, which are out of scope).
#include <array>
#include <algorithm>
#include <bitset>
#include <cassert>
#include <cctype>
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

// ---- packed_reads.cpp:66-72,91-159,178-207 restated: only what count_kmers touches ----
PackedRead::PackedRead()
    : read_id(0)
    , read_len(0)
    , bytes(nullptr) {}
PackedRead::PackedRead(const PackedRead &copy)
    : read_id(copy.read_id)
    , read_len(copy.read_len)
    , bytes(new unsigned char[read_len]) {
  memcpy(bytes, copy.bytes, read_len);
}
PackedRead::PackedRead(PackedRead &&move)
    : read_id(move.read_id)
    , read_len(move.read_len)
    , bytes(move.bytes) {
  move.bytes = nullptr;
  move.read_len = 0;
}
PackedRead::~PackedRead() { clear(); }
void PackedRead::clear() {
  if (bytes) delete[] bytes;
  bytes = nullptr;
  read_len = 0;
  read_id = 0;
}
void PackedRead::unpack(string &read_id_str, string &seq, string &quals, int qual_offset) const {
  char pair_id = (read_id < 0 ? '1' : '2');
  read_id_str = "@r" + to_string(labs(read_id)) + '/' + pair_id;
  seq.resize(read_len);
  quals.resize(read_len);
  for (int i = 0; i < read_len; i++) {
    seq[i] = nucleotide_map[bytes[i] & 7];
    quals[i] = qual_offset + (bytes[i] >> 3);
  }
}
uint16_t PackedRead::get_read_len() { return read_len; }
unsigned char *PackedRead::get_raw_bytes() { return bytes; }

PackedReads::PackedReads(int qual_offset, vector<PackedRead> &new_packed_reads)
    : packed_reads(new_packed_reads)
    , index(0)
    , qual_offset(qual_offset)
    , fname("")
    , str_ids(false) {
  max_read_len = 0;
  for (auto &packed_read : new_packed_reads) max_read_len = max((unsigned)packed_read.get_read_len(), max_read_len);
}
PackedReads::~PackedReads() {}
bool PackedReads::get_next_read(string &id, string &seq, string &quals) {
  if (index == packed_reads.size()) return false;
  packed_reads[index].unpack(id, seq, quals, qual_offset);
  index++;
  return true;
}
void PackedReads::reset() { index = 0; }
int64_t PackedReads::get_local_num_reads() { return packed_reads.size(); }

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
  std::vector<uint64_t> keys;
  std::vector<uint16_t> counts;
  std::vector<char> left, right;
};

template <int MAX_K>
int run(int k, int qual_offset, int dmin_thres, const uint8_t *bytes, const uint64_t *offs, uint64_t n_reads, const char *cseqs,
        const uint64_t *coffs, const double *cdepths, uint64_t n_ctgs, Result *res) {
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
    analyze_kmers<MAX_K>((unsigned)k, 0u, qual_offset, packed_reads_list, dmin_thres, ctgs, kmer_dht, false);
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

/* kcount_b200.h — C ABI of the B200-native kcount (k-mer analysis) hot path of MetaHipMer2.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types, int status codes (0 = ok), no
 * exceptions. Every entry point names the reference interface it replaces (paths relative to the reference's src/).
 * The reference-side binding (implementing SeqBlockInserter<MAX_K>/HashTableInserter<MAX_K> over this header) is
 * integration/kcount_b200.cpp; INTEGRATION.md shows how it is added to the reference's build.
 *
 * Semantics = the reference CPU path (kcount/kcount_cpu.cpp), bit-exact: canonical k-mer, saturating u16 count,
 * chosen left/right extension char (A,C,G,T,F,X); 'N' packs as 'G' in k-mers and never counts as an extension; first
 * and last k-mer of every sequence are not counted. See DESIGN.md.
 */
#ifndef KCOUNT_B200_H
#define KCOUNT_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KC_OK 0
#define KC_ERR_INVALID 1   /* bad argument / configuration */
#define KC_ERR_CUDA 2      /* CUDA runtime error; text in kc_last_error */
#define KC_ERR_STATE 3     /* call out of order (e.g. add after finish) */
#define KC_ERR_NOMEM 4     /* device or host allocation failed */
#define KC_ERR_BADCHAR 5   /* contig character outside ACGTNacgtn (reference DIEs: kcount_cpu.cpp:455-459) */
#define KC_ERR_COMM 6      /* NCCL error */
#define KC_ERR_OVERFLOW 7  /* survivor output capacity exceeded (reported, never silent) */

typedef struct kc_handle kc_handle;

/* Values of the reference's run-time options and compile-time constants that parameterise the path
 * (options.cpp:260-304, CMakeDefinitions.txt:46-63). Zero-initialise, then call kc_config_default(). */
typedef struct kc_config {
  int32_t kmer_len;          /* k, 1..127; N_LONGS = k/32 + 1 words per key as in main.cpp:170 / kmer.hpp:64 */
  int32_t qual_cutoff;       /* KCOUNT_QUAL_CUTOFF = 20: packed quality below this => base is low quality */
  int32_t dmin_thres;        /* --min-depth-thres, default 2 (options.hpp:96); must be >= 1 */
  double dyn_min_depth;      /* DYN_MIN_DEPTH = 0.9 (kcount_cpu.cpp:178) */
  int32_t num_shards;        /* = upcxx::rank_n(): table is sharded by quick_hash(minimizer) % num_shards */
  int32_t shard_id;          /* = upcxx::rank_me(): which shard this handle owns */
  int32_t device;            /* CUDA device ordinal (reference: rank % ndev, gpu_utils.cpp:68-75) */
  int32_t num_partitions;    /* 0 = auto. Sub-tables per shard, each sized to live in one CTA's shared memory while it is
                                built (internal, result-invariant; all shards of a job must use the same value) */
  uint64_t expected_instances; /* hint: total k-mer instances this shard's caller will add (0 = size from first add);
                                  plays the role of estimate_num_kmers (contigging.cpp:61-91) */
  int32_t keep_prefilter;    /* 1 = also keep every distinct k-mer's counters after the reads pass (debug parity) */
  int32_t batch_reads;       /* 0 = auto. reads per internal batch (H2D copy / parse / exchange pipelining unit) */
  int32_t reserved[6];
} kc_config;

/* counters the reference logs (kcount.cpp:93-97, kcount_cpu.cpp:107-112,467-487,525-527) */
typedef struct kc_stats {
  uint64_t num_seqs;            /* reads + contigs added through this handle */
  uint64_t num_bases;
  uint64_t instances_parsed;    /* k-mer instances parsed locally: sum max(0, len-k-1) */
  uint64_t instances_inserted;  /* k-mer instances inserted into this shard's table (reads + contigs) */
  uint64_t ctg_instances_inserted;
  uint64_t supermers_out;       /* supermers produced locally */
  uint64_t supermer_bytes_out;  /* bytes of packed supermers produced locally */
  uint64_t supermer_bytes_sent; /* of those, bytes that left this GPU */
  uint64_t supermers_in;        /* supermers inserted here */
  uint64_t distinct;            /* distinct canonical k-mers in this shard before the filter */
  uint64_t purged;              /* distinct - survivors */
  uint64_t survivors;
  uint64_t table_retries;       /* work items re-run (split into sub-passes, or with wide counters) */
  uint64_t max_probe;
  uint64_t num_partitions;
  uint64_t supermer_bytes_recv; /* bytes of supermer records + directories received from other GPUs */
  uint64_t exchange_rounds;     /* exchange rounds this shard took part in (data rounds + drain rounds) */
  /* CUDA-event times on the library's own streams, summed over launches. ms_parse = parse + fill kernels plus the
   * bucket-layout round trip between them; ms_insert = the count kernel (insert + filter + compaction fused: the table
   * lives in shared memory and is filtered where it was built). ms_exchange = exchange time the main stream had to WAIT
   * for (what the overlap with parsing did not hide); ms_comm = time the exchange groups occupied the communication
   * stream (send + receive of supermer_bytes_sent / _recv: the figure NVLink GB/s is computed from). */
  double ms_h2d, ms_parse, ms_exchange, ms_insert, ms_comm, ms_total;
  double ms_parse_count, ms_parse_fill, ms_finish;
  uint64_t gpu_launches;        /* kernels launched by this handle so far */
  uint64_t n_parse_launches, n_insert_launches;
  uint64_t table_bytes_max;     /* shared memory per CTA of the count kernel (index table + key store + staging) */
  uint64_t reserved[4];
} kc_stats;

void kc_config_default(kc_config *cfg, int32_t kmer_len);
/* sizeof(kc_config) and sizeof(kc_stats) as the library was compiled: bindings in other languages check their mirrors */
void kc_struct_sizes(uint32_t *config_bytes, uint32_t *stats_bytes);

/* KmerDHT<MAX_K> ctor + HashTableInserter<MAX_K>::init (kmer_dht.cpp:105-154, kcount_cpu.cpp:425-443) */
int kc_create(const kc_config *cfg, kc_handle **out);
/* ~KmerDHT / ~HashTableInserter (kmer_dht.cpp:161-166, kcount_cpu.cpp:417-423) */
int kc_destroy(kc_handle *h);
const char *kc_last_error(const kc_handle *h);
/* Device buffers freed by a handle are cached per process for the next kcount round (one round per k, main.cpp:169);
 * this returns them to the driver (e.g. before the alignment stage needs the GPU memory). */
int kc_release_cached_memory(void);

/* count_kmers' read loop -> SeqBlockInserter::process_seq(seq, 0) (kcount.cpp:72-88, kcount_cpu.cpp:73-103).
 * bytes = PackedRead bytes verbatim (packed_reads.cpp:73-109: base 0..4 | min(qual,31) << 3), offsets[n_reads+1].
 * HOST pointers (pinned memory copies faster); the caller keeps ownership; returns when the data has been consumed. */
int kc_add_reads(kc_handle *h, const uint8_t *bytes, const uint64_t *offsets, uint64_t n_reads);
/* same, DEVICE pointers already resident on cfg.device (no copy). total_bytes = offsets[n_reads]. */
int kc_add_reads_device(kc_handle *h, const uint8_t *d_bytes, const uint64_t *d_offsets, uint64_t n_reads, uint64_t total_bytes,
                        uint32_t max_read_len);
/* PackedReads kept in HBM across the k rounds of one job: contigging.cpp:115-118 runs analyze_kmers once per k over the same
 * packed_reads_list, which the reference keeps in host memory for the whole run (main.cpp:169, packed_reads.hpp:120-167).
 * kc_reads_create uploads bytes/offsets (same layout as kc_add_reads; offsets[0] must be 0) asynchronously in batches and
 * returns at once; kc_add_reads_resident parses them into a handle (any number of handles, one after the other) and, used
 * right after the create, overlaps the upload batch by batch. The host arrays must stay valid until the first
 * kc_add_reads_resident returned (or kc_reads_destroy). 180 GB of HBM hold the reads of any single-node run. */
typedef struct kc_reads kc_reads;
int kc_reads_create(int32_t device, const uint8_t *bytes, const uint64_t *offsets, uint64_t n_reads, int32_t batch_reads,
                    kc_reads **out);
int kc_add_reads_resident(kc_handle *h, kc_reads *reads);
int kc_reads_destroy(kc_reads *reads);
/* the step before the path: FASTQ text -> PackedRead bytes on the device, exactly PackedRead's constructor
 * (packed_reads.cpp:73-109: A0 C1 G2 T3, N and IUPAC codes 4, quality min(q - qual_offset, 31) in the high 5 bits; any other
 * character is the reference's DIE and comes back as KC_ERR_BADCHAR from the first kc_add_reads_resident /
 * kc_reads_copy_packed). seqs and quals are the concatenated sequence and quality lines, offsets[n+1] as in kc_add_reads.
 * The reads then never exist in packed form on the host unless kc_reads_copy_packed fetches them. */
int kc_reads_create_fastq(int32_t device, const char *seqs, const char *quals, const uint64_t *offsets, uint64_t n_reads,
                          int32_t qual_offset, int32_t batch_reads, kc_reads **out);
int kc_reads_copy_packed(kc_reads *reads, uint8_t *bytes_out);
/* add_ctg_kmers' loop -> process_seq(ctg.seq, ctg.get_uint16_t_depth()) (kcount.cpp:123-130, contigs.hpp:57-66).
 * seqs = ASCII ACGTN (host), offsets[n+1], depths = Contig::depth doubles. Any length; long contigs are tiled. */
int kc_add_contigs(kc_handle *h, const char *seqs, const uint64_t *offsets, const double *depths, uint64_t n_ctgs);
/* reads in the form count_kmers hands them to SeqBlockInserter::process_seq(seq, 0) (kcount.cpp:81-86): ASCII ACGTN with
 * LOWER CASE = quality below the cut-off. Used by the pimpl binding (integration/kcount_b200.cpp); kc_add_reads, which
 * takes the PackedRead bytes directly, is the fast path. */
int kc_add_reads_ascii(kc_handle *h, const char *seqs, const uint64_t *offsets, uint64_t n_reads);
/* the global _dmin_thres (kmer_dht.hpp:57) is assigned by analyze_kmers AFTER the KmerDHT / HashTableInserter exist
 * (kcount.cpp:145): set it any time before kc_finish */
int kc_set_dmin_thres(kc_handle *h, int32_t dmin_thres);

/* multi-GPU: bind an NCCL communicator built from a 128-byte ncclUniqueId shared by all shards
 * (replaces ThreeTierAggrStore<Supermer> / rpc_ff, kmer_dht.cpp:133,146-149,222-231). Not needed when num_shards==1.
 * Collective discipline once a communicator is bound (it mirrors the reference's, SURVEY.md 8b "Threading"):
 *   - kc_add_reads* / kc_add_contigs run one EXCHANGE ROUND per internal batch (an all-gather of sizes + one group of
 *     sends/receives). Rounds pair up by index across shards, whatever each shard brought, so shards may add different
 *     amounts of data in a different number of calls — but a shard inside an add call needs its peers to keep taking
 *     rounds: between the last add of a phase and any OTHER collective or host barrier every shard must call kc_flush
 *     (the reference has the same rule: SeqBlockInserter::done_processing -> flush_updates -> barrier, kcount.cpp:93-96,
 *     kmer_dht.cpp:227-231). kc_finish flushes implicitly.
 *   - kc_flush, kc_finish, kc_finish_to_host are collective. */
int kc_comm_unique_id(uint8_t id128[128]);
int kc_comm_init(kc_handle *h, const uint8_t id128[128]);
/* KmerDHT::flush_updates (kmer_dht.cpp:227-231): collective. Takes part in exchange rounds until EVERY shard has called
 * kc_flush (or kc_finish); when it returns no shard is inside a round and everything sent so far has been received.
 * No-op for num_shards == 1. */
int kc_flush(kc_handle *h);
/* number of CUDA devices visible to this process (the binding's default device is rank_me() % this, as the reference's
 * gpu_utils.cpp:68-75 does) */
int kc_device_count(int32_t *n);

/* several shards inside ONE process (handles[i] = shard i of n, any devices): exchange by peer copies, then finish all.
 * Same role as kc_comm_init + kc_finish on every rank; used by tests and by single-process multi-GPU drivers. */
int kc_finish_local_group(kc_handle **handles, int32_t n);

/* KmerDHT::flush_updates + finish_updates -> HashTableInserter::insert_into_local_hashtable
 * (kmer_dht.cpp:227-236, kcount_cpu.cpp:490-528). Collective across shards when num_shards > 1. */
int kc_finish(kc_handle *h);
/* kc_finish + kc_copy_survivors in one call, for callers that want the table on the host (insert_into_local_hashtable fills
 * the host-side local_kmers, kcount_cpu.cpp:490-528): the count stage runs in slices and the survivors of one slice travel
 * to the host arrays while the next slice is counted. capacity = entries the arrays can hold (kmer_words: capacity *
 * N_LONGS); *n_out = survivors. More survivors than capacity -> KC_ERR_OVERFLOW with the first `capacity` entries delivered
 * and the whole table still on the device (kc_num_survivors / kc_copy_survivors). Pinned host arrays overlap best. */
int kc_finish_to_host(kc_handle *h, uint64_t *kmer_words, uint16_t *counts, char *left, char *right, uint64_t capacity,
                      uint64_t *n_out);

/* local_kmers of this shard (kmer_dht.hpp:119, dbjg_traversal.cpp:301): n rows of N_LONGS words, u16 count, left, right */
int kc_num_survivors(kc_handle *h, uint64_t *n);
int kc_copy_survivors(kc_handle *h, uint64_t *kmer_words, uint16_t *counts, char *left, char *right);
/* KmerDHT::dump_kmers (kmer_dht.cpp:238-266, what --dump-kmers / analyze_kmers(..., dump_kmers=true) writes): one line per
 * surviving k-mer, "KMER count left right"; a path ending in .gz is gzip-compressed like the reference's kmers-<k>.txt.gz.
 * Line order is unspecified (the reference iterates a hash map). */
int kc_dump_kmers(kc_handle *h, const char *path);

/* Order-independent digest of this shard's surviving table, computed on the device: out4 = {n, sum of counts,
 * sum over rows of mix(row) mod 2^64, xor over rows of mix(row)} with mix(row): x = 0x9E3779B97F4A7C15; for every key word
 * w: x = fmix64(x ^ w); x = fmix64(x ^ (count | left << 16 | right << 24)) (fmix64 = MurmurHash3's finaliser,
 * hash_funcs.c:65-74). Two tables are equal iff (with overwhelming probability) their digests are: bench.py compares it with
 * the same digest of the reference's local_kmers at full workload size. */
int kc_table_digest(kc_handle *h, uint64_t out4[4]);

/* No counterpart in the reference: added for callers that cannot allocate (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). The same digest as kc_table_digest over a table
 * the caller holds in DEVICE memory on `device` (e.g. the arrays of kc_survivors_device): n rows of n_longs key words, u16
 * counts, left and right chars; d_out4 (DEVICE memory, 4 x u64) receives {n, sum of counts, sum of mix, xor of mix}.
 * The call enqueues one 256-byte memset and ONE kernel on cuda_stream (a cudaStream_t; NULL = the default stream) and
 * returns without synchronising; it calls no cudaMalloc / cudaFree and no allocating library routine. Its scratch (one partial
 * result per CTA + a ticket) is a file-scope __device__ array of 128.25 KB that every GPU loading the library holds until
 * exit. LIMIT: n <= KC_DIGEST_STATIC_MAX_ROWS (2^28 rows = 4096 CTAs x 65536 rows); more is rejected with KC_ERR_INVALID and a
 * message in errbuf (optional, errbuf_len bytes). All calls on one device share the scratch: calls that overlap on the GPU
 * (different streams, no ordering between them) race — order them. The first call per device looks the array's address up
 * (cudaGetSymbolAddress): make it once before starting a stream capture. */
#define KC_DIGEST_STATIC_MAX_ROWS 268435456ull
int kc_table_digest_static(int32_t device, const uint64_t *d_kmer_words, const uint16_t *d_counts, const char *d_left,
                           const char *d_right, uint64_t n, int32_t n_longs, uint64_t *d_out4, void *cuda_stream, char *errbuf,
                           uint64_t errbuf_len);

/* device-resident view of the same arrays (valid until kc_destroy) */
int kc_survivors_device(kc_handle *h, const uint64_t **d_kmer_words, const uint16_t **d_counts, const char **d_left,
                        const char **d_right);
/* pre-filter counters after the reads pass (KmerMapExts contents, kcount_cpu.cpp:191-200): vals = n x 9 u16
 * {count, left A,C,G,T, right A,C,G,T}, saturated at 65535. Needs cfg.keep_prefilter. */
int kc_num_prefilter(kc_handle *h, uint64_t *n);
int kc_copy_prefilter(kc_handle *h, uint64_t *kmer_words, uint16_t *vals9);

int kc_get_stats(kc_handle *h, kc_stats *st);

/* test knob, not part of the product interface: upper bound on the distinct k-mers one shared-memory table holds
 * (0 = what fits), to force the sub-pass path on small inputs; result-invariant. Call before the first add. */
int kc_debug_set_key_cap(kc_handle *h, int32_t key_cap);

/* ---- primitives, exported for known-answer tests of the device code (each runs a tiny kernel) ----
 * kmers: n ASCII k-mers of length k back to back. Outputs per k-mer: words[N_LONGS], rc_words[N_LONGS],
 * minimizer (kmer.cpp:345-393), hash = Kmer::hash = MurmurHash3_x64_64 (kmer.cpp:465-468; the count stage's own slot hash is internal), owner = quick_hash(minimizer) %
 * num_shards (kmer_dht.cpp:192-196). */
int kc_prim_kmers(int32_t device, int32_t k, const char *kmers, uint64_t n, int32_t num_shards, uint64_t *words, uint64_t *rc_words,
                  uint64_t *minimizer, uint64_t *hash, int32_t *owner);
/* ExtCounts::get_ext (kcount_cpu.cpp:173-182) for n rows of {A,C,G,T,count} u16 */
int kc_prim_get_ext(int32_t device, const uint16_t *acgt_count5, uint64_t n, int32_t dmin_thres, double dyn_min_depth, char *out);

#ifdef __cplusplus
}
#endif
#endif

/* include/links_gpu.h -- C-ABI of the B200-native LINKS hot path (liblinks_gpu.so).
 *
 * The reference (bcgsc/LINKS v2.0.1) has no plugin/FFI layer: its hot path is the four public
 * stage methods of `class LINKS` called by LINKS_CPP's main():
 *     init_bloom_filter(), start_read_fasta(), start_read_contig(), pair_contigs()
 *     (reference src/LINKS.hpp:32-36, src/main.cpp:28-59),
 * configured by InputParser's public fields (src/InputParser.hpp:68-91) and observable only
 * through <b>.bloom, <b>.pairing_issues and <b>.tigpair_checkpoint.tsv.  This header is the
 * boundary a maintainer binds instead of those four bodies; each entry point cites what it
 * replaces.  Plain pointers and sizes only; no C++/torch types; no exceptions cross it.
 *
 * Conventions
 *  - Every function returns 0 on success or a negative LINKS_E_* code; links_last_error(ctx)
 *    gives the message.  There is NO CPU fallback: without a CUDA device links_create fails.
 *  - Caller owns host buffers passed in.  *_push calls return once the host buffer may be reused;
 *    device work continues asynchronously on the context's stream.
 *  - Sequences are raw ASCII as btllib::SeqReader hands them to LINKS in LONG_MODE (no case
 *    folding): `seq` is the concatenation of n records, record i = seq[off[i] .. off[i+1]).
 *  - Order of calls: create -> bf_alloc|bf_load -> contigs_push* -> reads_push* -> finalize.
 *    (The reference reads the reads before the contigs; results do not depend on that order
 *    because first-wins is defined by (read index, distance index, k-mer rank) = its -j 1 order.)
 *  - One context per process per GPU; calls on one context must not be concurrent.
 */
#ifndef LINKS_GPU_H
#define LINKS_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LINKS_OK 0
#define LINKS_E_INVAL (-1)   /* bad argument / call order */
#define LINKS_E_CUDA (-2)    /* CUDA runtime error (message has the details) */
#define LINKS_E_NOMEM (-3)   /* host or device allocation failed */
#define LINKS_E_LIMIT (-4)   /* input exceeds a documented limit */
#define LINKS_E_NODEVICE (-5)/* no usable CUDA device */

#define LINKS_MAX_K 512u          /* ntHash tables are staged in shared memory */
#define LINKS_MAX_DIST 256u       /* order key keeps 8 bits of distance index */
#define LINKS_MAX_HASH_NUM 1024u  /* btllib MAX_HASH_VALUES (bloom_filter.hpp) */
#define LINKS_MAX_TIGS ((1u << 29) - 1u) /* row key packs (tig, orientation) in 30 bits */

typedef struct links_ctx links_ctx;

/* Options that reach the hot path (reference src/InputParser.hpp:68-91; LINKS ctor src/LINKS.hpp:333-346). */
typedef struct links_params {
  uint32_t k;              /* -k */
  uint32_t n_dist;         /* number of -d values */
  const uint32_t* dist;    /* -d list */
  const uint16_t* step;    /* -t list ALREADY padded with its last value to n_dist (InputParser.hpp:224-229) */
  uint64_t offset;         /* -o */
  float insert_stdev;      /* -e */
  int32_t device;          /* CUDA device ordinal */
} links_params;

/* One aggregated row of <b>.tigpair_checkpoint.tsv (reference src/LINKS.hpp:1023-1038):
 *   bin \t [f|r]tig_a \t [f|r]tig_b \t links \t gaps
 * tig == 0 stands for the reference's empty contig id "" (an absent k-mer default-inserted by
 * track_all_test[...] at src/LINKS.hpp:827-830); print it as the empty string. */
typedef struct links_row {
  int32_t bin;
  uint32_t tig_a, tig_b;
  uint8_t rev_a, rev_b;
  uint8_t pad_[2];
  uint64_t links;
  int64_t gaps;
} links_row;

/* One record of <b>.pairing_issues (reference src/LINKS.hpp:915-949). */
typedef struct links_issue {
  uint32_t tig;
  uint8_t o1, o2;
  uint8_t pad_[2];
  uint64_t tig_len;
  uint64_t s1, e1, s2, e2;
  uint64_t pet_size;
  int64_t insert_size;
} links_issue;

/* Counters.  The ct_* ones are the order-independent subset of the reference's
 * "PAIRED K-MER STATS" block (src/LINKS.hpp:981-1012). */
typedef struct links_stats {
  uint64_t n_candidates;   /* (A,B) k-mer pairs visited (before the Bloom test) */
  uint64_t n_survivors;    /* passed both Bloom tests (src/LINKS.hpp:554-555) */
  uint64_t n_unique_pairs; /* distinct ordered (h_a,h_b) = mate_pair.size() */
  uint64_t n_bf_probes;    /* Bloom bit tests of the reference's early-exit loop (bloom_filter.hpp:648-659) over these k-mers (roofline accounting) */
  uint64_t n_contig_kmers; /* valid contig k-mer positions hashed into the table (x2 strands) */
  uint64_t ct_ok_pairs, ct_problem_pairs, ct_ok_contig, ct_iz_issues, ct_illogical, ct_gate_fail;
  uint64_t n_rows;
  uint64_t n_bf_loads;       /* Bloom bit tests the kernels issued (proven members skip theirs, see links_bf_covers_contigs) */
  uint64_t n_member_sectors; /* 32-byte sectors of the exact member set read instead */
} links_stats;

/* ---- lifetime ---- */
int links_create(links_ctx** out, const links_params* p);
void links_destroy(links_ctx* ctx);
const char* links_last_error(const links_ctx* ctx); /* ctx may be NULL: last create error */
int links_sync(links_ctx* ctx);                      /* wait for all queued device work */
/* Run the compute work on a caller-owned cudaStream_t (e.g. torch's current stream) so the caller's
 * own CUDA events bracket it.  Call before any push. */
int links_set_stream(links_ctx* ctx, void* cuda_stream);
/* Forget the current job (filter, contigs, pair set, counters) but keep all device allocations, so
 * a long-lived process can run job after job without re-allocating. */
int links_reset(links_ctx* ctx);

/* Pinned host buffers for callers that want full-speed H2D (optional). */
void* links_host_alloc(size_t bytes);
void links_host_free(void* p);

/* ---- stage 1: Bloom filter.  Replaces LINKS::init_bloom_filter / make_bf (src/LINKS.hpp:284-352) ---- */

/* Sizing arithmetic of make_bf (src/LINKS.hpp:297-302) + BloomFilter's 8-byte round-up
 * (btllib bloom_filter.hpp:617-621); pure host helper.  m_bits is the "- m:" value it prints. */
int links_bf_size(uint64_t assembly_file_bytes, float fpr, uint64_t* m_bits, uint64_t* bytes, uint32_t* hash_num);
/* new KmerBloomFilter(bytes, hash_num, k): zeroed filter on the device; contigs_push will insert. */
int links_bf_alloc(links_ctx* ctx, uint64_t bytes, uint32_t hash_num);
/* new KmerBloomFilter(path) (-r): caller parsed the file; bits are uploaded, contigs_push will NOT insert. */
int links_bf_load(links_ctx* ctx, const uint8_t* bits, uint64_t bytes, uint32_t hash_num);
/* The device keeps, beside the filter, the exact set of canonical hashes this context inserted into it;
 * a read k-mer found there is known to pass BloomFilter::contains (no false negatives,
 * bloom_filter.hpp:639-659) and skips its hash_num probes.  After links_bf_load* the contigs are not
 * inserted, so that shortcut is off unless the caller vouches that the loaded bits were built from (a
 * superset of) the contigs about to be pushed: yes = 1, before the first links_contigs_push.
 * Results are identical either way; wrongly vouching is the caller's error. */
int links_bf_covers_contigs(links_ctx* ctx, int yes);
/* Raw bit array for KmerBloomFilter::save (bloom_filter.hpp:866-888); dst holds `bytes`. */
int links_bf_download(links_ctx* ctx, uint8_t* dst);
/* Exact header text of the .bloom file; returns its length (0 if cap too small). */
size_t links_bf_header(char* dst, size_t cap, uint32_t k, uint32_t hash_num, uint64_t bytes);

/* ---- stage 3 (and the insert half of stage 1): contigs.
 * Replaces the SeqReader loop of make_bf (src/LINKS.hpp:318-324) and
 * start_read_contig / populate_mate_info / write_from_block_to_set (src/LINKS.hpp:379-402,487-513,569-601,640-680).
 * Record i of this call is contig id first_tig_id + i (1-based ordinal over ALL records, :657-659). */
int links_contigs_push(links_ctx* ctx, const char* seq, const uint64_t* off, uint32_t n, uint32_t first_tig_id);
/* Same, with the concatenated ASCII already resident in device memory (offsets stay on the host). */
int links_contigs_push_dev(links_ctx* ctx, const void* dev_seq, const uint64_t* off, uint32_t n, uint32_t first_tig_id);

/* The contig k-mer table (LINKS::track_all_test, src/LINKS.hpp:379-402) comes in two builds with identical results:
 *   FULL   every contig k-mer, built on a side stream while the reads are processed (64 bytes per contig position);
 *   MATES  only the hashes that occur in a distinct pair -- the reference's own rule, populate_mate_info stores a contig
 *          k-mer only `if (mates.find(hash) != mates.end())` (src/LINKS.hpp:580,591) -- built by links_finalize once the
 *          pairs are known (at most ~50 bytes per distinct pair, whatever the assembly's size; with an attached
 *          exchange: per distinct pair this rank OWNS).
 * AUTO (default; LINKS_TABLE_MODE=auto|full|mates overrides at links_create) takes FULL while that table is under an
 * eighth of the device's memory.  Call before the job's first links_reads_push; the choice survives links_reset. */
#define LINKS_TABLE_AUTO 0
#define LINKS_TABLE_FULL 1
#define LINKS_TABLE_MATES 2
int links_table_mode(links_ctx* ctx, int mode);

/* ---- stage 2: reads.  Replaces start_read_fasta / extract_mate_pair / write_from_block_to_map
 * (src/LINKS.hpp:354-378,470-486,515-567,602-638).  first_read_idx = ordinal of record 0 over
 * all files of the .fof (defines first-wins order when n_dist > 1). */
int links_reads_push(links_ctx* ctx, const char* seq, const uint64_t* off, uint32_t n, uint64_t first_read_idx);
int links_reads_push_dev(links_ctx* ctx, const void* dev_seq, const uint64_t* off, uint32_t n, uint64_t first_read_idx);

/* ---- stage 4.  Replaces LINKS::pair_contigs (src/LINKS.hpp:780-1041).
 * rows: sorted by (bin, tig_a, rev_a, tig_b, rev_b), each key once.  issues: sorted.
 * Arrays are malloc'd by the library; release with links_free_result. */
int links_finalize(links_ctx* ctx, links_row** rows, uint64_t* n_rows, links_issue** issues,
                   uint64_t* n_issues, links_stats* stats);
void links_free_result(links_row* rows, links_issue* issues);

/* Formatting exactly as the reference prints (src/LINKS.hpp:1032-1035 and :915-927). */
int links_format_row(const links_row* r, char* dst, size_t cap);
int links_format_issue(const links_issue* r, char* dst, size_t cap);

/* ---- multi-GPU (reads sharded over ranks, contigs replicated).  The reference has no
 * counterpart: its mate_pair map (src/LINKS.hpp:138-147,354-378) lives in one address space.
 * After every rank has pushed its share of the reads, each distinct ordered (h_a,h_b) must end
 * up on exactly ONE rank before link counting (set semantics, first distance wins):
 *   export : the rank's distinct pairs as 24-byte records {h_a, h_b, order_key}, grouped by owner
 *            rank (a hash of (h_a,h_b)); counts[r] = records owned by rank r
 *   clear  : empty this rank's pair set (allocations kept)
 *   import : insert received records (device memory) into the pair set, minimum order key wins
 * then links_finalize per rank and sum the ranks' rows by key (rows of different ranks overlap).
 * The transport between export and import is the caller's: one all-to-all (NCCL in bench.py).
 * Global read ordinals (first_read_idx) must be disjoint across ranks when n_dist > 1. */
#define LINKS_PAIR_RECORD_BYTES 24u
#define LINKS_MAX_RANKS 256u
/* dev_dst == NULL: only fill counts[0..n_ranks).  Otherwise dev_dst (device memory, room for
 * cap_records records) receives the records, bucket r starting at sum(counts[0..r)). */
int links_pairs_export(links_ctx* ctx, uint32_t n_ranks, void* dev_dst, uint64_t cap_records, uint64_t* counts);
int links_pairs_import(links_ctx* ctx, const void* dev_records, uint64_t n_records);
int links_pairs_clear(links_ctx* ctx);
/* Sum rows that share (bin, tig_a, rev_a, tig_b, rev_b) -- the ranks' partial link tables after a
 * gather -- on the device; *out is malloc'd (links_free_result(*out, NULL)), sorted like links_finalize's. */
int links_rows_merge(links_ctx* ctx, const links_row* rows, uint64_t n, links_row** out, uint64_t* n_out);
/* Owner rank of a pair, the same function the export kernel uses (for hosts that route themselves). */
uint32_t links_pair_owner(uint64_t h_a, uint64_t h_b, uint32_t n_ranks);
/* Bloom filter replication: copy the filter's bit array device-to-device out of / into the
 * context, so a rank that built it can broadcast it (NCCL) and the others adopt it instead of
 * rebuilding.  links_bf_load_dev == links_bf_load with a device source. */
int links_bf_download_dev(links_ctx* ctx, void* dev_dst);
int links_bf_load_dev(links_ctx* ctx, const void* dev_bits, uint64_t bytes, uint32_t hash_num);

/* Bloom filter built in shares (large assemblies).  A filter of GBs is DRAM-resident and its build -- hash_num random
 * read-modify-writes per contig k-mer -- would be repeated in full on every rank.  Instead every rank owns a
 * peer-addressable buffer of `bytes` (bloom[r] = rank r's, as addressable from THIS process; same bytes / hash_num on all
 * ranks), inserts only its 1/n_ranks share of the k-mers of the contigs it is given (all ranks push ALL contigs, they stay
 * resident for stage 3), and links_bf_allreduce makes every copy the bitwise OR of all of them through peer memory:
 *   every rank: links_bf_alloc_shared -> links_contigs_push* ... -> links_sync;   barrier (caller's: all shares inserted)
 *   every rank: links_bf_allreduce (returns when this rank's slice is stored in every copy);   barrier;   links_reads_push* ...
 * The result is bit-identical to links_bf_alloc + the same contigs on one rank.  The exact member set
 * (links_bf_covers_contigs) is not kept in this mode. */
int links_bf_alloc_shared(links_ctx* ctx, uint64_t bytes, uint32_t hash_num, uint32_t rank, uint32_t n_ranks, void* const* bloom);
int links_bf_allreduce(links_ctx* ctx);

/* ---- multi-GPU over peer memory (NVLink P2P / NVSwitch): the exchange above without the stop in the middle.
 * Every rank owns an INBOX, a device buffer of inbox_bytes that all ranks of the node can address (CUDA
 * IPC, VMM export or torch symmetric memory -- the caller's business); inbox[r] is rank r's buffer as
 * addressable from THIS process.  Once attached (before the first job, on every rank, followed by any
 * barrier), links_reads_push routes the pairs each extract launch found new straight into their owners'
 * inboxes and inserts what has arrived here while the next chunk is extracted; links_exchange_finish
 * (after the last reads_push, on every rank, does not block the host) publishes "done", and
 * links_finalize then waits for the other ranks, classifies the pairs this rank OWNS, sends its partial
 * link rows to rank 0 and -- on rank 0 -- returns the summed table (other ranks return 0 rows).
 * All ranks must run the same sequence of jobs (links_reset advances a shared epoch).  A region that is
 * too small for a job fails that job with LINKS_E_LIMIT; size it for the distinct pairs of one rank
 * (bench.py: 24 bytes x 2 x expected pairs). */
int links_exchange_attach(links_ctx* ctx, uint32_t rank, uint32_t n_ranks, void* const* inbox, uint64_t inbox_bytes);
int links_exchange_finish(links_ctx* ctx);

/* ---- inspection utilities (used by the parity tests; thin wrappers over the same kernels) ---- */
/* fwd/rev ntHash2 of every k-mer start of one sequence (btllib NtHash, nthash.hpp:1265-1338);
 * valid[i] = 1 iff the k-mer at i has only ACGTU/acgtu.  Arrays hold max(len-k+1,0) entries. */
int links_kmer_hashes(links_ctx* ctx, const char* seq, uint64_t len, uint64_t* fwd, uint64_t* rev, uint8_t* valid);
/* links_kmer_hashes for callers that may not allocate device memory (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same results, same kernels; every intermediate (staged sequence, 2-bit packing, invalid-base mask,
 * result staging) lives in one static __device__ array of ~2.4 MB that each GPU loading the library holds until the
 * process exits, so the call makes no cudaMalloc / cudaFree.  len > LINKS_STATIC_MAX_BASES fails with LINKS_E_LIMIT.
 * All calls share that array, and it keeps the last call's contents (the call re-zeroes what its kernels expect zeroed).
 * Two calls that overlapped on the GPU would RACE on it: the library therefore serialises the calls of one process (a
 * mutex held until the call's stream has drained); different processes on one GPU each have their own copy. */
#define LINKS_STATIC_MAX_BASES (1u << 17)
int links_kmer_hashes_static(links_ctx* ctx, const char* seq, uint64_t len, uint64_t* fwd, uint64_t* rev, uint8_t* valid);
/* KmerBloomFilter membership of given (fwd,rev) hash pairs against the context's filter. */
int links_bf_contains(links_ctx* ctx, const uint64_t* fwd, const uint64_t* rev, uint64_t n, uint8_t* out);
/* Distinct pairs currently held: h_a[i], h_b[i], d_idx[i]; call with NULL arrays to get the count. */
int links_pairs_download(links_ctx* ctx, uint64_t* h_a, uint64_t* h_b, uint32_t* d_idx, uint64_t cap, uint64_t* n);
/* Device time (ms) and launch counts accumulated per kernel family since create / last reset. */
typedef struct links_timing {
  float ms_h2d, ms_pack, ms_bf_build, ms_table_build, ms_extract, ms_classify, ms_sort_reduce, ms_d2h, ms_grow, ms_exchange;
  uint64_t launches;         /* all kernels launched by this library */
  uint64_t extract_launches; /* launches of the dominant kernel (extract_kernel) */
} links_timing;
int links_get_timing(links_ctx* ctx, links_timing* t, int reset);

#ifdef __cplusplus
}
#endif
#endif

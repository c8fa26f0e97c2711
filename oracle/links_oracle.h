/* oracle/links_oracle.h -- CPU restatement of the LINKS hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this.  The product (links_b200/csrc, include/links_gpu.h) never links or calls it.
 *
 * PARITY: PINNED.  tests/test_oracle_*.py check this restatement against
 *   - btllib's ntHash known-answer test (reference btllib/tests/nthash.cpp:10-28),
 *   - the reference installation-test goldens (.bloom bytes, .tigpair_checkpoint.tsv,
 *     .pairing_issues; reference tests/expected_outputs/),
 *   - outputs of the unmodified reference binary oracle/_ref/LINKS_CPP on adversarial inputs
 *     (fixtures under tests/golden/, generator tools/make_golden.py).
 *
 * All file:line citations are into /root/reference.
 */
#ifndef LINKS_ORACLE_H
#define LINKS_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- ntHash2 primitives (btllib/include/btllib/nthash.hpp) ---- */

/* 1 iff c hashes to a non-zero seed: ACGTU acgtu (nthash.hpp:38-71).
 * Domain note: the reference table also gives seeds to the control bytes 1,3,4,5,7 (a side
 * effect of its `c & 7` complement trick); no sequence reader produces them and this restatement
 * (and the GPU path) treats them as invalid like any other non-ACGTU byte. */
int lo_base_valid(unsigned char c);

/* fwd / rev ntHash of one k-mer of valid bases, computed from scratch (nthash.hpp:897-930 base loop). */
void lo_nthash_kmer(const char* kmer, unsigned k, uint64_t* fwd, uint64_t* rev);

/* hashes()[0..m): h0 = min(fwd,rev); h_i = h0 * (i ^ k*MULTISEED); h_i ^= h_i >> 27
 * (nthash.hpp:21-27,922-928). */
void lo_multihash(uint64_t fwd, uint64_t rev, unsigned k, unsigned m, uint64_t* out);

/* Iterator restating btllib::NtHash init()/roll() including the skip rules for invalid bases
 * (nthash.hpp:1392-1410,1493-1550). */
typedef struct {
  const char* seq;
  uint64_t len;
  unsigned k;
  uint64_t pos;
  int initialized;
  uint64_t fwd, rev;
} lo_nthash;
void lo_nthash_begin(lo_nthash* it, const char* seq, uint64_t len, unsigned k, uint64_t pos);
int lo_nthash_roll(lo_nthash* it);

/* Walk (for tests): mirrors btl_probe.cpp:btl_nthash_walk. */
uint64_t lo_nthash_walk(const char* seq, uint64_t len, unsigned hash_num, unsigned k, uint64_t start,
                        uint64_t cap, uint64_t* pos, uint64_t* fwd, uint64_t* rev, uint64_t* hashes);

/* ---- Bloom filter (btllib/include/btllib/bloom_filter.hpp, src/LINKS.hpp:284-331) ---- */

/* Sizing from the assembly FILE SIZE in bytes and fpr, exactly as LINKS::make_bf
 * (src/LINKS.hpp:297-302) followed by BloomFilter's round-up to 8 bytes (bloom_filter.hpp:617-621).
 * m_bits_out gets the "- m:" value the reference prints. */
void lo_bf_size(uint64_t file_bytes, float fpr, uint64_t* m_bits_out, uint64_t* bytes_out,
                unsigned* hash_num_out);
void lo_bf_insert_seq(uint8_t* bits, uint64_t bytes, unsigned hash_num, unsigned k, const char* seq,
                      uint64_t len);
int lo_bf_contains(const uint8_t* bits, uint64_t bytes, unsigned hash_num, unsigned k, uint64_t fwd,
                   uint64_t rev);
/* Header text of KmerBloomFilter::save (bloom_filter.hpp:766-783,866-888): returns length written. */
size_t lo_bf_header(char* dst, size_t cap, unsigned k, unsigned hash_num, uint64_t bytes);

/* ---- whole hot path ---- */

typedef struct {
  unsigned k;
  unsigned n_dist;
  const uint32_t* dist;  /* -d list */
  const uint16_t* step;  /* -t list, already padded to n_dist (src/InputParser.hpp:224-229) */
  uint64_t offset;       /* -o */
  float insert_stdev;    /* -e */
} lo_params;

/* One aggregated link row (src/LINKS.hpp:1023-1038): "bin \t [fr]tigA \t [fr]tigB \t links \t gaps".
 * tig == 0 is the reference's empty tig id "" (default-inserted absent k-mer, :827-830). */
typedef struct {
  int32_t bin;
  uint32_t tig_a, tig_b;
  uint8_t rev_a, rev_b; /* 0 = 'f', 1 = 'r' */
  uint64_t links;
  int64_t gaps;
} lo_row;

/* One ".pairing_issues" record (src/LINKS.hpp:915-949). */
typedef struct {
  uint32_t tig;
  uint64_t tig_len;
  uint64_t s1, e1, s2, e2;
  uint8_t o1, o2;
  uint64_t pet_size;
  int64_t insert_size;
} lo_issue;

/* One Bloom-surviving observation in -j 1 processing order (src/LINKS.hpp:546-558). */
typedef struct {
  uint64_t h_a, h_b;
  uint32_t d_idx;
  uint32_t read_idx;
  uint64_t rank; /* n: index among valid k-mers of stream A from the offset */
} lo_obs;

typedef struct {
  uint64_t n_candidates;   /* (A,B) pairs visited before the Bloom test */
  uint64_t n_survivors;    /* passed both Bloom tests */
  uint64_t n_unique_pairs; /* distinct ordered (h_a,h_b) */
  uint64_t n_mates;
  uint64_t n_tracked;      /* entries in track_all after the contig pass */
  uint64_t ct_ok_pairs, ct_problem_pairs, ct_ok_contig, ct_iz_issues, ct_illogical;
  uint64_t ct_gate_fail;   /* pairs with a multiplicity>1 k-mer (reference: ct_single+ct_multiple) */
} lo_stats;

typedef struct lo_result {
  lo_row* rows;     uint64_t n_rows;     /* sorted by (bin, tig_a, rev_a, tig_b, rev_b) */
  lo_issue* issues; uint64_t n_issues;   /* sorted by all fields */
  lo_obs* obs;      uint64_t n_obs;      /* only if keep_obs */
  lo_stats stats;
} lo_result;

/* Run stages 2-4 of LINKS_CPP (src/main.cpp:39-64) given a filled Bloom filter.
 * contigs / reads: concatenated ASCII with n+1 offsets.  Contig ids are 1-based ordinals
 * counting every record (src/LINKS.hpp:657-659).  Returns NULL on allocation failure. */
lo_result* lo_run(const lo_params* p, const uint8_t* bf_bits, uint64_t bf_bytes, unsigned hash_num,
                  const char* contigs, const uint64_t* contig_off, uint64_t n_contigs,
                  const char* reads, const uint64_t* read_off, uint64_t n_reads, int keep_obs);
/* Stages 3-4 only, on a given set of distinct ordered pairs (h_a, h_b, winning distance index):
 * the share of pair_contigs (src/LINKS.hpp:780-1041) one rank of the multi-GPU path owns. */
lo_result* lo_run_from_pairs(const lo_params* p, const char* contigs, const uint64_t* contig_off, uint64_t n_contigs,
                             const uint64_t* h_a, const uint64_t* h_b, const uint32_t* d_idx, uint64_t n_pairs);
void lo_free(lo_result* r);

/* Formatting helpers: write one row / issue exactly as the reference prints them. */
int lo_format_row(const lo_row* r, char* dst, size_t cap);
int lo_format_issue(const lo_issue* r, char* dst, size_t cap);

#ifdef __cplusplus
}
#endif
#endif

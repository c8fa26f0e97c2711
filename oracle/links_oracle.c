/* oracle/links_oracle.c -- CPU restatement of the LINKS hot path in plain C.
 * TEST INFRASTRUCTURE ONLY; see links_oracle.h for the parity pins.  It follows the reference's
 * own control flow (two rolling iterators per distance, first-wins maps, default-insert gate)
 * rather than the GPU path's rank/select reformulation, so that agreement between the two is a
 * real check.  Citations are into /root/reference.
 */
#define _POSIX_C_SOURCE 200809L
#include "links_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ ntHash2 */

/* btllib/include/btllib/nthash.hpp:21-34 */
#define LO_MULTISHIFT 27
static const uint64_t LO_MULTISEED = 0x90b45d39fb6da1faULL;
static const uint64_t LO_SEED_A = 0x3c8bfbb395c60474ULL;
static const uint64_t LO_SEED_C = 0x3193c18562a02b4cULL;
static const uint64_t LO_SEED_G = 0x20323ed082572324ULL;
static const uint64_t LO_SEED_T = 0x295549f54be24456ULL;

static uint64_t seed_fwd(unsigned char c)
{
  switch (c) {
  case 'A': case 'a': return LO_SEED_A;
  case 'C': case 'c': return LO_SEED_C;
  case 'G': case 'g': return LO_SEED_G;
  case 'T': case 't': case 'U': case 'u': return LO_SEED_T;
  default: return 0;
  }
}
/* seed of the complement base: SEED_TAB[c & CP_OFF] (nthash.hpp:21,38-40) */
static uint64_t seed_cmp(unsigned char c)
{
  switch (c) {
  case 'A': case 'a': return LO_SEED_T;
  case 'C': case 'c': return LO_SEED_G;
  case 'G': case 'g': return LO_SEED_C;
  case 'T': case 't': case 'U': case 'u': return LO_SEED_A;
  default: return 0;
  }
}

int lo_base_valid(unsigned char c) { return seed_fwd(c) != 0; }

/* rol1 + swapbits033 (nthash.hpp:443-447,480-486): rotates the low 33 bits and the high 31 bits
 * left by one, independently. */
static uint64_t srol(uint64_t v)
{
  v = (v << 1) | (v >> 63);
  uint64_t x = (v ^ (v >> 33)) & 1;
  return v ^ (x | (x << 33));
}
/* ror1 + swapbits3263 (nthash.hpp:457-461,488-493): inverse of srol. */
static uint64_t sror(uint64_t v)
{
  v = (v >> 1) | (v << 63);
  uint64_t x = ((v >> 32) ^ (v >> 63)) & 1;
  return v ^ ((x << 32) | (x << 63));
}
/* srol applied n times: what MS_TAB_31L[c][n%31] | MS_TAB_33R[c][n%33] tabulate (nthash.hpp:576-603). */
static uint64_t srol_n(uint64_t v, unsigned n)
{
  uint64_t hi = v >> 33, lo = v & 0x1FFFFFFFFULL;
  unsigned a = n % 31, b = n % 33;
  if (a) hi = ((hi << a) | (hi >> (31 - a))) & 0x7FFFFFFFULL;
  if (b) lo = ((lo << b) | (lo >> (33 - b))) & 0x1FFFFFFFFULL;
  return (hi << 33) | lo;
}

void lo_nthash_kmer(const char* kmer, unsigned k, uint64_t* fwd, uint64_t* rev)
{
  uint64_t f = 0, r = 0;
  for (unsigned i = 0; i < k; ++i) { /* nthash.hpp:907-919 */
    f = srol(f) ^ seed_fwd((unsigned char)kmer[i]);
    r = srol(r) ^ seed_cmp((unsigned char)kmer[k - 1 - i]);
  }
  *fwd = f;
  *rev = r;
}

void lo_multihash(uint64_t fwd, uint64_t rev, unsigned k, unsigned m, uint64_t* out)
{
  uint64_t b = rev < fwd ? rev : fwd; /* nthash.hpp:920-921 */
  if (m) out[0] = b;
  for (unsigned i = 1; i < m; ++i) { /* nthash.hpp:922-927 */
    uint64_t t = b * ((uint64_t)i ^ ((uint64_t)k * LO_MULTISEED));
    t ^= t >> LO_MULTISHIFT;
    out[i] = t;
  }
}

void lo_nthash_begin(lo_nthash* it, const char* seq, uint64_t len, unsigned k, uint64_t pos)
{
  it->seq = seq; it->len = len; it->k = k; it->pos = pos; it->initialized = 0; it->fwd = it->rev = 0;
}

/* BTLLIB_NTHASH_INIT (nthash.hpp:1493-1513): from pos, find the first window without an invalid
 * base; on failure the base-hash routine reports the LAST invalid index in the window (it scans
 * from k-1 down, :905-910) and init skips just past it. */
static int nthash_init(lo_nthash* it)
{
  if (it->k > it->len) { it->pos = UINT64_MAX; return 0; }
  while (it->pos < it->len - it->k + 1) {
    int bad = -1;
    for (int i = (int)it->k - 1; i >= 0; --i)
      if (!lo_base_valid((unsigned char)it->seq[it->pos + (uint64_t)i])) { bad = i; break; }
    if (bad < 0) {
      lo_nthash_kmer(it->seq + it->pos, it->k, &it->fwd, &it->rev);
      break;
    }
    it->pos += (uint64_t)bad + 1;
  }
  if (it->pos > it->len - it->k) { it->pos = UINT64_MAX; return 0; }
  it->initialized = 1;
  return 1;
}

/* BTLLIB_NTHASH_ROLL (nthash.hpp:1516-1533) with the rolling updates of :576-603. */
int lo_nthash_roll(lo_nthash* it)
{
  if (!it->initialized) return nthash_init(it);
  ++it->pos;
  if (it->pos > it->len - it->k) return 0;
  unsigned char in = (unsigned char)it->seq[it->pos + it->k - 1];
  if (!lo_base_valid(in)) {
    it->pos += it->k;
    return nthash_init(it);
  }
  unsigned char out = (unsigned char)it->seq[it->pos - 1];
  it->fwd = srol(it->fwd) ^ seed_fwd(in) ^ srol_n(seed_fwd(out), it->k);
  it->rev = sror(it->rev ^ srol_n(seed_cmp(in), it->k) ^ seed_cmp(out));
  return 1;
}

uint64_t lo_nthash_walk(const char* seq, uint64_t len, unsigned hash_num, unsigned k, uint64_t start,
                        uint64_t cap, uint64_t* pos, uint64_t* fwd, uint64_t* rev, uint64_t* hashes)
{
  lo_nthash it;
  lo_nthash_begin(&it, seq, len, k, start);
  uint64_t n = 0;
  while (n < cap && lo_nthash_roll(&it)) {
    pos[n] = it.pos; fwd[n] = it.fwd; rev[n] = it.rev;
    if (hashes) lo_multihash(it.fwd, it.rev, k, hash_num, hashes + n * hash_num);
    ++n;
  }
  return n;
}

/* ------------------------------------------------------------------ Bloom filter */

void lo_bf_size(uint64_t file_bytes, float fpr, uint64_t* m_bits_out, uint64_t* bytes_out,
                unsigned* hash_num_out)
{
  /* src/LINKS.hpp:297-302.  `log(fpr)` there is the double libm log of the float promoted to
   * double (global ::log from <math.h>), checked against the reference binary's "- m:" output. */
  uint64_t m = (uint64_t)ceil((-1 * (double)file_bytes * log((double)fpr)) / (log(2) * log(2)));
  m = ((uint64_t)(m / 8) + 1) * 8;
  unsigned hash_fct = (unsigned)floor(((double)m / (double)file_bytes) * log(2));
  uint64_t bytes = m / 8;
  /* bloom_filter.hpp:617-621: round the byte budget up to a multiple of sizeof(uint64_t) */
  bytes = (uint64_t)(ceil((double)bytes / 8.0) * 8.0);
  if (m_bits_out) *m_bits_out = m;
  if (bytes_out) *bytes_out = bytes;
  if (hash_num_out) *hash_num_out = hash_fct;
}

static void bf_insert(uint8_t* bits, uint64_t array_bits, const uint64_t* h, unsigned m)
{
  for (unsigned i = 0; i < m; ++i) { /* bloom_filter.hpp:639-646 */
    uint64_t n = h[i] % array_bits;
    bits[n >> 3] |= (uint8_t)(1u << (n & 7));
  }
}
static int bf_contains(const uint8_t* bits, uint64_t array_bits, const uint64_t* h, unsigned m)
{
  for (unsigned i = 0; i < m; ++i) { /* bloom_filter.hpp:648-659 */
    uint64_t n = h[i] % array_bits;
    if (!(bits[n >> 3] & (uint8_t)(1u << (n & 7)))) return 0;
  }
  return 1;
}

#define LO_MAX_HASH 1024 /* bloom_filter.hpp MAX_HASH_VALUES */

void lo_bf_insert_seq(uint8_t* bits, uint64_t bytes, unsigned hash_num, unsigned k, const char* seq,
                      uint64_t len)
{
  /* KmerBloomFilter::insert (bloom_filter.hpp:814-821) */
  uint64_t h[LO_MAX_HASH];
  lo_nthash it;
  lo_nthash_begin(&it, seq, len, k, 0);
  while (lo_nthash_roll(&it)) {
    lo_multihash(it.fwd, it.rev, k, hash_num, h);
    bf_insert(bits, bytes * 8, h, hash_num);
  }
}

int lo_bf_contains(const uint8_t* bits, uint64_t bytes, unsigned hash_num, unsigned k, uint64_t fwd,
                   uint64_t rev)
{
  uint64_t h[LO_MAX_HASH];
  lo_multihash(fwd, rev, k, hash_num, h);
  return bf_contains(bits, bytes * 8, h, hash_num);
}

size_t lo_bf_header(char* dst, size_t cap, unsigned k, unsigned hash_num, uint64_t bytes)
{
  /* cpptoml emits the table keys sorted... the reference's golden header
   * (tests/expected_outputs/installation_test.bloom) reads, with TAB indentation:
   *   [BTLKmerBloomFilter_v6]\n\tk = 15\n\thash_fn = "ntHash_v1"\n\thash_num = 9\n\tbytes = N\n
   * followed by [HeaderEnd]\n and 50 newlines with "  <binary data>" before the 2nd
   * (bloom_filter.hpp:766-783). */
  int n = snprintf(dst, cap,
                   "[BTLKmerBloomFilter_v6]\n\tk = %u\n\thash_fn = \"ntHash_v1\"\n\thash_num = %u\n"
                   "\tbytes = %llu\n[HeaderEnd]\n\n  <binary data>\n",
                   k, hash_num, (unsigned long long)bytes);
  if (n < 0 || (size_t)n + 48 >= cap) return 0;
  for (int i = 0; i < 48; ++i) dst[n++] = '\n';
  dst[n] = 0;
  return (size_t)n;
}

/* ------------------------------------------------------------------ small hash maps */

static uint64_t mix64(uint64_t x)
{
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

/* u64 -> index map with explicit occupancy flags (keys may take any value, 0 included) */
typedef struct { uint64_t* key; uint32_t* val; uint8_t* used; uint64_t cap, n; } u64map;

static int u64map_init(u64map* m, uint64_t cap_pow2)
{
  m->cap = cap_pow2; m->n = 0;
  m->key = (uint64_t*)malloc(cap_pow2 * sizeof(uint64_t));
  m->val = (uint32_t*)malloc(cap_pow2 * sizeof(uint32_t));
  m->used = (uint8_t*)calloc(cap_pow2, 1);
  return m->key && m->val && m->used;
}
static void u64map_free(u64map* m) { free(m->key); free(m->val); free(m->used); }
static int u64map_grow(u64map* m);
/* returns slot; *found tells whether the key was present.  Inserts (val untouched) if absent. */
static uint64_t u64map_upsert(u64map* m, uint64_t k, int* found)
{
  if ((m->n + 1) * 10 > m->cap * 7) u64map_grow(m);
  uint64_t i = mix64(k) & (m->cap - 1);
  while (m->used[i]) {
    if (m->key[i] == k) { *found = 1; return i; }
    i = (i + 1) & (m->cap - 1);
  }
  m->used[i] = 1; m->key[i] = k; m->n++;
  *found = 0;
  return i;
}
static int64_t u64map_find(const u64map* m, uint64_t k)
{
  uint64_t i = mix64(k) & (m->cap - 1);
  while (m->used[i]) {
    if (m->key[i] == k) return (int64_t)i;
    i = (i + 1) & (m->cap - 1);
  }
  return -1;
}
static int u64map_grow(u64map* m)
{
  u64map n;
  if (!u64map_init(&n, m->cap * 2)) abort();
  for (uint64_t i = 0; i < m->cap; ++i)
    if (m->used[i]) { int f; uint64_t s = u64map_upsert(&n, m->key[i], &f); n.val[s] = m->val[i]; }
  u64map_free(m);
  *m = n;
  return 1;
}

/* (a,b) -> d_idx, first wins (mate_pair, src/LINKS.hpp:138-147,370-373) */
typedef struct { uint64_t a, b; uint32_t d_idx; uint8_t used; } pairslot;
typedef struct { pairslot* s; uint64_t cap, n; } pairmap;

static int pairmap_init(pairmap* m, uint64_t cap) { m->cap = cap; m->n = 0; m->s = (pairslot*)calloc(cap, sizeof(pairslot)); return m->s != NULL; }
static int pairmap_insert(pairmap* m, uint64_t a, uint64_t b, uint32_t d_idx);
static void pairmap_grow(pairmap* m)
{
  pairmap n;
  if (!pairmap_init(&n, m->cap * 2)) abort();
  for (uint64_t i = 0; i < m->cap; ++i)
    if (m->s[i].used) pairmap_insert(&n, m->s[i].a, m->s[i].b, m->s[i].d_idx);
  free(m->s);
  *m = n;
}
/* returns 1 if newly inserted */
static int pairmap_insert(pairmap* m, uint64_t a, uint64_t b, uint32_t d_idx)
{
  if ((m->n + 1) * 10 > m->cap * 7) pairmap_grow(m);
  uint64_t i = mix64(a ^ mix64(b)) & (m->cap - 1);
  while (m->s[i].used) {
    if (m->s[i].a == a && m->s[i].b == b) return 0;
    i = (i + 1) & (m->cap - 1);
  }
  m->s[i].a = a; m->s[i].b = b; m->s[i].d_idx = d_idx; m->s[i].used = 1; m->n++;
  return 1;
}

/* ------------------------------------------------------------------ stage 4 helpers */

typedef struct { uint32_t tig; uint64_t start, end, multiple; uint8_t orient; } kmerinfo; /* LINKS.hpp:123-136 */

/* LINKS::get_distance (src/LINKS.hpp:772-779), same integer types and truncations */
static int get_distance(uint64_t insert_size, uint64_t length_i, uint64_t start_i, uint64_t start_j)
{
  int insert_span = (int)(uint32_t)((length_i - start_i) + start_j);
  int gap_or_overlap = (int)(uint32_t)(insert_size - (uint64_t)(int64_t)insert_span);
  return gap_or_overlap;
}
/* LINKS::get_distance_bin (src/LINKS.hpp:764-770) */
static int get_distance_bin(int distance)
{
  return distance < 0 ? -1 : distance == 10 ? 10 : distance < 500 ? 500 : distance < 5000 ? 5000 : 10000;
}

typedef struct { lo_row* r; uint64_t n, cap; u64map idx; } rowagg;

static uint64_t row_key_hash(int32_t bin, uint32_t ta, uint8_t ra, uint32_t tb, uint8_t rb)
{
  uint64_t x = ((uint64_t)ta << 32) | tb;
  uint64_t y = ((uint64_t)(uint32_t)bin << 2) | ((uint64_t)ra << 1) | rb;
  return mix64(x) ^ mix64(y * 0x9e3779b97f4a7c15ULL + 1);
}
static void row_add(rowagg* g, int32_t bin, uint32_t ta, uint8_t ra, uint32_t tb, uint8_t rb, int distance)
{
  /* pair[tigA][bin][tigB] += (1, distance)  (src/LINKS.hpp:738-761).  The u64map is keyed by a
   * hash of the row key; collisions are resolved by chaining through a linear scan of equal-hash
   * rows, so the key itself is always compared in full. */
  uint64_t h = row_key_hash(bin, ta, ra, tb, rb);
  for (;; ++h) {
    int found;
    uint64_t s = u64map_upsert(&g->idx, h, &found);
    if (!found) {
      if (g->n == g->cap) { g->cap = g->cap ? g->cap * 2 : 1024; g->r = (lo_row*)realloc(g->r, g->cap * sizeof(lo_row)); if (!g->r) abort(); }
      g->idx.val[s] = (uint32_t)g->n;
      lo_row* r = &g->r[g->n++];
      r->bin = bin; r->tig_a = ta; r->tig_b = tb; r->rev_a = ra; r->rev_b = rb; r->links = 1; r->gaps = distance;
      return;
    }
    lo_row* r = &g->r[g->idx.val[s]];
    if (r->bin == bin && r->tig_a == ta && r->tig_b == tb && r->rev_a == ra && r->rev_b == rb) {
      r->links += 1; r->gaps += distance;
      return;
    }
  }
}

/* LINKS::add_to_pair_map (src/LINKS.hpp:699-762) */
static void add_to_pair_map(rowagg* g, int isz, int distance, uint32_t t1, uint32_t t2, unsigned orient_enum)
{
  switch (orient_enum) {
  case 0: row_add(g, isz, t1, 0, t2, 0, distance); row_add(g, isz, t2, 1, t1, 1, distance); break; /* fA fB | rB rA */
  case 1: row_add(g, isz, t1, 0, t2, 1, distance); row_add(g, isz, t2, 0, t1, 1, distance); break; /* fA rB | fB rA */
  case 2: row_add(g, isz, t1, 1, t2, 0, distance); row_add(g, isz, t2, 1, t1, 0, distance); break; /* rA fB | rB fA */
  case 3: row_add(g, isz, t1, 1, t2, 1, distance); row_add(g, isz, t2, 0, t1, 0, distance); break; /* rA rB | fB fA */
  default: break;
  }
}

static int row_cmp(const void* pa, const void* pb)
{
  const lo_row *a = (const lo_row*)pa, *b = (const lo_row*)pb;
  if (a->bin != b->bin) return a->bin < b->bin ? -1 : 1;
  if (a->tig_a != b->tig_a) return a->tig_a < b->tig_a ? -1 : 1;
  if (a->rev_a != b->rev_a) return a->rev_a < b->rev_a ? -1 : 1;
  if (a->tig_b != b->tig_b) return a->tig_b < b->tig_b ? -1 : 1;
  if (a->rev_b != b->rev_b) return a->rev_b < b->rev_b ? -1 : 1;
  return 0;
}
static int issue_cmp(const void* pa, const void* pb)
{
  const lo_issue *a = (const lo_issue*)pa, *b = (const lo_issue*)pb;
#define C(f) if (a->f != b->f) return a->f < b->f ? -1 : 1
  C(tig); C(s1); C(s2); C(e1); C(e2); C(o1); C(o2); C(pet_size); C(insert_size); C(tig_len);
#undef C
  return 0;
}

/* ------------------------------------------------------------------ whole path */

static lo_result* run_impl(const lo_params* p, const uint8_t* bf_bits, uint64_t bf_bytes, unsigned hash_num,
                           const char* contigs, const uint64_t* contig_off, uint64_t n_contigs,
                           const char* reads, const uint64_t* read_off, uint64_t n_reads, int keep_obs,
                           const uint64_t* preset_a, const uint64_t* preset_b, const uint32_t* preset_d, uint64_t n_preset);

lo_result* lo_run(const lo_params* p, const uint8_t* bf_bits, uint64_t bf_bytes, unsigned hash_num,
                  const char* contigs, const uint64_t* contig_off, uint64_t n_contigs,
                  const char* reads, const uint64_t* read_off, uint64_t n_reads, int keep_obs)
{
  return run_impl(p, bf_bits, bf_bytes, hash_num, contigs, contig_off, n_contigs, reads, read_off, n_reads, keep_obs,
                  NULL, NULL, NULL, 0);
}

/* Stages 3-4 on a GIVEN mate_pair content (distinct ordered pairs with their winning distance
 * index) instead of extracting it from reads: what one rank of the multi-GPU path computes after
 * the pair exchange.  Test infrastructure for tests/test_multi_rank.py. */
lo_result* lo_run_from_pairs(const lo_params* p, const char* contigs, const uint64_t* contig_off, uint64_t n_contigs,
                             const uint64_t* h_a, const uint64_t* h_b, const uint32_t* d_idx, uint64_t n_pairs)
{
  static const uint8_t no_bits[8] = {0};
  return run_impl(p, no_bits, 8, 1, contigs, contig_off, n_contigs, NULL, NULL, 0, 0, h_a, h_b, d_idx, n_pairs);
}

static lo_result* run_impl(const lo_params* p, const uint8_t* bf_bits, uint64_t bf_bytes, unsigned hash_num,
                           const char* contigs, const uint64_t* contig_off, uint64_t n_contigs,
                           const char* reads, const uint64_t* read_off, uint64_t n_reads, int keep_obs,
                           const uint64_t* preset_a, const uint64_t* preset_b, const uint32_t* preset_d, uint64_t n_preset)
{
  lo_result* res = (lo_result*)calloc(1, sizeof(lo_result));
  if (!res) return NULL;
  const unsigned k = p->k;
  const uint64_t array_bits = bf_bytes * 8;
  uint64_t hA[LO_MAX_HASH], hB[LO_MAX_HASH];
  if (hash_num > LO_MAX_HASH) { free(res); return NULL; }

  pairmap mate_pair;
  u64map mates; /* set: val unused */
  if (!pairmap_init(&mate_pair, 1u << 16) || !u64map_init(&mates, 1u << 16)) { free(res); return NULL; }
  uint64_t obs_cap = 0;

  /* ---- stage 2: start_read_fasta (src/LINKS.hpp:470-486) at -j 1:
   * read order -> distance order -> position order. */
  for (uint64_t r = 0; r < n_reads; ++r) {
    const char* seq = reads + read_off[r];
    const uint64_t len = read_off[r + 1] - read_off[r];
    if (!(k <= len)) continue; /* :619-623 */
    for (unsigned di = 0; di < p->n_dist; ++di) { /* extract_mate_pair :515-567 */
      const unsigned step = p->step[di];
      const uint64_t delta = p->dist[di];
      lo_nthash a, b;
      lo_nthash_begin(&a, seq, len, k, p->offset);
      lo_nthash_begin(&b, seq, len, k, delta + p->offset);
      uint64_t rolls = 0; /* successful rolls of stream A so far */
      /* NB: `ntHash.roll() && ntHash_lead.roll()` short-circuits, and the inner loop rolls the
       * LEAD first (`!ntHash_lead.roll() || !ntHash.roll()`); any failure ends this distance,
       * so the asymmetry is unobservable. */
      while (lo_nthash_roll(&a) && lo_nthash_roll(&b)) {
        ++rolls;
        int brk = 0;
        for (unsigned j = 1; j < step; ++j) {
          if (!lo_nthash_roll(&b) || !lo_nthash_roll(&a)) brk = 1;
          else ++rolls;
        }
        if (brk) break;
        res->stats.n_candidates++;
        uint64_t h_a, h_b;
        if (a.fwd < b.rev) { h_a = b.rev; h_b = a.rev; } /* :546-552 */
        else { h_a = a.fwd; h_b = b.fwd; }
        lo_multihash(a.fwd, a.rev, k, hash_num, hA);
        if (!bf_contains(bf_bits, array_bits, hA, hash_num)) continue; /* :554-555 */
        lo_multihash(b.fwd, b.rev, k, hash_num, hB);
        if (!bf_contains(bf_bits, array_bits, hB, hash_num)) continue;
        res->stats.n_survivors++;
        if (keep_obs) {
          if (res->n_obs == obs_cap) { obs_cap = obs_cap ? obs_cap * 2 : 4096; res->obs = (lo_obs*)realloc(res->obs, obs_cap * sizeof(lo_obs)); if (!res->obs) abort(); }
          lo_obs* o = &res->obs[res->n_obs++];
          o->h_a = h_a; o->h_b = h_b; o->d_idx = di; o->read_idx = (uint32_t)r; o->rank = rolls - 1;
        }
        /* write_from_block_to_map :354-378: first (h_a,h_b) wins; both hashes join `mates` */
        if (pairmap_insert(&mate_pair, h_a, h_b, di)) {
          int f;
          u64map_upsert(&mates, h_a, &f);
          u64map_upsert(&mates, h_b, &f);
        }
      }
    }
  }
  for (uint64_t i = 0; i < n_preset; ++i) {
    if (pairmap_insert(&mate_pair, preset_a[i], preset_b[i], preset_d[i])) {
      int f;
      u64map_upsert(&mates, preset_a[i], &f);
      u64map_upsert(&mates, preset_b[i], &f);
    }
  }
  res->stats.n_unique_pairs = mate_pair.n;
  res->stats.n_mates = mates.n;

  /* ---- stage 3: start_read_contig (src/LINKS.hpp:487-513,569-601,640-680,379-402) */
  u64map track; /* hash -> index into info[] */
  if (!u64map_init(&track, 1u << 16)) abort();
  kmerinfo* info = NULL; uint64_t n_info = 0, cap_info = 0;
  uint64_t* tig_length = (uint64_t*)calloc(n_contigs + 1, sizeof(uint64_t)); /* [0] is tig "" -> 0 */
  if (!tig_length) abort();
  for (uint64_t c = 0; c < n_contigs; ++c) {
    const char* seq = contigs + contig_off[c];
    const uint64_t len = contig_off[c + 1] - contig_off[c];
    if (!(k <= len)) continue; /* :655 */
    const uint32_t tig = (uint32_t)(c + 1); /* to_string(record.num + 1) :657-659 */
    tig_length[tig] = len;
    lo_nthash it;
    lo_nthash_begin(&it, seq, len, k, 0);
    while (lo_nthash_roll(&it)) {
      for (int strand = 0; strand < 2; ++strand) { /* :580-598 */
        const uint64_t h = strand ? it.rev : it.fwd;
        if (u64map_find(&mates, h) < 0) continue;
        int found;
        uint64_t s = u64map_upsert(&track, h, &found);
        if (!found) { /* :394-397 */
          if (n_info == cap_info) { cap_info = cap_info ? cap_info * 2 : 4096; info = (kmerinfo*)realloc(info, cap_info * sizeof(kmerinfo)); if (!info) abort(); }
          track.val[s] = (uint32_t)n_info;
          kmerinfo* ki = &info[n_info++];
          ki->tig = tig; ki->start = it.pos; ki->end = it.pos + k; ki->multiple = 1; ki->orient = (uint8_t)strand;
        } else {
          info[track.val[s]].multiple += 1; /* :399 */
        }
      }
    }
  }
  res->stats.n_tracked = track.n;

  /* ---- stage 4: pair_contigs (src/LINKS.hpp:780-1041) */
  rowagg agg; memset(&agg, 0, sizeof agg);
  if (!u64map_init(&agg.idx, 1u << 12)) abort();
  uint64_t issue_cap = 0;
  static const kmerinfo absent = {0, 0, 0, 1, 0}; /* default KmerInfo() via operator[] :827-830,123-136 */
  for (uint64_t i = 0; i < mate_pair.cap; ++i) {
    if (!mate_pair.s[i].used) continue;
    const uint64_t h1 = mate_pair.s[i].a, h2 = mate_pair.s[i].b;
    int64_t s1 = u64map_find(&track, h1), s2 = u64map_find(&track, h2);
    const kmerinfo* k1 = s1 < 0 ? &absent : &info[track.val[s1]];
    const kmerinfo* k2 = s2 < 0 ? &absent : &info[track.val[s2]];
    if (!(k1->multiple == 1 && k2->multiple == 1)) { res->stats.ct_gate_fail++; continue; }
    const int64_t insert_size = p->dist[mate_pair.s[i].d_idx];
    const int min_allowed = (int)(-1 * (p->insert_stdev * insert_size)); /* :838 float math, trunc */
    const unsigned low_iz = (unsigned)(insert_size + min_allowed);       /* :839 */
    const unsigned up_iz = (unsigned)(insert_size - min_allowed);        /* :840 */
    if (k1->tig != k2->tig) { /* :851 */
      int distance; unsigned oe;
      if (!k1->orient) {
        if (!k2->orient) { distance = get_distance((uint64_t)insert_size, tig_length[k1->tig], k1->start, k2->start); oe = 0; }
        else { distance = get_distance((uint64_t)insert_size, tig_length[k1->tig], k1->start, tig_length[k2->tig] - k2->end); oe = 1; }
      } else {
        if (!k2->orient) { distance = get_distance((uint64_t)insert_size, tig_length[k1->tig], tig_length[k1->tig] - k1->end, k2->start); oe = 2; }
        else { distance = get_distance((uint64_t)insert_size, tig_length[k2->tig], k2->end, k1->end); oe = 3; }
      }
      if (distance > min_allowed && (int64_t)distance < insert_size) {
        res->stats.ct_ok_pairs++;
        add_to_pair_map(&agg, get_distance_bin(distance), distance, k1->tig, k2->tig, oe);
      } else {
        res->stats.ct_problem_pairs++;
      }
    } else { /* same tig :900-954 */
      if (k1->orient == k2->orient) {
        uint64_t pet_size = k1->start > k2->start ? k1->start - k2->start : k2->start - k1->start;
        if (pet_size >= low_iz && pet_size <= up_iz) {
          res->stats.ct_ok_contig++;
        } else {
          res->stats.ct_iz_issues++;
          if (res->n_issues == issue_cap) { issue_cap = issue_cap ? issue_cap * 2 : 256; res->issues = (lo_issue*)realloc(res->issues, issue_cap * sizeof(lo_issue)); if (!res->issues) abort(); }
          lo_issue* is = &res->issues[res->n_issues++];
          memset(is, 0, sizeof *is);
          is->tig = k1->tig; is->tig_len = tig_length[k1->tig];
          is->s1 = k1->start; is->e1 = k1->end; is->o1 = k1->orient;
          is->s2 = k2->start; is->e2 = k2->end; is->o2 = k2->orient;
          is->pet_size = pet_size; is->insert_size = insert_size;
        }
      } else {
        res->stats.ct_illogical++;
      }
    }
  }

  qsort(agg.r, agg.n, sizeof(lo_row), row_cmp);
  if (res->n_issues) qsort(res->issues, res->n_issues, sizeof(lo_issue), issue_cmp);
  res->rows = agg.r; res->n_rows = agg.n;

  u64map_free(&agg.idx); u64map_free(&track); u64map_free(&mates);
  free(mate_pair.s); free(info); free(tig_length);
  return res;
}

void lo_free(lo_result* r)
{
  if (!r) return;
  free(r->rows); free(r->issues); free(r->obs); free(r);
}

int lo_format_row(const lo_row* r, char* dst, size_t cap)
{
  /* src/LINKS.hpp:1032-1035; tig 0 prints as the empty id */
  char a[16] = "", b[16] = "";
  if (r->tig_a) snprintf(a, sizeof a, "%u", r->tig_a);
  if (r->tig_b) snprintf(b, sizeof b, "%u", r->tig_b);
  return snprintf(dst, cap, "%d\t%c%s\t%c%s\t%llu\t%lld\n", r->bin, r->rev_a ? 'r' : 'f', a,
                  r->rev_b ? 'r' : 'f', b, (unsigned long long)r->links, (long long)r->gaps);
}

int lo_format_issue(const lo_issue* r, char* dst, size_t cap)
{
  /* src/LINKS.hpp:915-927 */
  char t[16] = "";
  if (r->tig) snprintf(t, sizeof t, "%u", r->tig);
  return snprintf(dst, cap,
                  "Pairs unsatisfied in distance within a contig.  Pairs on contig id: %s (length: %llu)\n"
                  "kmer start: %llu kmer end: %llu orientation: %d\n"
                  "kmer start: %llu kmer end: %llu orientation: %d\n"
                  "Calculated distance apart on contig: %llu - Distance on read: %lld\n",
                  t, (unsigned long long)r->tig_len, (unsigned long long)r->s1, (unsigned long long)r->e1,
                  (int)r->o1, (unsigned long long)r->s2, (unsigned long long)r->e2, (int)r->o2,
                  (unsigned long long)r->pet_size, (long long)r->insert_size);
}

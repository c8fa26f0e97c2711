// links_b200/csrc/kernels.cuh -- the hot-path kernels (sm_100a).
//
//   pack_kernel          ASCII -> 2-bit codes + invalid-base bitmap (+ per-record "has invalid" flag)
//   rank_build_kernel    per flagged record: valid-k-mer-start bitmap + prefix ranks (rank/select)
//   bf_build_kernel      K2: contig k-mers -> Bloom bits            (ref src/LINKS.hpp:316-324)
//   table_build_kernel   K4a: contig k-mers -> open-addressing table (ref :569-601,379-402)
//   extract_kernel       K3+K5a (extract.cuh): read k-mer pairs -> Bloom probe -> pair-set insert (ref :515-567,354-378)
//   classify_kernel      K4b: unique pairs -> table lookups -> link rows / issues (ref :824-968)
//
// HBM layout (all device arrays are flat, 256-byte aligned cudaMalloc blocks):
//   packed   u32[ceil(B/16)+pad]   2 bits per base, base j of a word at bits 2j..2j+1
//   inv      u32[ceil(B/32)+pad]   1 bit per base, 1 = not ACGTU
//   bloom    u32[bytes/4]          btllib byte array viewed as little-endian words
//   table    {u64 key,u64 val}[cap+1]   key = ntHash (fwd or rev) of a contig k-mer, EMPTY = ~0
//                                  val = pos<<32 | tig<<3 | orient<<2 | multiple<<1 | present
//   pairs    {u64 h_a,u64 h_b}[cap]         16 B slots, two per DRAM sector; 128-bit CAS claims a slot.  With several distances
//            {h_a, h_b, okey, -}[cap]       32 B slots: the order key (first distance wins) sits in the sector its pair is in
//   rows     u64 key[] + i32 distance[]  (dense packed row key, see pack_row_key)
#pragma once
#include "device_prims.cuh"
#include "nthash.cuh"

#include <cstdint>
#include <cuda_runtime.h>

namespace links {

constexpr uint64_t kEmpty = ~0ULL;
constexpr int kTile = 1024;        // candidates (or contig positions) per scheduling tile
constexpr int kCtaThreads = 256;

struct TableSlot { uint64_t key, val; };
struct PairSlot { uint64_t a, b; };

// occ (MATES build only): one bit per 32-byte sector of the table (two slots), set when a slot of it has been claimed
struct TableView { TableSlot* slots; uint64_t mask; int shift; /* 64 - log2(cap) */ uint32_t* occ = nullptr; };
struct PairView {
  PairSlot* slots; uint64_t mask; int shift;
  ulonglong2* list;            // dense (h_a,h_b) of every distinct pair, in insertion order
  uint32_t wide;               // 1: 32-byte slots {h_a, h_b, minimum order key, unused} (several distances); 0: 16-byte slots
};
__device__ __forceinline__ PairSlot* pair_slot(const PairView& pv, uint64_t i)
{
  return reinterpret_cast<PairSlot*>(reinterpret_cast<char*>(pv.slots) + (i << (4 + pv.wide)));
}
__device__ __forceinline__ unsigned long long* pair_okey_ptr(const PairView& pv, uint64_t i)  // wide slots only
{
  return reinterpret_cast<unsigned long long*>(pair_slot(pv, i)) + 2;
}

// counters accumulated on the device (one atomic per warp at kernel end)
struct Counters {
  unsigned long long n_candidates, n_survivors, n_unique, n_bf_probes, n_contig_kmers, n_bf_loads, n_member_sectors;
  unsigned long long ct_ok_pairs, ct_problem_pairs, ct_ok_contig, ct_iz_issues, ct_illogical, ct_gate_fail;
  unsigned long long n_rows, n_issues, issue_overflow;
};

struct IssueRec { // mirrors links_issue
  uint32_t tig; uint8_t o1, o2, pad_[2];
  uint64_t tig_len, s1, e1, s2, e2, pet_size;
  int64_t insert_size;
};

__device__ __forceinline__ unsigned long long warp_sum_ull(unsigned long long v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ uint64_t slot_hash1(uint64_t h) { return h * 0x9E3779B97F4A7C15ULL; }
__device__ __forceinline__ uint64_t slot_hash2(uint64_t a, uint64_t b)
{
  uint64_t x = a * 0x9E3779B97F4A7C15ULL + b;
  x ^= x >> 32;
  return x * 0xD6E8FEB86659FD93ULL;
}

// ------------------------------------------------------------------ exact member set of the Bloom filter's keys
//
// A Bloom filter has no false negatives: every canonical hash h0 inserted by the contig pass
// answers `contains` with true (bloom_filter.hpp:639-659).  The set below holds exactly those h0,
// so "h0 in set" proves the filter's answer with ONE 32-byte sector instead of hash_num probes;
// only k-mers outside the set (read errors, false-positive candidates) take the real probes.
// Results are identical by construction; the set exists only when this context inserted the
// contigs into the filter itself (or the caller vouched for it, links_bf_assume_covers_contigs).
// Layout: buckets of four u64 keys = one DRAM/L2 sector, home bucket = mulhi(slot_hash1(h0), n_buckets)
// (any bucket count, so the set is exactly as big as its load factor asks -- at 5 Mbp it stays
// L2-sized), overflow spills to the next bucket (load <= 0.5, so ~1.15 sectors per lookup).  ~0 marks a free
// slot; the value ~0 itself is never stored (it simply takes the probe path).
struct __align__(32) MemberBucket { unsigned long long s[4]; };
struct MemberView { MemberBucket* buckets; uint64_t n_buckets; };

__device__ __forceinline__ MemberBucket ld_bucket(const MemberBucket* p)
{
  MemberBucket r;
  asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.s[0]), "=l"(r.s[1]), "=l"(r.s[2]), "=l"(r.s[3]) : "l"(p));
  return r;
}

#ifndef LINKS_MEMBER_INSERT_VEC
#define LINKS_MEMBER_INSERT_VEC 1
#endif
__device__ __forceinline__ void member_insert(const MemberView& mv, uint64_t h0)
{
  if (h0 == kEmpty) return;
  uint64_t b = __umul64hi(slot_hash1(h0), mv.n_buckets);
  for (;;) {
    unsigned long long* s = mv.buckets[b].s;
#if LINKS_MEMBER_INSERT_VEC
    // one 32-byte load of the bucket instead of up to four dependent 8-byte ones.  It may be stale, but only towards
    // "empty" (slots are never cleared during a job), and a slot that looks empty is claimed with a CAS whose return
    // value is the truth.
    const MemberBucket v = ld_bucket(mv.buckets + b);
#endif
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#if LINKS_MEMBER_INSERT_VEC
      unsigned long long cur = v.s[j];
#else
      unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(s + j);
#endif
      if (cur == kEmpty) cur = atomicCAS(s + j, kEmpty, h0);
      if (cur == kEmpty || cur == h0) return;
    }
    b = b + 1 == mv.n_buckets ? 0 : b + 1;
  }
}

__device__ __forceinline__ bool member_contains(const MemberView& mv, uint64_t h0, uint32_t& sectors)
{
  uint64_t b = __umul64hi(slot_hash1(h0), mv.n_buckets);
  for (;;) {
    const MemberBucket v = ld_bucket(mv.buckets + b);
    ++sectors;
    if (v.s[0] == h0 || v.s[1] == h0 || v.s[2] == h0 || v.s[3] == h0) return h0 != kEmpty;
    if (v.s[3] == kEmpty) return false;  // slots fill front to back: a free last slot ends the chain
    b = b + 1 == mv.n_buckets ? 0 : b + 1;
  }
}

// ------------------------------------------------------------------ pack

__device__ __forceinline__ uint32_t find_record(const uint64_t* __restrict__ off, uint32_t n, uint64_t base0, uint64_t g)
{
  // largest r with off[r]-base0 <= g
  uint32_t lo = 0, hi = n;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (off[mid] - base0 <= g) lo = mid; else hi = mid;
  }
  return lo;
}

// one thread = 32 bases -> two packed words + one invalid-mask word
__global__ void __launch_bounds__(kCtaThreads) pack_kernel(const uint8_t* __restrict__ ascii, uint64_t n_bases,
                                                           uint32_t* __restrict__ packed, uint32_t* __restrict__ inv,
                                                           const uint64_t* __restrict__ off, uint32_t n_rec, uint64_t base0,
                                                           uint32_t* __restrict__ flags)
{
  const uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint64_t g0 = t * 32;
  if (g0 >= n_bases) return;
  uint32_t bytes[8];
  if (g0 + 32 <= n_bases && (reinterpret_cast<uintptr_t>(ascii) & 15) == 0) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(ascii + g0));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(ascii + g0) + 1);
    bytes[0] = a.x; bytes[1] = a.y; bytes[2] = a.z; bytes[3] = a.w;
    bytes[4] = b.x; bytes[5] = b.y; bytes[6] = b.z; bytes[7] = b.w;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      uint32_t v = 0;
      for (int q = 0; q < 4; ++q) {
        const uint64_t g = g0 + j * 4 + q;
        v |= uint32_t(g < n_bases ? ascii[g] : uint8_t('N')) << (8 * q);
      }
      bytes[j] = v;
    }
  }
  // four bases per step, SWAR on the 32-bit word (bytes b0..b3 = bases j..j+3):
  //   code   = (byte >> 1) & 3; the multiply gathers the four 2-bit codes into one byte
  //            (terms land on disjoint even bit positions, so no carries)
  //   valid  = byte & 0xDF in {A,C,G,T,U}: with v = (byte & 0xDF) ^ 0x40 that is v in {1,3,7,20,21},
  //            i.e. top three bits clear and  (~b4 & ~b3 & b0 & (~b2 | b1)) | (b4 & ~b3 & b2 & ~b1)
  uint32_t p[2] = {0, 0};
  uint32_t m = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t w = bytes[j];
    const uint32_t x = (w >> 1) & 0x03030303u;
    p[j >> 2] |= (((x * 0x00041041u) >> 18) & 0xFFu) << (8 * (j & 3));
    const uint32_t v = (w & 0xDFDFDFDFu) ^ 0x40404040u;
    const uint32_t b0 = v, b1 = v >> 1, b2 = v >> 2, b3 = v >> 3, b4 = v >> 4;
    const uint32_t hi = (v >> 5) | (v >> 6) | (v >> 7);
    const uint32_t ok = ((~b4 & ~b3 & b0 & (~b2 | b1)) | (b4 & ~b3 & b2 & ~b1)) & ~hi;
    const uint32_t bad = ~ok & 0x01010101u;
    m |= (((bad * 0x01020408u) >> 24) & 0xFu) << (4 * j);
  }
  packed[t * 2] = p[0];
  packed[t * 2 + 1] = p[1];
  inv[t] = m;
  // flag the records that own an invalid base (rare): they take the rank/select path
  if (g0 + 32 > n_bases) m &= (n_bases - g0 >= 32) ? 0xFFFFFFFFu : ((1u << (n_bases - g0)) - 1u);
  while (m) {
    const int j = __ffs(m) - 1;
    m &= m - 1;
    const uint32_t r = find_record(off, n_rec, base0, g0 + j);
    atomicOr(flags + r, 1u);
    // skip the other invalid bases of the same record inside this word
    const uint64_t end = off[r + 1] - base0;
    if (end >= g0 + 32) break;
    m &= ~((1u << (end - g0)) - 1u);
  }
}

// ------------------------------------------------------------------ rank/select for records with invalid bases
//
// btllib::NtHash::roll() visits, in increasing order, exactly the k-mer starts whose window holds
// no invalid base (nthash.hpp:1493-1533).  LINKS pairs the n-th such start of the stream begun
// at `offset` with the n-th of the stream begun at `offset+d` (src/LINKS.hpp:525-543), so a
// bit-exact kernel needs rank/select over "valid start" bits, not raw positions.
// For record r: words live at wb = (g0 >> 6) + 2r ; vs[wb+w] = valid-start bits of positions
// 64w..64w+63 ; vrank[wb+w] = number of valid starts before 64w ; vrank[wb+nwords] = total.

__global__ void __launch_bounds__(kCtaThreads) rank_build_kernel(const uint32_t* __restrict__ inv,
                                                                 const uint64_t* __restrict__ off, uint64_t base0,
                                                                 const uint32_t* __restrict__ flags, int k,
                                                                 unsigned long long* __restrict__ vs,
                                                                 uint32_t* __restrict__ vrank)
{
  const uint32_t r = blockIdx.x;
  if (!flags[r]) return;
  __shared__ uint32_t sm[9];
  const uint64_t g0 = off[r] - base0, len = off[r + 1] - off[r];
  const uint64_t wb = (g0 >> 6) + 2ull * r;
  const uint64_t npos = len >= uint64_t(k) ? len - k + 1 : 0;
  const uint64_t nwords = (npos + 63) >> 6;
  uint32_t carry = 0;
  for (uint64_t w0 = 0; w0 < nwords; w0 += kCtaThreads) {
    const uint64_t w = w0 + threadIdx.x;
    unsigned long long bits = 0;
    if (w < nwords) {
      const uint64_t p0 = w * 64;
      const uint64_t pend = p0 + 64 < npos ? p0 + 64 : npos; // positions [p0, pend)
      int run = 0;
      for (uint64_t x = p0; x < pend + k - 1; ++x) {
        const uint64_t g = g0 + x;
        const bool bad = (inv[g >> 5] >> (g & 31)) & 1u;
        run = bad ? 0 : run + 1;
        if (run >= k) bits |= 1ull << (x - (k - 1) - p0);
      }
      vs[wb + w] = bits;
    }
    uint32_t total;
    const uint32_t ex = block_excl_scan(uint32_t(__popcll(bits)), &total, sm);
    if (w < nwords) vrank[wb + w] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) vrank[wb + nwords] = carry;
}

struct RankView { const unsigned long long* vs; const uint32_t* vrank; uint64_t wb, nwords, npos; uint32_t total; };

__device__ __forceinline__ uint32_t rank_of(const RankView& rv, uint64_t p)
{
  if (p >= rv.npos) return rv.total;
  const uint64_t w = p >> 6;
  return rv.vrank[rv.wb + w] + __popcll(rv.vs[rv.wb + w] & ((1ull << (p & 63)) - 1ull));
}
__device__ __forceinline__ uint64_t select_of(const RankView& rv, uint32_t n) // n < total
{
  uint64_t lo = 0, hi = rv.nwords; // largest w with vrank[w] <= n
  while (hi - lo > 1) {
    const uint64_t mid = (lo + hi) >> 1;
    if (rv.vrank[rv.wb + mid] <= n) lo = mid; else hi = mid;
  }
  unsigned long long bits = rv.vs[rv.wb + lo];
  for (uint32_t i = rv.vrank[rv.wb + lo]; i < n; ++i) bits &= bits - 1;
  return lo * 64 + (__ffsll((long long)bits) - 1);
}

// ------------------------------------------------------------------ tile scheduling

// seg_tile_start[s] = first tile of segment s (cumulative); returns s owning `tile`
__device__ __forceinline__ uint32_t find_segment(const uint32_t* __restrict__ seg_tile_start, uint32_t n_seg, uint32_t tile)
{
  uint32_t lo = 0, hi = n_seg;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (seg_tile_start[mid] <= tile) lo = mid; else hi = mid;
  }
  return lo;
}

// same answer, found by the whole warp: 32 partition points per round instead of one (3 dependent loads for
// 25k segments instead of 15).  All lanes must call it; the result is warp-uniform.
__device__ __forceinline__ uint32_t find_segment_warp(const uint32_t* __restrict__ seg_tile_start, uint32_t n_seg, uint32_t tile)
{
  const uint32_t lane = threadIdx.x & 31;
  uint32_t lo = 0, hi = n_seg;  // invariant: seg[lo] <= tile, (hi == n_seg or seg[hi] > tile)
  while (hi - lo > 1) {
    const uint32_t span = hi - lo;
    const uint32_t stride = (span + 32) / 33;            // 32 probes cut [lo,hi) into <= 33 pieces
    const uint32_t idx = lo + (lane + 1) * stride;
    const bool le = idx < hi && __ldg(seg_tile_start + idx) <= tile;
    const uint32_t c = __popc(__ballot_sync(0xffffffffu, le));  // probes are monotone: the first c are true
    const uint32_t nlo = lo + c * stride;
    const uint32_t nhi = lo + (c + 1) * stride;
    lo = nlo;
    if (nhi < hi) hi = nhi;
  }
  return lo;
}

// ------------------------------------------------------------------ contigs: Bloom build + table build

struct ContigBatch {
  const uint32_t* packed; const uint32_t* inv; const uint64_t* off; uint64_t base0;
  const uint32_t* seg_tile_start; uint32_t n_seq, n_tiles; uint32_t first_tig;
};

__global__ void table_init_kernel(TableSlot* __restrict__ slots, uint64_t n)
{
  const ulonglong2 e = make_ulonglong2(kEmpty, 0ull);
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x)
    reinterpret_cast<ulonglong2*>(slots)[i] = e;
}

#ifndef LINKS_TABLE_CAS128
#define LINKS_TABLE_CAS128 1
#endif
__device__ __forceinline__ ulonglong2 cas128_raw(void* slot, ulonglong2 cmp, ulonglong2 val)
{
  ulonglong2 old;
  asm volatile(
    "{\n\t.reg .b128 c, v, o;\n\t"
    "mov.b128 c, {%2, %3};\n\t"
    "mov.b128 v, {%4, %5};\n\t"
    "atom.global.cas.b128 o, [%6], c, v;\n\t"
    "mov.b128 {%0, %1}, o;\n\t}"
    : "=l"(old.x), "=l"(old.y)
    : "l"(cmp.x), "l"(cmp.y), "l"(val.x), "l"(val.y), "l"(slot)
    : "memory");
  return old;
}

__device__ __forceinline__ void table_insert(const TableView& tv, uint64_t h, uint64_t payload)
{
  if (h == kEmpty) { // the sentinel value itself is a legal hash: it owns the extra slot [cap]
    unsigned long long* v = reinterpret_cast<unsigned long long*>(&tv.slots[tv.mask + 1].val);
    const unsigned long long old = atomicOr(v, 1ull);
    atomicOr(v, (old & 1ull) ? 2ull : payload);
    return;
  }
  uint64_t i = slot_hash1(h) >> tv.shift;
  for (;;) {
#if LINKS_TABLE_CAS128
    // key and value claimed by ONE 16-byte CAS on the empty slot {~0, 0}: one memory round trip for a first
    // occurrence instead of three dependent ones (load key, CAS key, OR value)
    const ulonglong2 cur = cas128_raw(&tv.slots[i], make_ulonglong2(kEmpty, 0ull), make_ulonglong2(h, payload | 1ull));
    if (cur.x == kEmpty && cur.y == 0ull) return;  // first of this hash: (tig,pos,orient)
    if (cur.x == h) {  // seen again: multiple (ref :398-400)
      if (!(cur.y & 2ull)) atomicOr(reinterpret_cast<unsigned long long*>(&tv.slots[i].val), 2ull);
      return;
    }
#else
    unsigned long long* kp = reinterpret_cast<unsigned long long*>(&tv.slots[i].key);
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(kp);
    if (cur == kEmpty) cur = atomicCAS(kp, kEmpty, h);
    unsigned long long* vp = reinterpret_cast<unsigned long long*>(&tv.slots[i].val);
    if (cur == kEmpty) { atomicOr(vp, payload | 1ull); return; } // first of this hash: (tig,pos,orient)
    if (cur == h) { atomicOr(vp, 2ull); return; }                // seen again: multiple (ref :398-400)
#endif
    i = (i + 1) & tv.mask;
  }
}

__device__ __forceinline__ uint64_t table_lookup(const TableView& tv, uint64_t h)
{
  if (h == kEmpty) return tv.slots[tv.mask + 1].val;
  uint64_t i = slot_hash1(h) >> tv.shift;
  for (;;) {
    const ulonglong2 s = __ldg(reinterpret_cast<const ulonglong2*>(&tv.slots[i]));
    if (s.x == h) return s.y;
    if (s.x == kEmpty) return 0;
    i = (i + 1) & tv.mask;
  }
}

// ---- the table restricted to `mates` (reference src/LINKS.hpp:374-375: mates = every h_a, h_b of mate_pair;
// populate_mate_info :580,591 keeps a contig k-mer only if its hash is in there).  Same slots, two steps:
//   table_seed    every hash the (owned) pairs refer to claims a slot {h, 0}: "a mate, not seen in a contig yet"
//   table_update  a contig k-mer whose hash has a slot fills it in (first occurrence) or marks it multiple;
//                 a hash without a slot is no mate and leaves no trace
// classify's table_lookup then reads val = 0 for an absent mate -> the default KmerInfo, as in the full table.
// Size is bounded by twice the distinct pairs instead of twice the genome (3.1 Gbp: 198 GB full, replicated).
__device__ __forceinline__ void table_seed(const TableView& tv, uint64_t h)
{
  if (h == kEmpty) return;  // owns the extra slot [cap], which always exists
  uint64_t i = slot_hash1(h) >> tv.shift;
  for (;;) {
    ulonglong2 cur = __ldcg(reinterpret_cast<const ulonglong2*>(&tv.slots[i]));
    if (cur.x == kEmpty) {
      cur = cas128_raw(&tv.slots[i], make_ulonglong2(kEmpty, 0ull), make_ulonglong2(h, 0ull));  // (a stale "empty" is settled by the CAS)
      if (cur.x == kEmpty) {  // claimed: mark the sector occupied (read by the NEXT kernel only)
        if (tv.occ) atomicOr(tv.occ + (i >> 6), 1u << ((i >> 1) & 31));
        return;
      }
    }
    if (cur.x == h) return;
    i = (i + 1) & tv.mask;
  }
}

__device__ __forceinline__ void table_update(const TableView& tv, uint64_t h, uint64_t payload)
{
  if (h == kEmpty) { table_insert(tv, h, payload); return; }
  uint64_t i = slot_hash1(h) >> tv.shift;
  // Most contig k-mers are not mates of THIS rank's pairs (with the pairs split over N ranks: ~1 - 1/N of them), and each
  // such miss would cost a random DRAM sector.  The occupancy bitmap (1 bit per sector, 1/256 of the table: L2-resident)
  // answers the common case first: nothing was ever claimed in the home sector => the home slot is empty => the key,
  // which linear probing would have put there or after an unbroken run of occupied slots, is absent.
  if (tv.occ && !((__ldg(tv.occ + (i >> 6)) >> ((i >> 1) & 31)) & 1u)) return;
  for (;;) {
    // keys are final (seeded by the previous kernel); values move 0 -> payload|1 -> |2 only
    ulonglong2 s = __ldcg(reinterpret_cast<const ulonglong2*>(&tv.slots[i]));
    if (s.x == h) {
      if (s.y == 0ull) {
        s = cas128_raw(&tv.slots[i], make_ulonglong2(h, 0ull), make_ulonglong2(h, payload | 1ull));
        if (s.y == 0ull) return;  // first of this hash: (tig,pos,orient)
      }
      if (!(s.y & 2ull)) atomicOr(reinterpret_cast<unsigned long long*>(&tv.slots[i].val), 2ull);  // seen again: multiple (ref :398-400)
      return;
    }
    if (s.x == kEmpty) return;  // not in mates (ref :580,591)
    i = (i + 1) & tv.mask;
  }
}

__global__ void __launch_bounds__(kCtaThreads) table_seed_kernel(const ulonglong2* __restrict__ list, uint64_t n_pairs, TableView tv)
{
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_pairs; i += uint64_t(gridDim.x) * blockDim.x) {
    const ulonglong2 hk = __ldg(list + i);
    table_seed(tv, hk.x);
    table_seed(tv, hk.y);
  }
}

// MODE 0: Bloom insert (KmerBloomFilter::insert, bloom_filter.hpp:814-821) + member-set insert (mv.buckets != nullptr)
// MODE 1: table insert of fwd and rev hash of every valid position (populate_mate_info, :577-598)
// MODE 2: member-set insert only (the filter's bits came from links_bf_load*, the caller vouches they cover these contigs)
// MODE 3: update of the table restricted to `mates` (table_update): only hashes that already own a slot are recorded
// A filter far larger than L2 is built in several passes over the contigs (MODE 0): a pass sets only the bits whose
// 32-bit word lies in [w_lo, w_hi), so its atomics stay in L2 (3.2 G positions/s with a DRAM-resident 180 MB filter
// against 21 G/s when the filter fits); the member set is filled by the first pass only.  One pass = {0, ~0, 1}.
// share_n > 1 (links_bf_alloc_shared): this rank inserts only the tiles with tile % share_n == share_rank.
struct BloomPass { uint64_t w_lo, w_hi; int with_set; uint32_t share_rank = 0, share_n = 1; };

template <int MODE>
__global__ void __launch_bounds__(kCtaThreads) contig_kernel(ContigBatch cb, int k, uint64_t kseed,
                                                             const ulonglong2* __restrict__ g_tab, BloomView bf,
                                                             TableView tv, MemberView mv, Counters* __restrict__ ctr, BloomPass bp)
{
  extern __shared__ ulonglong2 s_tab[];
  for (int i = threadIdx.x; i < k * 4; i += blockDim.x) s_tab[i] = g_tab[i];
  __shared__ uint32_t s_seg;
  __syncthreads();
  unsigned long long n_kmers = 0;
  // MODE 0 with shares: this rank's tiles are share_rank, share_rank + share_n, ...; they are enumerated directly (striding
  // by the grid and skipping foreign tiles would leave only the CTAs with blockIdx % share_n == share_rank any work whenever
  // the grid size is a multiple of share_n -- and 148 x 8 is)
  const uint32_t t_step = MODE == 0 && bp.share_n > 1 ? bp.share_n : 1u;
  const uint32_t t_first = MODE == 0 && bp.share_n > 1 ? bp.share_rank : 0u;
  for (uint64_t tile64 = t_first + uint64_t(blockIdx.x) * t_step; tile64 < cb.n_tiles; tile64 += uint64_t(gridDim.x) * t_step) {
    const uint32_t tile = uint32_t(tile64);
    if (threadIdx.x == 0) s_seg = find_segment(cb.seg_tile_start, cb.n_seq, tile);
    __syncthreads();
    const uint32_t c = s_seg;
    const uint64_t g0 = cb.off[c] - cb.base0, len = cb.off[c + 1] - cb.off[c];
    const uint64_t npos = len - k + 1; // segments with len < k own zero tiles
    const uint64_t p_begin = uint64_t(tile - cb.seg_tile_start[c]) * kTile;
    // a thread walks kTile / kCtaThreads consecutive positions: the first k-mer is hashed from scratch (k lookups), the
    // others rolled from it (2 lookups each)
    constexpr int kRun = kTile / kCtaThreads;
    uint64_t f = 0, r = 0;
    for (int j = 0; j < kRun; ++j) {
      const uint64_t p = p_begin + uint64_t(threadIdx.x) * kRun + j;
      if (p >= npos) break;
      if (j == 0) hash_kmer(cb.packed, g0 + p, k, s_tab, f, r);
      else roll_kmer(cb.packed, g0 + p - 1, k, s_tab, f, r);
      if (!window_valid(cb.inv, g0 + p, k)) continue;
      if (MODE == 0) {
        bloom_insert_range(bf, f, r, kseed, bp.w_lo, bp.w_hi);
        if (bp.with_set && mv.buckets) member_insert(mv, r < f ? r : f);
      } else if (MODE == 2) {
        member_insert(mv, r < f ? r : f);
      } else if (MODE == 3) {
        const uint64_t payload = (p << 32) | (uint64_t(cb.first_tig + c) << 3);
        table_update(tv, f, payload);
        table_update(tv, r, payload | 4ull);
        n_kmers += 2;
      } else {
        const uint64_t payload = (p << 32) | (uint64_t(cb.first_tig + c) << 3);
        table_insert(tv, f, payload);
        table_insert(tv, r, payload | 4ull);
        n_kmers += 2;
      }
    }
    __syncthreads();
  }
  if (MODE == 1 || MODE == 3) {
    n_kmers = warp_sum_ull(n_kmers);
    if ((threadIdx.x & 31) == 0 && n_kmers) atomicAdd(&ctr->n_contig_kmers, n_kmers);
  }
}

// ------------------------------------------------------------------ Bloom build for filters beyond L2: partition, then apply
//
// hash_num random read-modify-writes per contig k-mer into a DRAM-resident filter cost two DRAM operations each (sector
// in, sector back) and DRAM delivers ~43 G random sectors/s (tools/gather_peak.cu): 3.2 G k-mers/s, the longest stage of a
// Gbp-scale job.  Restricting a pass to an L2-sized block of the filter (BloomPass) keeps the atomics in L2 but hashes
// every k-mer once per block.  Here every k-mer is hashed ONCE: its hash_num updates are appended, as 4-byte entries
// (word within the block << 5 | bit), to one bucket per L2-sized block of the filter -- sequential streams, which DRAM
// takes at copy speed: 2 x 4 B x hash_num streamed per k-mer instead of 2 x hash_num random sectors -- and a second
// kernel, launched once per bucket, replays it with the atomics landing in an L2-resident block.
struct BloomBuckets {
  uint32_t* entries;             // bucket b occupies [b * cap, (b + 1) * cap)
  unsigned long long* cursor;    // [n_buckets] entries appended (may exceed cap: the excess is dropped and flagged)
  uint64_t cap;
  uint32_t n_buckets, block_shift;  // filter word w belongs to bucket w >> block_shift
  unsigned int* overflow;
};
constexpr int kMaxBloomBuckets = 256;

template <bool WITH_SET>
__global__ void __launch_bounds__(kCtaThreads) bloom_partition_kernel(ContigBatch cb, uint32_t tile_lo, uint32_t tile_hi, int k, uint64_t kseed,
                                                                     const ulonglong2* __restrict__ g_tab, BloomView bf, MemberView mv,
                                                                     BloomBuckets bk, uint32_t share_rank, uint32_t share_n)
{
  constexpr int kWarps = kCtaThreads / 32;
  extern __shared__ ulonglong2 s_tab[];
  // per-warp counters: a warp counts its own entries per bucket (lanes with the same bucket are found with match.any, one
  // plain add per distinct bucket), so no shared-memory atomics and no contention however few buckets there are
  __shared__ uint32_t s_cnt[kWarps][kMaxBloomBuckets];
  __shared__ unsigned long long s_base[kMaxBloomBuckets];
  __shared__ uint32_t s_seg;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt = (1u << lane) - 1u;
  for (int i = threadIdx.x; i < k * 4; i += blockDim.x) s_tab[i] = g_tab[i];
  for (uint32_t i = threadIdx.x; i < uint32_t(kWarps) * kMaxBloomBuckets; i += blockDim.x) (&s_cnt[0][0])[i] = 0;
  __syncthreads();
  const uint32_t block_mask = (1u << bk.block_shift) - 1u;
  // this rank's tiles of [tile_lo, tile_hi): the first one congruent to share_rank, then every share_n-th (enumerated
  // directly: a grid-stride walk that skips foreign tiles idles all CTAs but every share_n-th when share_n divides the grid)
  const uint32_t t_first = tile_lo + (share_rank + share_n - tile_lo % share_n) % share_n;
  for (uint64_t tile64 = t_first + uint64_t(blockIdx.x) * share_n; tile64 < tile_hi; tile64 += uint64_t(gridDim.x) * share_n) {
    const uint32_t tile = uint32_t(tile64);
    if (threadIdx.x == 0) s_seg = find_segment(cb.seg_tile_start, cb.n_seq, tile);
    __syncthreads();
    const uint32_t c = s_seg;
    const uint64_t g0 = cb.off[c] - cb.base0, len = cb.off[c + 1] - cb.off[c];
    const uint64_t npos = len - k + 1;
    const uint64_t p_begin = uint64_t(tile - cb.seg_tile_start[c]) * kTile;
    constexpr int kRun = kTile / kCtaThreads;  // a thread walks kRun consecutive positions, one per round: hashed once, then rolled
    uint64_t f = 0, r = 0;
    for (int j = 0; j < kRun; ++j) {  // one round: every thread one position, hash_num entries
      const uint64_t p = p_begin + uint64_t(threadIdx.x) * kRun + j;
      if (p < npos) {
        if (j == 0) hash_kmer(cb.packed, g0 + p, k, s_tab, f, r);
        else roll_kmer(cb.packed, g0 + p - 1, k, s_tab, f, r);
      }
      const bool live = p < npos && window_valid(cb.inv, g0 + p, k);
      uint64_t h0 = 0;
      if (live) {
        h0 = r < f ? r : f;
        if (WITH_SET && mv.buckets) member_insert(mv, h0);
      }
      __syncwarp();
      for (uint32_t i = 0; i < bf.hash_num; ++i) {  // pass 1: count per bucket
        uint32_t b = 0xFFFFFFFFu;
        if (live) b = uint32_t(bloom_mod(multihash_i(h0, kseed, i), bf) >> 5) >> bk.block_shift;
        const uint32_t peers = __match_any_sync(0xffffffffu, b);
        if (live && lane == __ffs(peers) - 1) s_cnt[warp][b] += __popc(peers);
        __syncwarp();
      }
      __syncthreads();
      for (uint32_t b = threadIdx.x; b < bk.n_buckets; b += blockDim.x) {  // one reservation per bucket and round
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) { const uint32_t n = s_cnt[w][b]; s_cnt[w][b] = total; total += n; }  // -> each warp's offset
        s_base[b] = total ? atomicAdd(bk.cursor + b, (unsigned long long)total) : 0ull;
      }
      __syncthreads();
      for (uint32_t i = 0; i < bf.hash_num; ++i) {  // pass 2: write (the multi-hash is a multiply and a shift: cheaper to redo than to keep)
        uint32_t b = 0xFFFFFFFFu, w = 0, bit = 0;
        if (live) {
          const uint64_t n = bloom_mod(multihash_i(h0, kseed, i), bf);
          w = uint32_t(n >> 5); bit = uint32_t(n & 31); b = w >> bk.block_shift;
        }
        const uint32_t peers = __match_any_sync(0xffffffffu, b);
        const int leader = __ffs(peers) - 1;
        uint32_t first = 0;
        if (live && lane == leader) { first = s_cnt[warp][b]; s_cnt[warp][b] = first + __popc(peers); }
        first = __shfl_sync(0xffffffffu, first, leader);
        if (live) {
          const unsigned long long at = s_base[b] + first + __popc(peers & lt);
          if (at < bk.cap) bk.entries[uint64_t(b) * bk.cap + at] = ((w & block_mask) << 5) | bit;
          else *bk.overflow = 1u;
        }
        __syncwarp();
      }
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < uint32_t(kWarps) * bk.n_buckets; i += blockDim.x) s_cnt[i / bk.n_buckets][i % bk.n_buckets] = 0;
      __syncthreads();
    }
  }
}

// replay of bucket b: its atomics stay in one L2-sized block of the filter.  ONE LAUNCH PER BUCKET: a single kernel that
// walked bucket after bucket let the CTAs drift apart (nothing holds a fast CTA back: a RED does not wait for its result),
// so that several blocks were live at once and pushed each other out of L2 -- 47 G updates/s with 64 MB blocks, 83 G/s with
// 16 MB blocks, against the 181-202 G/s the same access pattern reaches on one block (tools/atomic_peak.cu,
// profiles/r02z_atomic_peak.jsonl).  The kernel boundary is the grid-wide barrier.
__global__ void __launch_bounds__(kCtaThreads) bloom_apply_kernel(BloomView bf, BloomBuckets bk, uint32_t b)
{
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const unsigned long long have = bk.cursor[b];
  const uint64_t n = have < bk.cap ? have : bk.cap;
  const uint32_t* e = bk.entries + uint64_t(b) * bk.cap;
  uint32_t* words = bf.words + (uint64_t(b) << bk.block_shift);
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t v = __ldcs(e + i);  // streamed once
    atomicOr(words + (v >> 5), 1u << (v & 31));
  }
}

// ------------------------------------------------------------------ reads: extract + probe + pair-set insert

__device__ __forceinline__ ulonglong2 cas128(PairSlot* slot, ulonglong2 cmp, ulonglong2 val)
{
  ulonglong2 old;
  asm volatile(
    "{\n\t.reg .b128 c, v, o;\n\t"
    "mov.b128 c, {%2, %3};\n\t"
    "mov.b128 v, {%4, %5};\n\t"
    "atom.global.cas.b128 o, [%6], c, v;\n\t"
    "mov.b128 {%0, %1}, o;\n\t}"
    : "=l"(old.x), "=l"(old.y)
    : "l"(cmp.x), "l"(cmp.y), "l"(val.x), "l"(val.y), "l"(slot)
    : "memory");
  return old;
}

// mate_pair first-wins insert (ref src/LINKS.hpp:370-373): set semantics on (a,b); the winner's
// distance index is recovered from the minimum order key.  Returns true if (a,b) was new.
__device__ __forceinline__ bool pair_insert(const PairView& pv, uint64_t a, uint64_t b, uint64_t okey, bool multi_d)
{
  uint64_t i = slot_hash2(a, b) >> pv.shift;
  bool fresh = false;
  for (;;) {
    PairSlot* s = pair_slot(pv, i);
    ulonglong2 cur = __ldcg(reinterpret_cast<const ulonglong2*>(s));
    if (cur.x == kEmpty || cur.y == kEmpty) { // empty (or a torn read of a slot being claimed)
      cur = cas128(s, make_ulonglong2(kEmpty, kEmpty), make_ulonglong2(a, b));
      if (cur.x == kEmpty && cur.y == kEmpty) { fresh = true; cur = make_ulonglong2(a, b); }
    }
    if (cur.x == a && cur.y == b) {
      if (multi_d) atomicMin(pair_okey_ptr(pv, i), (unsigned long long)okey);  // same sector as the pair: no second DRAM access
      return fresh;
    }
    i = (i + 1) & pv.mask;
  }
}

// order key of a pair known to be present.  L2 loads, not the read-only path: the incremental classify_kernel runs
// beside an extract_kernel that is claiming slots of the same table and lowering order keys.  What it reads is still
// final: a listed pair's slot never moves, the slots in front of it on its probe chain were occupied before it was
// inserted, and its order key can only be lowered by a read with a smaller global ordinal, which classify_increment
// rules out (reads arrive in increasing ordinal, else cls_valid is dropped).
__device__ __forceinline__ uint64_t pair_okey(const PairView& pv, uint64_t a, uint64_t b)
{
  uint64_t i = slot_hash2(a, b) >> pv.shift;
  for (;;) {
    const PairSlot* s = pair_slot(pv, i);
    const ulonglong2 cur = __ldcg(reinterpret_cast<const ulonglong2*>(s));
    if (cur.x == a && cur.y == b) return pv.wide ? __ldcg(pair_okey_ptr(pv, i)) : 0ull;
    if (cur.x == kEmpty && cur.y == kEmpty) return kEmpty; // not present (cannot happen for listed pairs)
    i = (i + 1) & pv.mask;
  }
}

constexpr int kOkeyRankBits = 28, kOkeyDistBits = 8;
__host__ __device__ __forceinline__ uint64_t make_okey(uint64_t read_idx, uint32_t d_idx, uint64_t rank)
{
  return (read_idx << (kOkeyRankBits + kOkeyDistBits)) | (uint64_t(d_idx) << kOkeyRankBits) | rank;
}
__host__ __device__ __forceinline__ uint32_t okey_d_idx(uint64_t okey)
{
  return uint32_t(okey >> kOkeyRankBits) & ((1u << kOkeyDistBits) - 1u);
}

// re-insert every occupied slot of `src` into `dst` (growth) -- also used by links_pairs_import
__global__ void pair_rehash_kernel(PairView src, uint64_t n_src, PairView dst, Counters* ctr, int count_unique)
{
  unsigned long long fresh = 0;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_src; i += uint64_t(gridDim.x) * blockDim.x) {
    const PairSlot s = *pair_slot(src, i);
    if (s.a == kEmpty && s.b == kEmpty) continue;
    if (pair_insert(dst, s.a, s.b, src.wide ? *pair_okey_ptr(src, i) : 0ull, src.wide != 0)) ++fresh;
  }
  if (count_unique) {
    fresh = warp_sum_ull(fresh);
    if ((threadIdx.x & 31) == 0 && fresh) atomicAdd(&ctr->n_unique, fresh);
  }
}

// ------------------------------------------------------------------ stage 4: classify unique pairs

// dense row key: bin_code << (2*tb) | A << tb | B with A/B = tig<<1 | rev ; tb = bits of (n_tigs<<1|1)
__host__ __device__ __forceinline__ uint64_t pack_row_key(uint32_t bin_code, uint32_t a, uint32_t b, int tb)
{
  return (uint64_t(bin_code) << (2 * tb)) | (uint64_t(a) << tb) | uint64_t(b);
}
__host__ __device__ __forceinline__ int bin_code_of(int distance) // get_distance_bin, ref :764-770
{
  return distance < 0 ? 0 : distance == 10 ? 1 : distance < 500 ? 2 : distance < 5000 ? 3 : 4;
}

struct ClassifyParams {
  const uint32_t* dist; const int32_t* min_allowed; uint32_t n_dist;
  const uint32_t* tig_len; // [0] = 0 : the empty tig id
  int tb;
  uint64_t* row_keys; int32_t* row_dist; uint64_t row_cap;
  IssueRec* issues; uint64_t issue_cap;
};

struct KInfo { uint32_t tig; uint32_t start, end; bool orient; bool multiple; };
__device__ __forceinline__ KInfo decode_kinfo(uint64_t v, int k)
{
  KInfo ki;
  if (!(v & 1ull)) { ki.tig = 0; ki.start = 0; ki.end = 0; ki.orient = false; ki.multiple = false; return ki; } // KmerInfo() ref :123-136
  ki.tig = uint32_t(v >> 3) & 0x1FFFFFFFu; ki.start = uint32_t(v >> 32); ki.end = ki.start + k;
  ki.orient = (v >> 2) & 1ull; ki.multiple = (v >> 1) & 1ull;
  return ki;
}

// `range` != nullptr: classify list entries [range[0], range[1]) (bounds live on the device: the host
// launches this beside the extraction without knowing how many pairs exist yet); else [0, n_pairs).
__global__ void __launch_bounds__(kCtaThreads) classify_kernel(PairView pv, uint64_t n_pairs, TableView tv, ClassifyParams cp,
                                                               int k, Counters* __restrict__ ctr,
                                                               const unsigned long long* __restrict__ range)
{
  uint64_t first = 0;
  if (range) { first = range[0]; n_pairs = range[1] > range[0] ? range[1] - range[0] : 0; }
  pv.list += first;
  __shared__ uint32_t sm_scan[9];
  __shared__ unsigned long long s_base;
  unsigned long long c_ok = 0, c_prob = 0, c_okc = 0, c_iz = 0, c_ill = 0, c_gate = 0;
  const int lane = threadIdx.x & 31;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t n_round = (n_pairs + stride - 1) / stride * stride; // whole blocks stay in the loop
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    bool emit = false;
    uint64_t rkey = 0; int32_t rdist = 0;
    if (i < n_pairs) {
      const ulonglong2 hk = __ldg(pv.list + i);
      {
        const uint32_t di = cp.n_dist > 1 ? okey_d_idx(pair_okey(pv, hk.x, hk.y)) : 0;
        const KInfo k1 = decode_kinfo(table_lookup(tv, hk.x), k);
        const KInfo k2 = decode_kinfo(table_lookup(tv, hk.y), k);
        if (k1.multiple || k2.multiple) { ++c_gate; } // ref :827-830
        else {
          const int64_t insert_size = cp.dist[di];
          const int min_allowed = cp.min_allowed[di]; // ref :838, evaluated on the host in float
          if (k1.tig != k2.tig) { // ref :851-899
            const uint32_t l1 = cp.tig_len[k1.tig], l2 = cp.tig_len[k2.tig];
            // get_distance (ref :772-779) in 32-bit wrap-around arithmetic == its uint64->int truncations
            uint32_t span; uint32_t oe;
            if (!k1.orient) {
              if (!k2.orient) { span = (l1 - k1.start) + k2.start; oe = 0; }
              else { span = (l1 - k1.start) + (l2 - k2.end); oe = 1; }
            } else {
              if (!k2.orient) { span = (l1 - (l1 - k1.end)) + k2.start; oe = 2; }
              else { span = (l2 - k2.end) + k1.end; oe = 3; }
            }
            const int distance = int(uint32_t(insert_size) - span);
            if (distance > min_allowed && int64_t(distance) < insert_size) {
              ++c_ok;
              // add_to_pair_map (ref :717-733): first row X->Y, second row flip(Y)->flip(X).
              uint32_t A, B;
              switch (oe) {
              case 0: A = k1.tig << 1; B = k2.tig << 1; break;
              case 1: A = k1.tig << 1; B = (k2.tig << 1) | 1u; break;
              case 2: A = (k1.tig << 1) | 1u; B = k2.tig << 1; break;
              default: A = (k1.tig << 1) | 1u; B = (k2.tig << 1) | 1u; break;
              }
              const uint32_t bc = bin_code_of(distance);
              const uint64_t k_first = pack_row_key(bc, A, B, cp.tb);
              const uint64_t k_second = pack_row_key(bc, B ^ 1u, A ^ 1u, cp.tb);
              rkey = k_first < k_second ? k_first : k_second; // the two rows always travel together
              rdist = distance;
              emit = true;
            } else { ++c_prob; }
          } else { // same contig, ref :900-954
            if (k1.orient == k2.orient) {
              const unsigned low_iz = unsigned(insert_size + min_allowed), up_iz = unsigned(insert_size - min_allowed);
              const uint64_t pet = k1.start > k2.start ? uint64_t(k1.start - k2.start) : uint64_t(k2.start - k1.start);
              if (pet >= low_iz && pet <= up_iz) { ++c_okc; }
              else {
                ++c_iz;
                const unsigned long long slot = atomicAdd(&ctr->n_issues, 1ull);
                if (slot < cp.issue_cap) {
                  IssueRec is;
                  is.tig = k1.tig; is.o1 = k1.orient; is.o2 = k2.orient; is.pad_[0] = is.pad_[1] = 0;
                  is.tig_len = cp.tig_len[k1.tig];
                  is.s1 = k1.start; is.e1 = k1.end; is.s2 = k2.start; is.e2 = k2.end;
                  is.pet_size = pet; is.insert_size = insert_size;
                  cp.issues[slot] = is;
                }
              }
            } else { ++c_ill; }
          }
        }
      }
    }
    // block-level compaction of the emitted rows: one global atomic per block iteration
    uint32_t total;
    const uint32_t ex = block_excl_scan(emit ? 1u : 0u, &total, sm_scan);
    if (total) {
      if (threadIdx.x == 0) s_base = atomicAdd(&ctr->n_rows, (unsigned long long)total);
      __syncthreads();
      if (emit) {
        const uint64_t dst = s_base + ex;
        if (dst < cp.row_cap) { cp.row_keys[dst] = rkey; cp.row_dist[dst] = rdist; }
      }
      __syncthreads();
    }
  }
  c_ok = warp_sum_ull(c_ok); c_prob = warp_sum_ull(c_prob); c_okc = warp_sum_ull(c_okc);
  c_iz = warp_sum_ull(c_iz); c_ill = warp_sum_ull(c_ill); c_gate = warp_sum_ull(c_gate);
  if (lane == 0) {
    if (c_ok) atomicAdd(&ctr->ct_ok_pairs, c_ok);
    if (c_prob) atomicAdd(&ctr->ct_problem_pairs, c_prob);
    if (c_okc) atomicAdd(&ctr->ct_ok_contig, c_okc);
    if (c_iz) atomicAdd(&ctr->ct_iz_issues, c_iz);
    if (c_ill) atomicAdd(&ctr->ct_illogical, c_ill);
    if (c_gate) atomicAdd(&ctr->ct_gate_fail, c_gate);
  }
}

// After reduce-by-key: every canonical row i stands for two rows of the tsv (add_to_pair_map,
// ref :717-761).  Emit both keys with payload = i (second row: i | 0x40000000 is not needed, the
// payload only indexes links/gaps) so a final small sort puts all rows in output order.
__global__ void mirror_expand_kernel(const uint64_t* __restrict__ keys, uint32_t n, int tb, uint64_t* __restrict__ out_keys,
                                     int32_t* __restrict__ out_idx)
{
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t key = keys[i];
  const uint64_t tmask = (1ull << tb) - 1;
  const uint32_t B = uint32_t(key & tmask), A = uint32_t((key >> tb) & tmask), bc = uint32_t(key >> (2 * tb));
  out_keys[2 * i] = key;
  out_keys[2 * i + 1] = pack_row_key(bc, B ^ 1u, A ^ 1u, tb);
  out_idx[2 * i] = int32_t(i);
  out_idx[2 * i + 1] = int32_t(i);
}

// same layout as links_row (include/links_gpu.h): the rows leave the device ready for the caller
struct FinalRow { int32_t bin; uint32_t tig_a, tig_b; uint8_t rev_a, rev_b, pad_[2]; unsigned long long links; long long gaps; };
__global__ void gather_rows_kernel(const uint64_t* __restrict__ keys, const int32_t* __restrict__ idx, uint32_t n, int tb,
                                   const unsigned long long* __restrict__ links, const unsigned long long* __restrict__ gaps,
                                   FinalRow* __restrict__ out)
{
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t key = keys[i];
  const uint64_t tmask = (1ull << tb) - 1;
  const uint32_t B = uint32_t(key & tmask), A = uint32_t((key >> tb) & tmask), bc = uint32_t(key >> (2 * tb));
  FinalRow r;
  r.bin = bc == 0 ? -1 : bc == 1 ? 10 : bc == 2 ? 500 : bc == 3 ? 5000 : 10000;  // get_distance_bin's values, ref :764-770
  r.tig_a = A >> 1; r.rev_a = A & 1u; r.tig_b = B >> 1; r.rev_b = B & 1u; r.pad_[0] = r.pad_[1] = 0;
  r.links = links[idx[i]]; r.gaps = (long long)gaps[idx[i]];
  out[i] = r;
}

// ------------------------------------------------------------------ inspection helpers

__global__ void kmer_hashes_kernel(const uint32_t* __restrict__ packed, const uint32_t* __restrict__ inv, uint64_t npos,
                                   int k, const ulonglong2* __restrict__ g_tab, uint64_t* __restrict__ fwd,
                                   uint64_t* __restrict__ rev, uint8_t* __restrict__ valid)
{
  extern __shared__ ulonglong2 s_tab[];
  for (int i = threadIdx.x; i < k * 4; i += blockDim.x) s_tab[i] = g_tab[i];
  __syncthreads();
  for (uint64_t p = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; p < npos; p += uint64_t(gridDim.x) * blockDim.x) {
    const bool ok = window_valid(inv, p, k);
    uint64_t f = 0, r = 0;
    if (ok) hash_kmer(packed, p, k, s_tab, f, r);
    fwd[p] = f; rev[p] = r; valid[p] = ok;
  }
}

__global__ void bf_contains_kernel(BloomView bf, uint64_t kseed, const uint64_t* __restrict__ fwd,
                                   const uint64_t* __restrict__ rev, uint64_t n, uint8_t* __restrict__ out)
{
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t probes = 0;
  out[i] = bloom_contains(bf, fwd[i], rev[i], kseed, probes);
}

// ------------------------------------------------------------------ multi-GPU exchange
//
// Reads are sharded over ranks, so the same ordered (h_a,h_b) may sit in several ranks' pair sets;
// mate_pair's set semantics (ref src/LINKS.hpp:370-373) need one owner per pair.  Owner =
// dest_rank(h_a,h_b).  The low half of the mixed hash picks the rank, the high bits pick the slot
// (slot_hash2 >> shift), so an owner's imported pairs still spread over its whole table.
__host__ __device__ __forceinline__ uint32_t dest_rank_of_hash(uint64_t mixed, uint32_t n_ranks)
{
#ifdef __CUDA_ARCH__
  // explicit mul.hi: nvcc 12.9 folded the 64-bit form ((lo32 * n) >> 32) << 2 of the shared-memory
  // index into lo32 * ((n << 2) >> 32) == 0 in pair_bucket_count_kernel (seen in the SASS)
  return __umulhi(uint32_t(mixed), n_ranks);
#else
  return uint32_t((uint64_t(uint32_t(mixed)) * uint64_t(n_ranks)) >> 32);
#endif
}
__device__ __forceinline__ uint32_t dest_rank(uint64_t a, uint64_t b, uint32_t n_ranks)
{
  return dest_rank_of_hash(slot_hash2(a, b), n_ranks);
}

constexpr int kMaxRanks = 256;

// pass 1: records per destination rank (block-local histogram, one global atomic per block and rank)
__global__ void __launch_bounds__(kCtaThreads) pair_bucket_count_kernel(const ulonglong2* __restrict__ list, uint64_t n,
                                                                        uint32_t n_ranks, unsigned long long* __restrict__ counts)
{
  __shared__ uint32_t s_hist[kMaxRanks];
  for (uint32_t i = threadIdx.x; i < n_ranks; i += blockDim.x) s_hist[i] = 0;
  __syncthreads();
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x) {
    const ulonglong2 hk = __ldg(list + i);
    atomicAdd(&s_hist[dest_rank(hk.x, hk.y, n_ranks)], 1u);
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < n_ranks; i += blockDim.x)
    if (s_hist[i]) atomicAdd(counts + i, (unsigned long long)s_hist[i]);
}

// pass 2: scatter {a, b, okey} records into per-rank buckets.  cursors[] start at the bucket
// offsets; every block reserves one range per rank per tile, lanes take block-local offsets.
__global__ void __launch_bounds__(kCtaThreads) pair_bucket_scatter_kernel(PairView pv, uint64_t n, uint32_t n_ranks, int with_okey,
                                                                          unsigned long long* __restrict__ cursors,
                                                                          uint64_t* __restrict__ out3)
{
  __shared__ uint32_t s_hist[kMaxRanks];
  __shared__ unsigned long long s_base[kMaxRanks];
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t n_round = (n + stride - 1) / stride * stride;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    for (uint32_t r = threadIdx.x; r < n_ranks; r += blockDim.x) s_hist[r] = 0;
    __syncthreads();
    ulonglong2 hk = make_ulonglong2(0, 0);
    uint32_t dst = 0, local = 0;
    if (i < n) {
      hk = __ldg(pv.list + i);
      dst = dest_rank(hk.x, hk.y, n_ranks);
      local = atomicAdd(&s_hist[dst], 1u);
    }
    __syncthreads();
    for (uint32_t r = threadIdx.x; r < n_ranks; r += blockDim.x)
      if (s_hist[r]) s_base[r] = atomicAdd(cursors + r, (unsigned long long)s_hist[r]);
    __syncthreads();
    if (i < n) {
      const uint64_t o = (s_base[dst] + local) * 3;
      out3[o] = hk.x; out3[o + 1] = hk.y;
      out3[o + 2] = with_okey ? pair_okey(pv, hk.x, hk.y) : 0ull;
    }
    __syncthreads();
  }
}

// insert received {a, b, okey} records; fresh pairs are appended to the dense list (one global
// atomic per warp).  Same first-wins rule as extract_kernel: minimum order key.
__global__ void __launch_bounds__(kCtaThreads) pair_import_kernel(const uint64_t* __restrict__ rec3, uint64_t n, PairView pv,
                                                                  int multi_d, Counters* __restrict__ ctr)
{
  const int lane = threadIdx.x & 31;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t n_round = (n + stride - 1) / stride * stride; // whole warps stay in the loop
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    bool fresh = false;
    uint64_t a = 0, b = 0;
    if (i < n) {
      a = __ldg(rec3 + i * 3); b = __ldg(rec3 + i * 3 + 1);
      fresh = pair_insert(pv, a, b, __ldg(rec3 + i * 3 + 2), multi_d != 0);
    }
    const uint32_t m = __ballot_sync(0xffffffffu, fresh);
    if (m) {
      unsigned long long base = 0;
      if (lane == __ffs(m) - 1) base = atomicAdd(&ctr->n_unique, (unsigned long long)__popc(m));
      base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
      if (fresh) pv.list[base + __popc(m & ((1u << lane) - 1u))] = make_ulonglong2(a, b);
    }
  }
}

// export the dense pair list with order keys: out[] = {a, b, okey}
__global__ void pair_export_kernel(PairView pv, uint64_t n_pairs, uint64_t* __restrict__ out3)
{
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_pairs; i += uint64_t(gridDim.x) * blockDim.x) {
    const ulonglong2 hk = pv.list[i];
    out3[i * 3] = hk.x; out3[i * 3 + 1] = hk.y; out3[i * 3 + 2] = pair_okey(pv, hk.x, hk.y);
  }
}

} // namespace links

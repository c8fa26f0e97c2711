// links_b200/csrc/nthash.cuh -- ntHash2 on 2-bit packed bases (device) + table generation (host).
//
// Restates btllib's ntHash2 (reference btllib/include/btllib/nthash.hpp):
//   fwd(kmer) = XOR_i srol^(k-1-i)(seed[c_i])        (:897-930 base loop; rolling form :576-589)
//   rev(kmer) = XOR_i srol^(i)(seed[comp(c_i)])      (:591-603)
// where srol = rol1 + swapbits033 (:443-447,480-486) rotates the low 33 and high 31 bits of the
// word independently.  Because the hash is XOR-linear in per-position contributions, a k-mer is
// hashed from scratch with k table lookups -- no serial dependence between neighbouring
// positions, which is what lets one thread own one (read, distance, rank) candidate.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <vector>

namespace links {

// 2-bit code = (ascii >> 1) & 3 : A/a=0 C/c=1 T/t/U/u=2 G/g=3 ; complement = code ^ 2
__host__ __device__ __forceinline__ uint32_t base_code(uint32_t c) { return (c >> 1) & 3u; }
__host__ __device__ __forceinline__ bool base_valid(uint32_t c)
{
  // SEED_TAB[c] != 0 for ACGTU/acgtu (nthash.hpp:38-71); control bytes 1,3,4,5,7 are out of domain
  const uint32_t u = c & 0xDFu;
  return u == 'A' || u == 'C' || u == 'G' || u == 'T' || u == 'U';
}

constexpr uint64_t kSeedA = 0x3c8bfbb395c60474ULL; // nthash.hpp:30-33
constexpr uint64_t kSeedC = 0x3193c18562a02b4cULL;
constexpr uint64_t kSeedG = 0x20323ed082572324ULL;
constexpr uint64_t kSeedT = 0x295549f54be24456ULL;
constexpr uint64_t kMultiSeed = 0x90b45d39fb6da1faULL; // nthash.hpp:27
constexpr int kMultiShift = 27;                        // nthash.hpp:24

inline uint64_t srol_n_host(uint64_t v, unsigned n)
{
  uint64_t hi = v >> 33, lo = v & 0x1FFFFFFFFULL;
  const unsigned a = n % 31, b = n % 33;
  if (a) hi = ((hi << a) | (hi >> (31 - a))) & 0x7FFFFFFFULL;
  if (b) lo = ((lo << b) | (lo >> (33 - b))) & 0x1FFFFFFFFULL;
  return (hi << 33) | lo;
}

// tab[i*4 + code] = { contribution of base `code` at k-mer offset i to fwd , to rev }
inline void build_nthash_table(unsigned k, ulonglong2* tab)
{
  const uint64_t seed_by_code[4] = {kSeedA, kSeedC, kSeedT, kSeedG};
  for (unsigned i = 0; i < k; ++i)
    for (unsigned c = 0; c < 4; ++c) {
      tab[i * 4 + c].x = srol_n_host(seed_by_code[c], k - 1 - i);
      tab[i * 4 + c].y = srol_n_host(seed_by_code[c ^ 2], i);
    }
}

// Hash the k-mer starting at global base index g of a 2-bit packed stream (16 bases per u32,
// base j of a word at bits 2j..2j+1).  `tab` is the shared-memory copy of the table.  Reads one
// word past the last needed one (buffers are padded).
__device__ __forceinline__ void hash_kmer(const uint32_t* __restrict__ packed, uint64_t g, int k,
                                          const ulonglong2* __restrict__ tab, uint64_t& fwd, uint64_t& rev)
{
  const uint64_t w = g >> 4;
  const int sh = int(g & 15) * 2;
  uint32_t cur = __ldg(packed + w);
  uint64_t f = 0, r = 0;
  for (int i = 0; i < k; i += 16) {
    const uint32_t nxt = __ldg(packed + w + 1 + (i >> 4));
    const uint32_t win = __funnelshift_r(cur, nxt, sh);
    const int nb = min(16, k - i);
#pragma unroll
    for (int b = 0; b < 16; ++b) {
      if (b < nb) {
        const ulonglong2 e = tab[(i + b) * 4 + ((win >> (2 * b)) & 3u)];
        f ^= e.x;
        r ^= e.y;
      }
    }
    cur = nxt;
  }
  fwd = f;
  rev = r;
}

// ---- rolling by one base (contig-side kernels: a thread walks a few consecutive positions).  The hash is XOR-linear in
// per-position contributions, so  fwd' = srol(fwd ^ c_out(0)) ^ c_in(k-1)  and  rev' = sror(rev ^ c_out(0)) ^ c_in(k-1)
// with c(i) the table entries of build_nthash_table (btllib rolls the same way, nthash.hpp:576-603).  A base that is not
// ACGT enters and leaves with the same (arbitrary) 2-bit code, so a window that has moved past it hashes exactly as from
// scratch; windows that still hold it are skipped by the caller (window_valid).
__host__ __device__ __forceinline__ uint64_t srol1(uint64_t v)
{
  const uint64_t hi = v >> 33, lo = v & 0x1FFFFFFFFULL;
  return ((((hi << 1) | (hi >> 30)) & 0x7FFFFFFFULL) << 33) | (((lo << 1) | (lo >> 32)) & 0x1FFFFFFFFULL);
}
__host__ __device__ __forceinline__ uint64_t sror1(uint64_t v)
{
  const uint64_t hi = v >> 33, lo = v & 0x1FFFFFFFFULL;
  return ((((hi >> 1) | (hi << 30)) & 0x7FFFFFFFULL) << 33) | (((lo >> 1) | (lo << 32)) & 0x1FFFFFFFFULL);
}
__device__ __forceinline__ uint32_t packed_code(const uint32_t* __restrict__ packed, uint64_t g)
{
  return (__ldg(packed + (g >> 4)) >> (int(g & 15) * 2)) & 3u;
}
// (fwd, rev) of the k-mer at g  ->  of the k-mer at g + 1
__device__ __forceinline__ void roll_kmer(const uint32_t* __restrict__ packed, uint64_t g, int k, const ulonglong2* __restrict__ tab,
                                          uint64_t& fwd, uint64_t& rev)
{
  const ulonglong2 out = tab[packed_code(packed, g)];                        // offset 0 of the old window
  const ulonglong2 in = tab[(k - 1) * 4 + packed_code(packed, g + k)];       // offset k-1 of the new window
  fwd = srol1(fwd ^ out.x) ^ in.x;
  rev = sror1(rev ^ out.y) ^ in.y;
}

// ---- two bases per lookup.  Because the hash is XOR-linear, the contributions of k-mer offsets
// 2p and 2p+1 can be pre-XORed: tab2[p*16 + c4] with c4 = code(2p) | code(2p+1)<<2 (the 4 bits
// the packed stream already holds).  Halves the shared-memory loads and the XORs per k-mer.
// For odd k the last entry ignores its high two bits (the base after the k-mer).
inline unsigned nthash_pair_count(unsigned k) { return (k + 1) / 2; }
inline void build_nthash_pair_table(unsigned k, ulonglong2* tab2)
{
  std::vector<ulonglong2> one(size_t(k) * 4);
  build_nthash_table(k, one.data());
  for (unsigned p = 0; p < nthash_pair_count(k); ++p)
    for (unsigned c4 = 0; c4 < 16; ++c4) {
      ulonglong2 e = one[(2 * p) * 4 + (c4 & 3)];
      if (2 * p + 1 < k) {
        const ulonglong2 e2 = one[(2 * p + 1) * 4 + (c4 >> 2)];
        e.x ^= e2.x;
        e.y ^= e2.y;
      }
      tab2[p * 16 + c4] = e;
    }
}

__device__ __forceinline__ void hash_kmer2(const uint32_t* __restrict__ packed, uint64_t g, int k,
                                           const ulonglong2* __restrict__ tab2, uint64_t& fwd, uint64_t& rev)
{
  const uint64_t w = g >> 4;
  const int sh = int(g & 15) * 2;
  const int npairs = (k + 1) >> 1;
  uint32_t cur = __ldg(packed + w);
  uint64_t f = 0, r = 0;
  for (int p0 = 0; p0 < npairs; p0 += 8) {
    const uint32_t nxt = __ldg(packed + w + 1 + (p0 >> 3));
    const uint32_t win = __funnelshift_r(cur, nxt, sh);
    const ulonglong2* t = tab2 + p0 * 16;
    if (npairs - p0 >= 8) {
#pragma unroll
      for (int p = 0; p < 8; ++p) {
        const ulonglong2 e = t[p * 16 + ((win >> (4 * p)) & 15u)];
        f ^= e.x;
        r ^= e.y;
      }
    } else {
      for (int p = 0; p < npairs - p0; ++p) {
        const ulonglong2 e = t[p * 16 + ((win >> (4 * p)) & 15u)];
        f ^= e.x;
        r ^= e.y;
      }
    }
    cur = nxt;
  }
  fwd = f;
  rev = r;
}

// Same hash from the table split into its forward and reverse halves (8-byte entries).  A 16-entry table of 8-byte
// entries spans all 32 banks exactly once, so a half-warp's LDS.64 is conflict-free whatever the bases are; with
// 16-byte entries two of the 16 entries share each bank group and the lookups ran at 1.8x their ideal wavefronts
// (ncu: 60 M shared wavefronts per launch against 33 M ideal), on the L1/shared pipe that bounds extract_kernel.
__device__ __forceinline__ void hash_kmer2_split(const uint32_t* __restrict__ packed, uint64_t g, int k,
                                                 const uint64_t* __restrict__ tf, const uint64_t* __restrict__ tr, uint64_t& fwd,
                                                 uint64_t& rev)
{
  const uint64_t w = g >> 4;
  const int sh = int(g & 15) * 2;
  const int npairs = (k + 1) >> 1;
  uint32_t cur = __ldg(packed + w);
  uint64_t f = 0, r = 0;
  for (int p0 = 0; p0 < npairs; p0 += 8) {
    const uint32_t nxt = __ldg(packed + w + 1 + (p0 >> 3));
    const uint32_t win = __funnelshift_r(cur, nxt, sh);
    const uint64_t* af = tf + p0 * 16;
    const uint64_t* ar = tr + p0 * 16;
    if (npairs - p0 >= 8) {
#pragma unroll
      for (int p = 0; p < 8; ++p) {
        const uint32_t e = p * 16 + ((win >> (4 * p)) & 15u);
        f ^= af[e];
        r ^= ar[e];
      }
    } else {
      for (int p = 0; p < npairs - p0; ++p) {
        const uint32_t e = p * 16 + ((win >> (4 * p)) & 15u);
        f ^= af[e];
        r ^= ar[e];
      }
    }
    cur = nxt;
  }
  fwd = f;
  rev = r;
}

// true iff none of the bases g .. g+k-1 is flagged in the invalid-base bitmap (32 bases per u32)
__device__ __forceinline__ bool window_valid(const uint32_t* __restrict__ inv, uint64_t g, int k)
{
  uint64_t w = g >> 5;
  int sh = int(g & 31);
  int left = k;
  uint32_t cur = __ldg(inv + w);
  while (left > 0) {
    const uint32_t nxt = __ldg(inv + w + 1);
    const uint32_t win = __funnelshift_r(cur, nxt, sh);
    const uint32_t mask = left >= 32 ? 0xFFFFFFFFu : ((1u << left) - 1u);
    if (win & mask) return false;
    left -= 32;
    ++w;
    cur = nxt;
  }
  return true;
}

// hashes()[i] of the canonical multi-hash (nthash.hpp:920-927)
__device__ __forceinline__ uint64_t multihash_i(uint64_t h0, uint64_t kseed /* k*MULTISEED */, uint32_t i)
{
  if (i == 0) return h0;
  uint64_t t = h0 * (uint64_t(i) ^ kseed);
  t ^= t >> kMultiShift;
  return t;
}

// Bloom filter view: bit n of the byte array = bit (n & 31) of little-endian word n >> 5
// (btllib BloomFilter::insert/contains, bloom_filter.hpp:639-659).
struct BloomView {
  uint32_t* words;
  uint64_t array_bits; // bytes * 8
  uint64_t magic;      // floor(2^64 / array_bits)
  uint32_t hash_num;
  uint32_t bits32;     // array_bits when it is below 2^31 (filters under 256 MB), else 0
};

// SMALL: the caller knows bf.bits32 != 0 (array_bits < 2^31, i.e. a filter under 256 MB).  A compile-time choice, not a
// branch: with both versions inlined at every probe site extract_kernel outgrew the instruction cache and ran 20 % slower.
template <bool SMALL = false>
__device__ __forceinline__ uint64_t bloom_mod(uint64_t h, const BloomView& bf)
{
  // exact h % array_bits: q = floor(h*magic/2^64) is floor(h/d) or one less
  const uint64_t q = __umul64hi(h, bf.magic);
  if (SMALL) {
    // h - q*d lies in [0, 2d) and 2d < 2^32: the low words alone give it exactly (a third of the 64-bit version's instructions)
    uint32_t r = uint32_t(h) - uint32_t(q) * bf.bits32;
    if (r >= bf.bits32) r -= bf.bits32;
    return r;
  }
  uint64_t r = h - q * bf.array_bits;
  if (r >= bf.array_bits) r -= bf.array_bits;
  return r;
}

__device__ __forceinline__ bool bloom_contains(const BloomView& bf, uint64_t fwd, uint64_t rev, uint64_t kseed,
                                               uint32_t& probes)
{
  const uint64_t h0 = rev < fwd ? rev : fwd;
  for (uint32_t i = 0; i < bf.hash_num; ++i) {
    const uint64_t n = bloom_mod(multihash_i(h0, kseed, i), bf);
    ++probes;
    if (!((__ldg(bf.words + (n >> 5)) >> (n & 31)) & 1u)) return false;
  }
  return true;
}

template <bool SMALL = false>
__device__ __forceinline__ uint32_t bloom_bit(const BloomView& bf, uint64_t h0, uint64_t kseed, uint32_t i)
{
  const uint64_t n = bloom_mod<SMALL>(multihash_i(h0, kseed, i), bf);
  return (__ldg(bf.words + (n >> 5)) >> (n & 31)) & 1u;
}

// N probes starting at i0, N known at compile time, issued together (independent loads in flight, no predicated-off
// bodies) instead of the reference's one-at-a-time early-exit loop (bloom_filter.hpp:648-659).  Same answer; `probes`
// is advanced by the number of probes the sequential loop would have made (roofline accounting).
template <int N, bool SMALL = false>
__device__ __forceinline__ bool bloom_probe_fixed(const BloomView& bf, uint64_t h0, uint64_t kseed, uint32_t i0, uint32_t& probes)
{
  if (N <= 0) return true;
  uint32_t miss = 0;
#pragma unroll
  for (int j = 0; j < N; ++j) miss |= (bloom_bit<SMALL>(bf, h0, kseed, i0 + j) ^ 1u) << j;
  if (miss) { probes += __ffs(miss); return false; }
  probes += N;
  return true;
}

// runtime count, groups of four then singles (generic hash_num)
template <bool SMALL = false>
__device__ __forceinline__ bool bloom_probe_loop(const BloomView& bf, uint64_t h0, uint64_t kseed, uint32_t i0, uint32_t i1,
                                                 uint32_t& probes)
{
  while (i0 + 4 <= i1) {
    if (!bloom_probe_fixed<4, SMALL>(bf, h0, kseed, i0, probes)) return false;
    i0 += 4;
  }
  for (; i0 < i1; ++i0) {
    ++probes;
    if (!bloom_bit<SMALL>(bf, h0, kseed, i0)) return false;
  }
  return true;
}

__device__ __forceinline__ void bloom_insert(const BloomView& bf, uint64_t fwd, uint64_t rev, uint64_t kseed)
{
  const uint64_t h0 = rev < fwd ? rev : fwd;
  for (uint32_t i = 0; i < bf.hash_num; ++i) {
    const uint64_t n = bloom_mod(multihash_i(h0, kseed, i), bf);
    atomicOr(bf.words + (n >> 5), 1u << (n & 31));
  }
}

// the same insert restricted to the filter words [w_lo, w_hi) (contig_kernel's BloomPass)
__device__ __forceinline__ void bloom_insert_range(const BloomView& bf, uint64_t fwd, uint64_t rev, uint64_t kseed, uint64_t w_lo,
                                                   uint64_t w_hi)
{
  const uint64_t h0 = rev < fwd ? rev : fwd;
  for (uint32_t i = 0; i < bf.hash_num; ++i) {
    const uint64_t n = bloom_mod(multihash_i(h0, kseed, i), bf);
    const uint64_t w = n >> 5;
    if (w >= w_lo && w < w_hi) atomicOr(bf.words + w, 1u << (n & 31));
  }
}

} // namespace links

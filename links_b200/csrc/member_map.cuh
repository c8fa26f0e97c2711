// links_b200/csrc/member_map.cuh -- "is this read k-mer in the Bloom filter?" answered ONCE per sampled read position.
//
// The reference asks the filter per candidate pair (src/LINKS.hpp:554-555: contains(A) && contains(B)).  But the k-mers
// a read contributes are few: for distance i the A k-mers sit at x = n, the B k-mers at x = n + d_i (x = position - offset,
// n = t_i-1, 2t_i-1, ...; src/LINKS.hpp:525-543), so with several -d values -- how LINKS is normally run: the installation
// test has 15, BASELINE's configs[2] three -- the same position is asked about again and again: as A of every distance whose
// step divides it, and as B of those again d_i further left.  Here every position that ANY (distance, role) can use is
// tested once, and the answer is kept in a byte map; extract_kernel then needs two byte loads per candidate instead of
// two filter walks.  (15 distances with one step: 1 filter walk per sampled position instead of ~22.)
//
//   which positions : x is needed iff for some i (x+1) % t_i == 0 [A] or (x+1-d_i) % t_i == 0 [B].  That is periodic with
//                     T = lcm(t_i); the host builds need_idx[T] (-1, or the dense index 0..c-1 of the residue among the
//                     needed ones) and need_res[c] (its inverse).  Entry e of a read <-> x = (e / c) * T + need_res[e % c].
//   the map         : u8 per entry, read r's entries start at mm_base[r]:  bit 7 = passes BloomFilter::contains,
//                     bits 0-6 = bit tests the reference's early-exit loop makes for this k-mer (bloom_filter.hpp:648-659),
//                     so that n_bf_probes stays the exact per-candidate count of the reference's loop.
// Only reads WITHOUT an invalid base are mapped (their k-mer positions are the closed form above); flagged reads pair by
// rank, not position (nthash.hpp:1493-1533), and keep extract_kernel's filter-walking stages.  The host chooses per chunk
// (links_gpu.cu map_pays): single-distance jobs on short reads walk the filter less often than the map would.
#pragma once
#include "kernels.cuh"

namespace links {

constexpr int kMapTile = 1024;      // map entries per scheduling tile
constexpr uint32_t kMapMaxPeriod = 4096;  // lcm of the step sizes beyond which the map is not used

struct MemberMap {
  uint8_t* bytes;                  // nullptr: no map for this launch
  const unsigned long long* base;  // [n_seq + 1] first entry of each read
  const int16_t* need_idx;         // [T]
  const uint16_t* need_res;        // [c]
  uint32_t T, c;
  uint64_t magic_T, magic_c;       // ceil(2^64 / T), ceil(2^64 / c): x / T == umul64hi(x, magic_T) exactly for x < 2^32 (T > 1)
};

// divisions by the period and by the entries per period happen once or twice per candidate: a 64-bit hardware division there
// (~100 instructions) cost more than the filter walks the map saves
__host__ __device__ inline uint64_t map_magic(uint32_t d) { return d > 1 ? ~0ull / d + 1 : 0; }
__device__ __forceinline__ uint32_t map_div(uint32_t x, uint32_t d, uint64_t magic) { return d > 1 ? uint32_t(__umul64hi(uint64_t(x), magic)) : x; }

struct MapBatch {
  const uint32_t* packed; const uint64_t* off; uint64_t base0;
  const uint32_t* seg_tile_start;  // [n_seq + 1] cumulative map tiles per read
  uint32_t n_seq, n_tiles;
  const uint32_t* flags;
};

__device__ __forceinline__ uint64_t map_entry_of(const MemberMap& mm, uint64_t x64)  // x < 2^32 (reads are shorter)
{
  const uint32_t x = uint32_t(x64), p = map_div(x, mm.T, mm.magic_T);
  return uint64_t(p) * mm.c + uint64_t(__ldg(mm.need_idx + (x - p * mm.T)));
}

// one thread per map entry
template <int H, bool SMALL>
__global__ void __launch_bounds__(kCtaThreads) member_map_kernel(MapBatch mb, MemberMap mm, int k, uint64_t offset, uint64_t kseed,
                                                                 const ulonglong2* __restrict__ g_tab2, BloomView bf, MemberView mv,
                                                                 Counters* __restrict__ ctr)
{
  extern __shared__ ulonglong2 s_dyn[];
  const int n_tab = ((k + 1) >> 1) * 16;
  uint64_t* s_tf = reinterpret_cast<uint64_t*>(s_dyn);
  uint64_t* s_tr = s_tf + n_tab;
  for (int i = threadIdx.x; i < n_tab; i += blockDim.x) { const ulonglong2 e = g_tab2[i]; s_tf[i] = e.x; s_tr[i] = e.y; }
  __shared__ uint32_t s_seg;
  __syncthreads();
  const uint32_t hn = H ? uint32_t(H) : bf.hash_num;
  const bool has_set = mv.buckets != nullptr;
  uint32_t n_load = 0, n_msec = 0;
  for (uint32_t tile = blockIdx.x; tile < mb.n_tiles; tile += gridDim.x) {
    if (threadIdx.x == 0) s_seg = find_segment(mb.seg_tile_start, mb.n_seq, tile);
    __syncthreads();
    const uint32_t r = s_seg;
    __syncthreads();
    if (mb.flags[r]) continue;  // (a flagged read owns tiles too: the host does not know the flags)
    const uint64_t g0 = mb.off[r] - mb.base0, len = mb.off[r + 1] - mb.off[r];
    const uint64_t npos = len - k + 1;                       // reads shorter than k own zero tiles
    const uint64_t X = npos > offset ? npos - offset : 0;    // x = position - offset in [0, X)
    const uint64_t n_entries = mm.base[r + 1] - mm.base[r];
    const uint64_t e_begin = uint64_t(tile - mb.seg_tile_start[r]) * kMapTile;
    uint8_t* out = mm.bytes + mm.base[r];
    for (int q = threadIdx.x; q < kMapTile; q += kCtaThreads) {
      const uint64_t e = e_begin + q;
      if (e >= n_entries) break;
      const uint32_t p = map_div(uint32_t(e), mm.c, mm.magic_c);  // e < 2^32: a read has fewer entries than bases
      const uint64_t x = uint64_t(p) * mm.T + __ldg(mm.need_res + (uint32_t(e) - p * mm.c));
      uint8_t v = 0;
      if (x < X) {
        uint64_t f, rr;
        hash_kmer2_split(mb.packed, g0 + offset + x, k, s_tf, s_tr, f, rr);
        const uint64_t h0 = rr < f ? rr : f;
        bool member = false;
        if (has_set) member = member_contains(mv, h0, n_msec);
        if (member) v = uint8_t(0x80u | hn);  // the filter holds every set key: all hash_num bits are set
        else {
          // the reference's own loop, one bit at a time with its early exit (bloom_filter.hpp:648-659).  A k-mer that is not in
          // the filter fails each bit with probability ~1/2, so this reads ~2 sectors where testing bits in groups reads 3-4;
          // random 32-byte sectors are what the DRAM delivers least of (tools/gather_peak.cu: 43 G/s on this GPU), and with one
          // thread per entry and no queues there are enough warps in flight to cover the extra round trips
          uint32_t np = 0;
          bool pass = true;
          for (uint32_t i = 0; i < hn; ++i) {
            ++np;
            if (!bloom_bit<SMALL>(bf, h0, kseed, i)) { pass = false; break; }
          }
          n_load += np;
          v = uint8_t((pass ? 0x80u : 0u) | np);
        }
      }
      out[e] = v;
    }
  }
  unsigned long long a = warp_sum_ull(n_load), b = warp_sum_ull(n_msec);
  if ((threadIdx.x & 31) == 0) {
    if (a) atomicAdd(&ctr->n_bf_loads, a);
    if (b) atomicAdd(&ctr->n_member_sectors, b);
  }
}

}  // namespace links

// links_b200/csrc/device_prims.cuh -- device-wide primitives written for this path:
// exclusive scan (u32), LSD radix sort of (u64 key, i32 payload), reduce-by-key.
// Used by stage 4's link aggregation (reference src/LINKS.hpp:699-762 does this with nested
// unordered_maps; here it is a sort + segmented reduction over packed row keys).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace links {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems; // 2048 elements per block

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v)
{
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// exclusive scan of one value per thread across a 256-thread block; returns exclusive prefix,
// *total gets the block sum (same for all threads).  smem: 8 u32 + 1.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total, uint32_t* smem /* >= 9 */)
{
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t incl = warp_incl_scan(v);
  if (lane == 31) smem[w] = incl;
  __syncthreads();
  if (w == 0) {
    uint32_t s = lane < (kScanThreads / 32) ? smem[lane] : 0;
    const uint32_t si = warp_incl_scan(s);
    if (lane < (kScanThreads / 32)) smem[lane] = si - s;
    if (lane == (kScanThreads / 32) - 1) smem[8] = si;
  }
  __syncthreads();
  const uint32_t r = smem[w] + incl - v;
  *total = smem[8];
  __syncthreads();
  return r;
}

// phase 1: per-block sums
__global__ void __launch_bounds__(kScanThreads) scan_sums_kernel(const uint32_t* __restrict__ in, uint64_t n,
                                                                 uint32_t* __restrict__ sums)
{
  __shared__ uint32_t sm[9];
  const uint64_t base = uint64_t(blockIdx.x) * kScanTile + uint64_t(threadIdx.x) * kScanItems;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i)
    if (base + i < n) s += in[base + i];
  uint32_t total;
  block_excl_scan(s, &total, sm);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// phase 3: rescan each tile with its carry-in (sums already exclusive-scanned); in place allowed
__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint32_t* __restrict__ in, uint64_t n,
                                                                  const uint32_t* __restrict__ sums,
                                                                  uint32_t* __restrict__ out)
{
  __shared__ uint32_t sm[9];
  const uint64_t base = uint64_t(blockIdx.x) * kScanTile + uint64_t(threadIdx.x) * kScanItems;
  uint32_t v[kScanItems];
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = base + i < n ? in[base + i] : 0;
    s += v[i];
  }
  uint32_t total;
  uint32_t run = block_excl_scan(s, &total, sm) + (sums ? sums[blockIdx.x] : 0);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = run;
    run += v[i];
  }
}

// single-block scan for small arrays (n <= kScanTile): writes exclusive scan and the total
__global__ void __launch_bounds__(kScanThreads) scan_small_kernel(const uint32_t* __restrict__ in, uint32_t n,
                                                                  uint32_t* __restrict__ out, uint32_t* __restrict__ total_out)
{
  __shared__ uint32_t sm[9];
  const uint32_t base = threadIdx.x * kScanItems;
  uint32_t v[kScanItems];
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = base + i < n ? in[base + i] : 0;
    s += v[i];
  }
  uint32_t total;
  uint32_t run = block_excl_scan(s, &total, sm);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = run;
    run += v[i];
  }
  if (threadIdx.x == 0 && total_out) *total_out = total;
}

// ---------------------------------------------------------------- radix sort (one 8-bit pass)

constexpr int kSortThreads = 256;
constexpr int kSortRounds = 8;
constexpr int kSortTile = kSortThreads * kSortRounds; // 2048 items per block
constexpr int kRadix = 256;

// hist[digit * n_blocks + block] = count of digit in this block's tile
__global__ void __launch_bounds__(kSortThreads) radix_hist_kernel(const uint64_t* __restrict__ keys, uint64_t n, int shift,
                                                                  uint32_t* __restrict__ hist, uint32_t n_blocks)
{
  __shared__ uint32_t h[kRadix];
  h[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t tile0 = uint64_t(blockIdx.x) * kSortTile;
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const uint64_t i = tile0 + uint64_t(r) * kSortThreads + threadIdx.x;
    if (i < n) atomicAdd(&h[(keys[i] >> shift) & (kRadix - 1)], 1u);
  }
  __syncthreads();
  hist[uint64_t(threadIdx.x) * n_blocks + blockIdx.x] = h[threadIdx.x];
}

// Stable scatter.  Warp w owns items [tile0 + w*256, +256) visited in 8 rounds of 32 consecutive
// items, so index order == (warp, round, lane) order and ranks computed per warp are stable.
__global__ void __launch_bounds__(kSortThreads) radix_scatter_kernel(const uint64_t* __restrict__ keys_in,
                                                                     const int32_t* __restrict__ vals_in, uint64_t n,
                                                                     int shift, const uint32_t* __restrict__ hist_scanned,
                                                                     uint32_t n_blocks, uint64_t* __restrict__ keys_out,
                                                                     int32_t* __restrict__ vals_out)
{
  __shared__ uint32_t cnt[kSortThreads / 32][kRadix];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (kSortThreads / 32) * kRadix; i += kSortThreads) (&cnt[0][0])[i] = 0;
  __syncthreads();
  const uint64_t warp0 = uint64_t(blockIdx.x) * kSortTile + uint64_t(w) * (32 * kSortRounds);
  uint64_t key[kSortRounds];
  int32_t val[kSortRounds];
  uint32_t rank[kSortRounds];
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const uint64_t i = warp0 + uint64_t(r) * 32 + lane;
    const bool ok = i < n;
    key[r] = ok ? keys_in[i] : 0;
    val[r] = ok ? vals_in[i] : 0;
    const uint32_t d = ok ? uint32_t((key[r] >> shift) & (kRadix - 1)) : 0xFFFFFFFFu;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    uint32_t base = 0;
    if (ok) base = cnt[w][d];
    __syncwarp();
    if (ok && lane == (__ffs(peers) - 1)) cnt[w][d] = base + __popc(peers);
    __syncwarp();
    rank[r] = base + __popc(peers & ((1u << lane) - 1u));
  }
  __syncthreads();
  {
    // exclusive prefix over warps for digit = threadIdx.x, plus the global base of this block
    const uint32_t d = threadIdx.x;
    uint32_t run = hist_scanned[uint64_t(d) * n_blocks + blockIdx.x];
#pragma unroll
    for (int ww = 0; ww < kSortThreads / 32; ++ww) {
      const uint32_t t = cnt[ww][d];
      cnt[ww][d] = run;
      run += t;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const uint64_t i = warp0 + uint64_t(r) * 32 + lane;
    if (i < n) {
      const uint32_t d = uint32_t((key[r] >> shift) & (kRadix - 1));
      const uint32_t dst = cnt[w][d] + rank[r];
      keys_out[dst] = key[r];
      vals_out[dst] = val[r];
    }
  }
}

// ---------------------------------------------------------------- reduce by key (sorted input)

__global__ void head_flags_kernel(const uint64_t* __restrict__ keys, uint64_t n, uint32_t* __restrict__ flags)
{
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) flags[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}

constexpr int kReduceItems = 16;
// seg_excl = exclusive scan of head flags; segment of item i = seg_excl[i] + flag[i] - 1.
// Each thread walks kReduceItems consecutive items, accumulating runs locally, one atomic per run.
__global__ void reduce_by_key_kernel(const uint64_t* __restrict__ keys, const int32_t* __restrict__ vals, uint64_t n,
                                     const uint32_t* __restrict__ seg_excl, uint64_t* __restrict__ out_keys,
                                     unsigned long long* __restrict__ out_links, unsigned long long* __restrict__ out_gaps)
{
  const uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint64_t i0 = t * kReduceItems;
  if (i0 >= n) return;
  const uint64_t i1 = i0 + kReduceItems < n ? i0 + kReduceItems : n;
  uint64_t cur_key = keys[i0];
  const bool head0 = (i0 == 0 || keys[i0 - 1] != cur_key);
  uint32_t seg = seg_excl[i0] + (head0 ? 1u : 0u) - 1u;
  if (head0) out_keys[seg] = cur_key;
  unsigned long long links = 0;
  long long gaps = 0;
  for (uint64_t i = i0; i < i1; ++i) {
    const uint64_t kk = keys[i];
    if (kk != cur_key) {
      atomicAdd(out_links + seg, links);
      atomicAdd(out_gaps + seg, (unsigned long long)gaps);
      ++seg;
      cur_key = kk;
      out_keys[seg] = kk;
      links = 0;
      gaps = 0;
    }
    links += 1;
    gaps += vals[i];
  }
  atomicAdd(out_links + seg, links);
  atomicAdd(out_gaps + seg, (unsigned long long)gaps);
}

// Sum already-aggregated rows that share a key (merging the partial link tables of several
// ranks): item i of the sorted order adds in_links/in_gaps[idx[i]] to its segment.
__global__ void rows_segment_sum_kernel(const uint64_t* __restrict__ keys, const int32_t* __restrict__ idx, uint64_t n,
                                        const uint32_t* __restrict__ seg_excl, const unsigned long long* __restrict__ in_links,
                                        const unsigned long long* __restrict__ in_gaps, uint64_t* __restrict__ out_keys,
                                        unsigned long long* __restrict__ out_links, unsigned long long* __restrict__ out_gaps)
{
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t key = keys[i];
  const bool head = (i == 0 || keys[i - 1] != key);
  const uint32_t seg = seg_excl[i] + (head ? 1u : 0u) - 1u;
  if (head) out_keys[seg] = key;
  atomicAdd(out_links + seg, in_links[idx[i]]);
  atomicAdd(out_gaps + seg, in_gaps[idx[i]]);
}

} // namespace links

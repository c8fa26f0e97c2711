// links_b200/csrc/exchange.cuh -- multi-GPU pair exchange over NVLink peer memory.
//
// The reference keeps mate_pair in one address space (src/LINKS.hpp:138-147,354-378).  With the
// reads sharded over ranks, every distinct ordered (h_a,h_b) must reach exactly one OWNER rank
// before links are counted (set semantics, minimum order key = first distance wins).  Instead of
// "export -> NCCL all-to-all -> import" after the last read, the exchange rides along with the
// extraction, chunk by chunk, through memory every rank can address (NVLink P2P / NVSwitch):
//
//   route_kernel    after each extract launch: the pairs that launch found new are written
//                   straight into their owner's INBOX (a peer pointer), one region per source rank
//   commit_kernel   publishes "records [0,n) of my region on rank d are complete" (epoch-tagged)
//   inbox_snapshot_kernel + inbox_import_kernel
//                   on the owner, on a side stream, while it keeps extracting: insert whatever
//                   has been committed so far into the owned pair set
//   inbox_wait_kernel  one block spins until every source has published "done" for this job
//   rows_send_kernel   the per-rank partial link rows go to rank 0's inbox the same way
//
// Inbox layout (identical on every rank; two halves used by alternating jobs, so a fast rank that
// already started the next job never writes where a slow rank is still reading):
//   half h:  [InboxHeader | pair regions: n_ranks x pair_cap records | row regions: n_ranks x row_cap x 24 B]
// Words in the header are tagged (epoch << 40) | count and only ever written by their source rank.
#pragma once
#include "kernels.cuh"

namespace links {

constexpr int kMaxP2PRanks = 64;
constexpr uint64_t kCountMask = (1ull << 40) - 1;

struct InboxHeader {
  unsigned long long committed[kMaxP2PRanks];  // [src] pair records complete in src's region
  unsigned long long done[kMaxP2PRanks];       // [src] epoch of the last job src finished routing
  unsigned long long rows[kMaxP2PRanks];       // [src] canonical link rows complete in src's row region
};

struct RowRec { uint64_t key; unsigned long long links, gaps; };

// what one rank knows about the exchange (device copy lives in ExchangeDev)
struct InboxLayout {
  uint64_t half_bytes, pair_off, row_off, pair_cap, row_cap;  // offsets inside a half; caps in records
  uint32_t rec_words;                                         // 2: {a,b}   3: {a,b,okey}
  uint32_t n_ranks, rank;
};

__host__ __device__ __forceinline__ char* inbox_half(char* base, const InboxLayout& L, uint64_t epoch)
{
  return base + (epoch & 1ull) * L.half_bytes;
}
__host__ __device__ __forceinline__ uint64_t* inbox_pair_region(char* base, const InboxLayout& L, uint64_t epoch, uint32_t src)
{
  return reinterpret_cast<uint64_t*>(inbox_half(base, L, epoch) + L.pair_off) + uint64_t(src) * L.pair_cap * L.rec_words;
}
__host__ __device__ __forceinline__ RowRec* inbox_row_region(char* base, const InboxLayout& L, uint64_t epoch, uint32_t src)
{
  return reinterpret_cast<RowRec*>(inbox_half(base, L, epoch) + L.row_off) + uint64_t(src) * L.row_cap;
}

// device-resident state of the exchange on one rank
struct ExchangeDev {
  char* inbox[kMaxP2PRanks];                    // inbox[r] = rank r's buffer as addressable from here
  unsigned long long cursor[kMaxP2PRanks];      // records routed to rank d so far (this job)
  unsigned long long imported[kMaxP2PRanks];    // records of src already inserted into the owned set
  unsigned long long snap[kMaxP2PRanks];        // upper bound of the import in flight
  unsigned long long routed;                    // prefix of the dense unique-pair list already routed
  unsigned long long route_hi;                  // list length when the last extract launch ended (copied in stream order)
  unsigned long long n_owned;                   // distinct pairs in the owned set
  unsigned long long route_overflow, owned_overflow;
  unsigned int blocks_done, row_blocks_done;    // last-block-publishes counters of route_kernel / rows_send_kernel
  unsigned int pad_[2];
  unsigned long long timed_out;                 // a wait below gave up (peer dead or never called in): the job fails, the GPU does not hang
};

__device__ __forceinline__ unsigned long long ld_sys(const unsigned long long* p)
{
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns()
{
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void st_sys(unsigned long long* p, unsigned long long v)
{
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// "last block publishes": every thread fences its own peer writes to system scope, the block then counts
// itself in; the block that completes the count knows every record of this launch has been made visible
// before it writes the word the owner polls.  (A publish from a separate, later kernel is NOT enough over
// NVLink: stream order makes the records visible to this device's next kernel, not yet to the peer.)
__device__ __forceinline__ bool last_block_after_fence(unsigned int* counter)
{
  __shared__ bool s_last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int done = atomicAdd(counter, 1u) + 1u;
    s_last = done == gridDim.x;
    if (s_last) *counter = 0;  // ready for the next launch (stream-ordered)
  }
  __syncthreads();
  if (s_last) __threadfence_system();
  return s_last;
}

__device__ __forceinline__ void publish_counts(ExchangeDev* x, const InboxLayout& L, uint64_t epoch, int finish)
{
  const uint32_t d = threadIdx.x;
  if (d < L.n_ranks) {
    InboxHeader* h = reinterpret_cast<InboxHeader*>(inbox_half(x->inbox[d], L, epoch));
    const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&x->cursor[d]);
    const unsigned long long cnt = cur < L.pair_cap ? cur : L.pair_cap;
    st_sys(&h->committed[L.rank], (epoch << 40) | cnt);
    if (finish) st_sys(&h->done[L.rank], epoch);
  }
}

// pairs [x->routed, x->route_hi) of the dense list -> their owners' inboxes.  Slots are handed out by
// this rank's own cursors (only this rank writes its regions), one atomic per (warp, owner).
// Runs on its own stream beside the NEXT extract launch, which keeps appending to the list: the range ends
// at route_hi, the length recorded when the previous launch ended, not at the live counter.
__global__ void __launch_bounds__(kCtaThreads) route_kernel(PairView pv, ExchangeDev* __restrict__ x, InboxLayout L, uint64_t epoch, int finish)
{
  const unsigned long long lo = x->routed, hi = x->route_hi;
  const int lane = threadIdx.x & 31;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t n = hi - lo;
  const uint64_t n_round = (n + 31) / 32 * 32;  // whole warps stay in the loop
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    const bool live = i < n;
    ulonglong2 hk = make_ulonglong2(0, 0);
    uint32_t dst = 0xFFFFFFFFu;
    if (live) {
      hk = __ldcg(pv.list + lo + i);
      dst = dest_rank(hk.x, hk.y, L.n_ranks);
    }
    const uint32_t peers = __match_any_sync(0xffffffffu, dst);
    const int leader = __ffs(peers) - 1;
    unsigned long long base = 0;
    if (live && lane == leader) base = atomicAdd(&x->cursor[dst], (unsigned long long)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (live) {
      const unsigned long long slot = base + __popc(peers & ((1u << lane) - 1u));
      if (slot < L.pair_cap) {
        uint64_t* rec = inbox_pair_region(x->inbox[dst], L, epoch, L.rank) + slot * L.rec_words;
        if (L.rec_words == 2) {
          *reinterpret_cast<ulonglong2*>(rec) = hk;
        } else {
          rec[0] = hk.x; rec[1] = hk.y; rec[2] = pair_okey(pv, hk.x, hk.y);
        }
      } else {
        x->route_overflow = 1;
      }
    }
  }
  // publish the new counts (and, with `finish`, that this rank has nothing more to send in this job)
  if (last_block_after_fence(&x->blocks_done)) {
    publish_counts(x, L, epoch, finish);
    if (threadIdx.x == 0) x->routed = hi;
  }
}

// a rank with nothing to route (no pair set yet) still has to say so
__global__ void commit_kernel(ExchangeDev* __restrict__ x, InboxLayout L, uint64_t epoch, int finish)
{
  __threadfence_system();
  publish_counts(x, L, epoch, finish);
}

// one block: wait until every source rank has finished routing this job
__global__ void inbox_wait_kernel(ExchangeDev* __restrict__ x, InboxLayout L, uint64_t epoch, unsigned long long timeout_ns)
{
  const uint32_t s = threadIdx.x;
  if (s < L.n_ranks) {
    const InboxHeader* h = reinterpret_cast<const InboxHeader*>(inbox_half(x->inbox[L.rank], L, epoch));
    const unsigned long long t0 = global_ns();
    while (ld_sys(&h->done[s]) != epoch) {
      __nanosleep(200);
      if (global_ns() - t0 > timeout_ns) { x->timed_out = 1; break; }
    }
  }
}

// one block: fix the range the next import covers.  Refuses (and flags) when the owned set could
// pass its load limit; the host then re-imports everything into a larger table.
__global__ void inbox_snapshot_kernel(ExchangeDev* __restrict__ x, InboxLayout L, uint64_t epoch, uint64_t owned_limit)
{
  __shared__ unsigned long long s_new[kMaxP2PRanks];
  __shared__ int s_ok;
  const uint32_t s = threadIdx.x;
  if (s < L.n_ranks) {
    x->imported[s] = x->snap[s];
    const InboxHeader* h = reinterpret_cast<const InboxHeader*>(inbox_half(x->inbox[L.rank], L, epoch));
    const unsigned long long w = ld_sys(&h->committed[s]);
    s_new[s] = (w >> 40) == epoch ? (w & kCountMask) : x->snap[s];
  }
  __syncthreads();
  if (s == 0) {
    unsigned long long pending = 0;
    for (uint32_t r = 0; r < L.n_ranks; ++r) pending += s_new[r] - x->snap[r];
    s_ok = x->n_owned + pending <= owned_limit && !x->owned_overflow;
    if (!s_ok) x->owned_overflow = 1;
  }
  __syncthreads();
  if (s < L.n_ranks && s_ok) x->snap[s] = s_new[s];
}

// records [imported[s], snap[s]) of every source region of MY inbox -> owned pair set (+ dense list)
__global__ void __launch_bounds__(kCtaThreads) inbox_import_kernel(ExchangeDev* __restrict__ x, InboxLayout L, uint64_t epoch,
                                                                   PairView owned, int multi_d)
{
  const int lane = threadIdx.x & 31;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint32_t s = 0; s < L.n_ranks; ++s) {
    const unsigned long long lo = x->imported[s], hi = x->snap[s];
    if (hi <= lo) continue;
    const uint64_t* region = inbox_pair_region(x->inbox[L.rank], L, epoch, s);
    const uint64_t n = hi - lo;
    const uint64_t n_round = (n + 31) / 32 * 32;
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_round; i += stride) {
      bool fresh = false;
      uint64_t a = 0, b = 0;
      if (i < n) {
        const uint64_t* rec = region + (lo + i) * L.rec_words;
        uint64_t okey = 0;
        if (L.rec_words == 2) {
          const ulonglong2 v = __ldcg(reinterpret_cast<const ulonglong2*>(rec));
          a = v.x; b = v.y;
        } else {
          a = __ldcg(rec); b = __ldcg(rec + 1); okey = __ldcg(rec + 2);
        }
        fresh = pair_insert(owned, a, b, okey, multi_d != 0);
      }
      const uint32_t m = __ballot_sync(0xffffffffu, fresh);
      if (m) {
        const int leader = __ffs(m) - 1;
        unsigned long long base = 0;
        if (lane == leader) base = atomicAdd(&x->n_owned, (unsigned long long)__popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (fresh) owned.list[base + __popc(m & ((1u << lane) - 1u))] = make_ulonglong2(a, b);
      }
    }
  }
}

// partial link rows (after this rank's reduce-by-key) -> rank 0's row region for this rank
__global__ void rows_send_kernel(const uint64_t* __restrict__ keys, const unsigned long long* __restrict__ links,
                                 const unsigned long long* __restrict__ gaps, uint32_t n, ExchangeDev* __restrict__ x, InboxLayout L,
                                 uint64_t epoch)
{
  RowRec* dst = inbox_row_region(x->inbox[0], L, epoch, L.rank);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    RowRec r;
    r.key = keys[i]; r.links = links[i]; r.gaps = gaps[i];
    dst[i] = r;
  }
  if (last_block_after_fence(&x->row_blocks_done) && threadIdx.x == 0) {
    InboxHeader* h = reinterpret_cast<InboxHeader*>(inbox_half(x->inbox[0], L, epoch));
    st_sys(&h->rows[L.rank], (epoch << 40) | n);
  }
}
// rank 0, one block: wait for every rank's rows; counts[s] = rows of rank s
__global__ void rows_wait_kernel(ExchangeDev* __restrict__ x, InboxLayout L, uint64_t epoch, unsigned long long* __restrict__ counts,
                                 unsigned long long timeout_ns)
{
  const uint32_t s = threadIdx.x;
  if (s < L.n_ranks) {
    const InboxHeader* h = reinterpret_cast<const InboxHeader*>(inbox_half(x->inbox[0], L, epoch));
    const unsigned long long t0 = global_ns();
    unsigned long long w;
    while (((w = ld_sys(&h->rows[s])) >> 40) != epoch) {
      __nanosleep(200);
      if (global_ns() - t0 > timeout_ns) { x->timed_out = 1; w = 0; break; }
    }
    counts[s] = w & kCountMask;
  }
}
// rank 0: concatenate the row regions; src_start[s] = first output index of rank s's rows
__global__ void rows_gather_kernel(ExchangeDev* __restrict__ x, InboxLayout L, uint64_t epoch, const unsigned long long* __restrict__ src_start,
                                   uint32_t n_total, uint64_t* __restrict__ keys, int32_t* __restrict__ idx,
                                   unsigned long long* __restrict__ links, unsigned long long* __restrict__ gaps)
{
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_total; i += gridDim.x * blockDim.x) {
    uint32_t s = 0;
    while (s + 1 < L.n_ranks && src_start[s + 1] <= i) ++s;
    const RowRec* src = inbox_row_region(x->inbox[0], L, epoch, s);
    const RowRec* rp = src + (i - src_start[s]);
    keys[i] = __ldcg(&rp->key); links[i] = __ldcg(&rp->links); gaps[i] = __ldcg(&rp->gaps);
    idx[i] = int32_t(i);
  }
}

// ------------------------------------------------------------------ Bloom filter built in shares, OR-ed over peer memory
//
// A filter for a Gbp-scale assembly is GBs of DRAM-resident bits and its build is bound by random read-modify-writes
// (hash_num per contig k-mer), the same on every rank if every rank builds its own copy.  With links_bf_alloc_shared each
// rank inserts 1/N of the k-mers into its own full-size copy; this kernel then makes all copies equal to the OR of all:
// rank r owns the r-th slice of the words, reads that slice from every peer (NVLink loads), ORs, and stores the result into
// every copy (NVLink stores) -- a reduce-scatter and an all-gather in one pass, 2 x (N-1)/N x filter bytes over the links.
// The barriers around it (everybody has inserted / everybody has stored) are the caller's: one per job, host side.
struct BloomPeers { uint4* p[kMaxP2PRanks]; };

__global__ void __launch_bounds__(kCtaThreads) bloom_or_reduce_kernel(BloomPeers bp, uint32_t n_ranks, uint64_t lo, uint64_t hi, uint64_t tail8,
                                                                     uint64_t n16)
{
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = lo + uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < hi; i += stride) {
    uint4 v = make_uint4(0, 0, 0, 0);
    for (uint32_t r = 0; r < n_ranks; ++r) {
      const uint4 w = __ldcg(bp.p[r] + i);
      v.x |= w.x; v.y |= w.y; v.z |= w.z; v.w |= w.w;
    }
    for (uint32_t r = 0; r < n_ranks; ++r) __stcg(bp.p[r] + i, v);
  }
  if (tail8 && blockIdx.x == 0 && threadIdx.x == 0) {  // a filter of 8 (mod 16) bytes: the last rank takes the odd half word pair
    unsigned long long v = 0;
    for (uint32_t r = 0; r < n_ranks; ++r) v |= __ldcg(reinterpret_cast<const unsigned long long*>(bp.p[r] + n16));
    for (uint32_t r = 0; r < n_ranks; ++r) __stcg(reinterpret_cast<unsigned long long*>(bp.p[r] + n16), v);
  }
  __threadfence_system();
}

}  // namespace links

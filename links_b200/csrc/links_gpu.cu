// links_b200/csrc/links_gpu.cu -- C-ABI implementation (include/links_gpu.h) over the sm_100a kernels.
//
// Host-side flow per context (one CUDA device, two streams):
//   contigs_push : H2D ascii -> pack -> [bf_build] ; packed contigs stay resident for the table build
//   reads_push   : chunked; per chunk  H2D (copy stream) -> pack -> rank_build -> extract (compute stream)
//                  with two chunk slots so chunk i+1's copy overlaps chunk i's kernels
//   finalize     : table_build over resident contigs -> classify over the pair set -> radix sort
//                  -> reduce-by-key -> D2H rows -> mirror expansion + final ordering on the host
// There is no CPU implementation of any of these stages in this library.
#include "../../include/links_gpu.h"
#include "exchange.cuh"
#include "extract.cuh"
#include "kernels.cuh"

#include <nvtx3/nvToolsExt.h>  // header-only; ranges cost nothing unless a profiler (nsys, ncu --nvtx) is attached

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <ctime>
#include <string>
#include <vector>

using namespace links;

namespace {

thread_local std::string g_create_error;

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes)
  {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { cudaGetLastError(); want = bytes; e = cudaMalloc(&p, want); }
    if (e == cudaSuccess) cap = want; else p = nullptr;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct SeqBatchDev { // device-resident packed sequences of one push/chunk
  DevBuf packed, inv, off, segs, flags, vs, vrank;
  DevBuf mm_bytes, mm_base, mm_segs;  // member map of a read chunk (member_map.cuh)
  uint64_t n_bases = 0, base0 = 0;
  uint32_t n_seq = 0, n_tiles = 0, first_id = 0;
  uint64_t first_read_idx = 0;
  void release() { packed.release(); inv.release(); off.release(); segs.release(); flags.release(); vs.release(); vrank.release(); mm_bytes.release(); mm_base.release(); mm_segs.release(); }
};

enum Family { F_H2D, F_PACK, F_BF, F_TABLE, F_EXTRACT, F_CLASSIFY, F_SORT, F_D2H, F_GROW, F_EXCHANGE, F_COUNT };

struct TimedSpan { int fam; cudaEvent_t a, b; };

} // namespace

struct links_ctx {
  int device = 0;
  int sm_count = 148;
  size_t dev_total_mem = 0;
  links_params prm{};
  std::vector<uint32_t> dist;
  std::vector<uint32_t> step;
  std::vector<int32_t> min_allowed;
  std::string err;
  cudaStream_t s_compute = nullptr, s_copy = nullptr, s_aux = nullptr, s_pack = nullptr;
  bool own_compute_stream = true;

  // ntHash table + distance params on device
  DevBuf d_tab, d_tab2, d_dist, d_step, d_minallowed, d_cursor;
  uint64_t kseed = 0;

  // Bloom filter
  DevBuf d_bloom;
  // links_bf_alloc_shared: the filter lives in a caller-provided, peer-addressable buffer; this rank inserts only its share
  // of the contig k-mers and links_bf_allreduce ORs the ranks' copies through peer memory
  void* bloom_ext = nullptr;
  void* bloom_peers[kMaxP2PRanks] = {nullptr};
  uint32_t bf_share_rank = 0, bf_share_n = 1;
  bool bf_reduced = true;  // false between links_bf_alloc_shared and links_bf_allreduce
  uint64_t bf_bytes = 0;
  uint32_t hash_num = 0;
  bool bf_ready = false, bf_insert = false;
  bool bf_covers_contigs = false;  // links_bf_covers_contigs(): a loaded filter is vouched to hold every pushed contig k-mer

  // member map (member_map.cuh): which x = position - offset any (distance, role) can use, periodic with map_T = lcm(steps)
  DevBuf d_need_idx, d_need_res;
  std::vector<int16_t> need_idx;
  uint32_t map_T = 0, map_c = 0;   // map_T == 0: no map for these options
  int map_mode = -1;               // LINKS_MEMBER_MAP=0/1 forces it off/on; -1: per chunk, where it saves filter walks

  // exact member set of the filter's keys (kernels.cuh MemberView); empty when disabled or not provable
  DevBuf d_members;
  uint64_t member_buckets = 0;   // buckets of the current job's set (0: none)
  uint64_t member_positions = 0; // contig positions inserted so far
  bool member_off = false;       // this job runs without the set (memory budget, LINKS_MEMBER_SET=0)

  // contigs (resident)
  std::vector<SeqBatchDev*> contig_batches;
  std::vector<SeqBatchDev*> contig_pool; // released by links_reset, buffers kept
  cudaEvent_t ev_contig = nullptr, ev_contig_packed = nullptr, ev_cleared = nullptr;
  bool contig_pack_pending = false;
  std::vector<uint32_t> tig_len; // [0] = 0
  uint64_t contig_positions = 0;
  uint32_t n_tigs_seen = 0;

  // read chunk slots
  static constexpr int kSlots = 5;
  SeqBatchDev slot[kSlots];
  DevBuf slot_ascii[kSlots];
  cudaEvent_t ev_copied[kSlots]{}, ev_done[kSlots]{}, ev_packed[kSlots]{};
  bool slot_used[kSlots] = {false, false};
  int next_slot = 0;
  DevBuf contig_ascii;

  // pair set
  DevBuf d_pairs, d_pair_list;  // slots of 16 bytes, or 32 with the order key inside when n_dist > 1 (pair_slot_bytes)
  uint64_t pair_cap = 0;
  int extract_ctas_per_sm = 4;
  uint64_t cand_total = 0;   // candidates enqueued so far (upper bound on inserts)
  // asynchronous occupancy snapshots: {n_unique read back after a chunk, cand_total at that point}
  static constexpr int kSnaps = 8;
  Counters* h_snap = nullptr; // pinned, kSnaps entries
  cudaEvent_t ev_snap[kSnaps]{};
  uint64_t snap_cand[kSnaps] = {0};
  bool snap_live[kSnaps] = {false};
  int snap_next = 0;
  uint64_t known_unique = 0, known_cand = 0; // latest completed snapshot

  // counters
  DevBuf d_ctr;

  // contig table (built on s_aux while the reads stream in; finalize waits on ev_table)
  DevBuf d_bf_stage, d_bf_cursor;  // Bloom build by partition + apply (kernels.cuh BloomBuckets): update entries, cursors + overflow flag
  DevBuf d_table, d_table_occ;  // occ: the MATES build's sector-occupancy bitmap (kernels.cuh TableView::occ)
  bool table_has_occ = false;
  uint64_t table_cap = 0;
  cudaEvent_t ev_table = nullptr, ev_fork = nullptr;
  size_t table_batches_built = 0;
  bool table_started = false;
  // LINKS_TABLE_* (links_gpu.h): FULL = every contig k-mer, built beside the extraction; MATES = only the hashes the
  // (owned) pairs refer to, built at finalize (reference src/LINKS.hpp:580,591); AUTO picks by the full table's size
  int table_mode = LINKS_TABLE_AUTO;
  bool table_mates = false;  // this job's choice

  // stage-4 scratch (kept across links_reset so repeated jobs do not re-allocate)
  DevBuf f_final;
  DevBuf f_keys[2], f_vals[2], f_issues, f_tiglen, f_hist, f_scan_tmp, f_flags, f_total, f_okeys, f_olinks, f_ogaps, f_exch, f_il, f_ig;
  uint64_t issue_cap_hint = 1ull << 16;

  // peer-memory exchange (exchange.cuh); attached once, survives links_reset
  bool x_attached = false, x_finished = false, x_settled = false;
  bool x_fused = true;          // pairs leave for their owner from inside extract_kernel (no local pair set); false: local de-duplication first, route_kernel after every launch (LINKS_EXCHANGE_DEDUP=local)
  bool x_job_started = false;   // something of this job may already sit in the owned set
  uint64_t x_cand_hint = 0;     // candidates of the largest reads_push seen: sizes the owned set before a job's first import
  InboxLayout x_lay{};
  uint64_t x_epoch = 0;
  DevBuf d_xdev;                                  // ExchangeDev
  DevBuf d_owned, d_owned_list;                   // the pairs this rank owns (imports land here)
  uint64_t owned_cap = 0;
  cudaStream_t s_import = nullptr, s_route = nullptr;
  cudaEvent_t ev_routed = nullptr, ev_imported = nullptr, ev_extracted = nullptr;
  DevBuf f_rowcnt;
  FinalRow* h_rows = nullptr;  // pinned staging for the rows' D2H
  size_t h_rows_cap = 0;

  // stage 4 runs along with the extraction (single GPU): after every extract launch the pairs that launch added to
  // the dense list are classified on s_cls, so only the last launch's share is left for links_finalize
  bool cls_on = true, cls_valid = false, cls_started = false;
  cudaStream_t s_cls = nullptr;
  cudaEvent_t ev_cls_in = nullptr, ev_cls_done = nullptr;
  DevBuf d_cls;               // bounds[r] = length of the pair list after round r-1 (bounds[0] = 0): round r classifies [bounds[r], bounds[r+1])
  uint32_t cls_round = 0;     // rounds enqueued so far in this job; every round has its own bound, so no stream waits for another
  static constexpr uint32_t kClsRounds = 8192;
  uint64_t cls_row_cap = 0, cls_pair_cap = 0, cls_issue_cap = 0, max_read_end = 0;

  // timing
  std::vector<TimedSpan> spans;
  std::vector<cudaEvent_t> ev_pool;
  float fam_ms[F_COUNT] = {0};
  uint64_t launches = 0, extract_launches = 0;
  bool timing = true;
  bool finalized = false;
};

namespace {

// Result arrays handed to the caller (released with links_free_result).  A job returns the same megabyte-sized
// arrays over and over; going back to the C library each time means fresh pages and ~1 us of page fault per 4 KB
// (0.4 ms per job measured).  The last few released blocks are kept and reused when large enough.
struct ResultHeader { uint64_t magic, cap; uint64_t pad_[2]; };
constexpr uint64_t kResultMagic = 0x4c494e4b53524f57ull;
std::mutex g_result_mu;
ResultHeader* g_result_cache[4] = {nullptr, nullptr, nullptr, nullptr};

void* result_alloc(size_t bytes)
{
  {
    std::lock_guard<std::mutex> lk(g_result_mu);
    int best = -1;
    for (int i = 0; i < 4; ++i)
      if (g_result_cache[i] && g_result_cache[i]->cap >= bytes && (best < 0 || g_result_cache[i]->cap < g_result_cache[best]->cap)) best = i;
    if (best >= 0 && g_result_cache[best]->cap <= 4 * bytes + (1u << 16)) {
      ResultHeader* h = g_result_cache[best];
      g_result_cache[best] = nullptr;
      return h + 1;
    }
  }
  ResultHeader* h = static_cast<ResultHeader*>(malloc(sizeof(ResultHeader) + bytes));
  if (!h) return nullptr;
  h->magic = kResultMagic; h->cap = bytes;
  return h + 1;
}

void result_free(void* p)
{
  if (!p) return;
  ResultHeader* h = static_cast<ResultHeader*>(p) - 1;
  if (h->magic != kResultMagic) return;  // not ours: leave it alone rather than corrupt the heap
  {
    std::lock_guard<std::mutex> lk(g_result_mu);
    int slot = -1;
    for (int i = 0; i < 4; ++i) if (!g_result_cache[i]) { slot = i; break; }
    if (slot < 0) {  // cache full: replace the smallest block if this one is larger
      int sm = 0;
      for (int i = 1; i < 4; ++i) if (g_result_cache[i]->cap < g_result_cache[sm]->cap) sm = i;
      if (g_result_cache[sm]->cap < h->cap) { ResultHeader* old = g_result_cache[sm]; g_result_cache[sm] = h; h = old; }
    } else { g_result_cache[slot] = h; h = nullptr; }
  }
  if (h) { h->magic = 0; free(h); }
}

// LINKS_DEBUG_HOST=1: host wall-clock marks (us since the first mark) at the milestones of a job
void host_mark(links_ctx* c, const char* what)
{
  static const bool on = getenv("LINKS_DEBUG_HOST") != nullptr;
  if (!on) return;
  static timespec t0{};
  timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  if (!t0.tv_sec) t0 = t;
  fprintf(stderr, "[links host dev%d] %10.1f us  %s\n", c ? c->device : -1, (t.tv_sec - t0.tv_sec) * 1e6 + (t.tv_nsec - t0.tv_nsec) * 1e-3, what);
}

int fail(links_ctx* c, int code, const char* fmt, ...)
{
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (c) c->err = buf; else g_create_error = buf;
  return code;
}

#define CU(c, call)                                                                                   \
  do {                                                                                                \
    cudaError_t e__ = (call);                                                                         \
    if (e__ != cudaSuccess) {                                                                         \
      cudaGetLastError();                                                                             \
      return fail((c), e__ == cudaErrorMemoryAllocation ? LINKS_E_NOMEM : LINKS_E_CUDA, "%s: %s (%s:%d)", #call, \
                  cudaGetErrorString(e__), __FILE__, __LINE__);                                       \
    }                                                                                                 \
  } while (0)

cudaEvent_t get_event(links_ctx* c)
{
  if (!c->ev_pool.empty()) { cudaEvent_t e = c->ev_pool.back(); c->ev_pool.pop_back(); return e; }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}

// NVTX range over a C-ABI call or a phase of one (host side: what was enqueued where; the kernels show under it in nsys)
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

const char* const kFamilyNames[F_COUNT] = {"links:h2d", "links:pack", "links:bloom_build", "links:table_build", "links:extract",
                                           "links:classify", "links:sort_reduce", "links:d2h", "links:grow", "links:exchange"};

struct Span {
  links_ctx* c; int fam; cudaStream_t s; cudaEvent_t a = nullptr;
  Span(links_ctx* c_, int fam_, cudaStream_t s_) : c(c_), fam(fam_), s(s_)
  {
    nvtxRangePushA(kFamilyNames[fam]);
    if (c->timing) { a = get_event(c); cudaEventRecord(a, s); }
  }
  ~Span()
  {
    if (c->timing) { cudaEvent_t b = get_event(c); cudaEventRecord(b, s); c->spans.push_back({fam, a, b}); }
    nvtxRangePop();
  }
};

void resolve_spans(links_ctx* c)
{
  static const bool dbg = getenv("LINKS_DEBUG_TIMELINE") != nullptr;  // device timeline of every span since the last resolve
  if (dbg && !c->spans.empty()) {
    static const char* names[] = {"h2d", "pack", "bf", "table", "extract", "classify", "sort", "d2h", "grow", "exchange"};
    cudaEvent_t t0 = c->spans.front().a;
    for (auto& sp : c->spans) {
      float a = 0, b = 0;
      cudaEventSynchronize(sp.b);
      cudaEventElapsedTime(&a, t0, sp.a);
      cudaEventElapsedTime(&b, t0, sp.b);
      fprintf(stderr, "[links timeline dev%d] %-8s %8.3f -> %8.3f ms\n", c->device, names[sp.fam], a, b);
    }
    fprintf(stderr, "[links timeline dev%d] ----\n", c->device);
  }
  for (auto& sp : c->spans) {
    float ms = 0;
    if (cudaEventSynchronize(sp.b) == cudaSuccess && cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) c->fam_ms[sp.fam] += ms;
    else cudaGetLastError();
    c->ev_pool.push_back(sp.a);
    c->ev_pool.push_back(sp.b);
  }
  c->spans.clear();
}

int grid_for(links_ctx* c, uint64_t work_items, int per_cta, int ctas_per_sm)
{
  uint64_t need = (work_items + per_cta - 1) / per_cta;
  uint64_t cap = uint64_t(c->sm_count) * ctas_per_sm;
  if (need < 1) need = 1;
  return int(std::min(need, cap));
}

int log2_ceil(uint64_t v)
{
  int b = 0;
  while ((1ull << b) < v) ++b;
  return b;
}

TableView table_view(links_ctx* c)
{
  TableView tv;
  tv.slots = c->d_table.as<TableSlot>();
  tv.mask = c->table_cap ? c->table_cap - 1 : 0;
  tv.shift = 64 - log2_ceil(c->table_cap ? c->table_cap : 2);
  tv.occ = c->table_has_occ ? c->d_table_occ.as<uint32_t>() : nullptr;
  return tv;
}
PairView pair_view(PairSlot* p, uint64_t cap, ulonglong2* list, bool wide)
{
  PairView pv;
  pv.slots = p; pv.mask = cap ? cap - 1 : 0; pv.shift = 64 - log2_ceil(cap ? cap : 2); pv.list = list; pv.wide = wide ? 1u : 0u;
  return pv;
}
size_t pair_slot_bytes(const links_ctx* c) { return c->prm.n_dist > 1 ? 32 : 16; }
PairView pair_view(links_ctx* c)
{
  return pair_view(c->d_pairs.as<PairSlot>(), c->pair_cap, c->d_pair_list.as<ulonglong2>(), c->prm.n_dist > 1);
}
uint64_t pair_list_capacity(uint64_t cap) { return cap / 10 * 7 + 1024; }
MemberView member_view(links_ctx* c)
{
  MemberView mv;
  mv.buckets = c->member_buckets ? c->d_members.as<MemberBucket>() : nullptr;
  mv.n_buckets = c->member_buckets;
  return mv;
}
BloomView bloom_view(links_ctx* c)
{
  BloomView bf;
  bf.words = static_cast<uint32_t*>(c->bloom_ext ? c->bloom_ext : c->d_bloom.p);
  bf.array_bits = c->bf_bytes * 8;
  bf.magic = bf.array_bits ? uint64_t((((unsigned __int128)1) << 64) / bf.array_bits) : 0;
  bf.hash_num = c->hash_num;
  bf.bits32 = bf.array_bits < (1ull << 31) ? uint32_t(bf.array_bits) : 0u;
  return bf;
}

// extract_kernel is instantiated for the Bloom hash counts LINKS produces in practice
// (hash_num = floor(-log2(fpr)): 9 for the default -p 0.001) and once with a runtime count.
using ExtractFn = void (*)(ReadBatch, DistParams, int, uint64_t, const ulonglong2*, BloomView, MemberView, PairView, Counters*,
                           ExchangeDev*, InboxLayout, uint64_t, MemberMap);
using MapFn = void (*)(MapBatch, MemberMap, int, uint64_t, uint64_t, const ulonglong2*, BloomView, MemberView, Counters*);
template <bool SMALL>
MapFn map_fn_sel(uint32_t hash_num)
{
  switch (hash_num) {
  case 6: return member_map_kernel<6, SMALL>;
  case 7: return member_map_kernel<7, SMALL>;
  case 8: return member_map_kernel<8, SMALL>;
  case 9: return member_map_kernel<9, SMALL>;
  case 10: return member_map_kernel<10, SMALL>;
  case 13: return member_map_kernel<13, SMALL>;
  default: return member_map_kernel<0, SMALL>;
  }
}
MapFn map_fn_for(uint32_t hash_num, bool small_filter) { return small_filter ? map_fn_sel<true>(hash_num) : map_fn_sel<false>(hash_num); }
template <bool ROUTED, bool SMALL, int MODE>
ExtractFn extract_fn_sel(uint32_t hash_num)
{
  switch (hash_num) {
  case 6: return extract_kernel<6, ROUTED, SMALL, MODE>;
  case 7: return extract_kernel<7, ROUTED, SMALL, MODE>;
  case 8: return extract_kernel<8, ROUTED, SMALL, MODE>;
  case 9: return extract_kernel<9, ROUTED, SMALL, MODE>;
  case 10: return extract_kernel<10, ROUTED, SMALL, MODE>;
  case 13: return extract_kernel<13, ROUTED, SMALL, MODE>;
  default: return extract_kernel<0, ROUTED, SMALL, MODE>;
  }
}
// small_filter: fewer than 2^31 bits (BloomView::bits32 set), the 32-bit remainder variant; mode: 0 grouped probes, 1 a member
// map comes with the launch, 2 one probe per stage visit (extract.cuh)
ExtractFn extract_fn_for(uint32_t hash_num, bool routed, bool small_filter, int mode)
{
  const int v = (routed ? 1 : 0) | (small_filter ? 2 : 0);
  switch (mode * 4 + v) {
  case 0: return extract_fn_sel<false, false, 0>(hash_num);
  case 1: return extract_fn_sel<true, false, 0>(hash_num);
  case 2: return extract_fn_sel<false, true, 0>(hash_num);
  case 3: return extract_fn_sel<true, true, 0>(hash_num);
  case 4: return extract_fn_sel<false, false, 1>(hash_num);
  case 5: return extract_fn_sel<true, false, 1>(hash_num);
  case 6: return extract_fn_sel<false, true, 1>(hash_num);
  case 7: return extract_fn_sel<true, true, 1>(hash_num);
  case 8: return extract_fn_sel<false, false, 2>(hash_num);
  case 9: return extract_fn_sel<true, false, 2>(hash_num);
  case 10: return extract_fn_sel<false, true, 2>(hash_num);
  default: return extract_fn_sel<true, true, 2>(hash_num);
  }
}
const uint32_t kExtractVariants[] = {0, 6, 7, 8, 9, 10, 13};

// exclusive scan of n u32 (in place allowed); total written to *d_total (device)
int device_scan(links_ctx* c, cudaStream_t s, uint32_t* d_in, uint64_t n, uint32_t* d_out, DevBuf& tmp, uint32_t* d_total)
{
  if (n <= uint64_t(kScanTile)) {
    scan_small_kernel<<<1, kScanThreads, 0, s>>>(d_in, uint32_t(n), d_out, d_total);
    c->launches++;
    return LINKS_OK;
  }
  // level sizes
  std::vector<uint64_t> lv;
  for (uint64_t m = n; m > uint64_t(kScanTile);) { m = (m + kScanTile - 1) / kScanTile; lv.push_back(m); }
  uint64_t tot = 0;
  for (auto m : lv) tot += m + 8;
  CU(c, tmp.ensure(tot * sizeof(uint32_t)));
  std::vector<uint32_t*> ptr;
  { uint32_t* p = tmp.as<uint32_t>(); for (auto m : lv) { ptr.push_back(p); p += m + 8; } }
  // up-sweep
  const uint32_t* src = d_in; uint64_t m = n;
  for (size_t l = 0; l < lv.size(); ++l) {
    scan_sums_kernel<<<unsigned(lv[l]), kScanThreads, 0, s>>>(src, m, ptr[l]);
    c->launches++;
    src = ptr[l]; m = lv[l];
  }
  scan_small_kernel<<<1, kScanThreads, 0, s>>>(ptr.back(), uint32_t(lv.back()), ptr.back(), d_total);
  c->launches++;
  // down-sweep
  for (size_t l = lv.size(); l-- > 0;) {
    uint32_t* data = l == 0 ? d_in : ptr[l - 1];
    uint32_t* out = l == 0 ? d_out : ptr[l - 1];
    const uint64_t cnt = l == 0 ? n : lv[l - 1];
    scan_apply_kernel<<<unsigned(lv[l]), kScanThreads, 0, s>>>(data, cnt, ptr[l], out);
    c->launches++;
  }
  return LINKS_OK;
}

// LSD radix sort of (u64 key, i32 payload) over the low `key_bits` bits; buffers ping-pong, `cur`
// names the one holding the result.  keys[1]/vals[1]/f_hist are grown as needed.
int radix_sort_pairs(links_ctx* c, cudaStream_t s, DevBuf (&keys)[2], DevBuf (&vals)[2], uint64_t n, int key_bits, int& cur)
{
  if (n == 0) return LINKS_OK;
  if (n >= (1ull << 32)) return fail(c, LINKS_E_LIMIT, "radix sort: more than 2^32 rows");
  const uint32_t n_blocks = uint32_t((n + kSortTile - 1) / kSortTile);
  CU(c, keys[cur ^ 1].ensure(n * 8)); CU(c, vals[cur ^ 1].ensure(n * 4));
  CU(c, c->f_hist.ensure(uint64_t(kRadix) * n_blocks * 4));
  for (int shift = 0; shift < key_bits; shift += 8) {
    radix_hist_kernel<<<n_blocks, kSortThreads, 0, s>>>(keys[cur].as<uint64_t>(), n, shift, c->f_hist.as<uint32_t>(), n_blocks);
    c->launches++;
    int rc = device_scan(c, s, c->f_hist.as<uint32_t>(), uint64_t(kRadix) * n_blocks, c->f_hist.as<uint32_t>(), c->f_scan_tmp,
                         c->f_total.as<uint32_t>());
    if (rc) return rc;
    radix_scatter_kernel<<<n_blocks, kSortThreads, 0, s>>>(keys[cur].as<uint64_t>(), vals[cur].as<int32_t>(), n, shift,
                                                          c->f_hist.as<uint32_t>(), n_blocks, keys[cur ^ 1].as<uint64_t>(),
                                                          vals[cur ^ 1].as<int32_t>());
    c->launches++;
    cur ^= 1;
  }
  CU(c, cudaGetLastError());
  return LINKS_OK;
}

// upload (unless the ASCII is already device-resident) + pack one batch of sequences into `b`.
// seg_tile_start is computed by the caller.  `dev_seq` != nullptr: device pointer to the same
// concatenated ASCII the host pointer would address (offsets are still absolute into it).
int upload_and_pack(links_ctx* c, SeqBatchDev& b, DevBuf& ascii, const char* seq, const uint8_t* dev_seq, const uint64_t* off,
                    uint32_t n, const std::vector<uint32_t>& seg_tile_start, bool want_rank, cudaStream_t s_copy,
                    cudaStream_t s_comp, cudaEvent_t ev_copied)
{
  const uint64_t base0 = off[0], n_bases = off[n] - off[0];
  b.n_bases = n_bases; b.base0 = base0; b.n_seq = n;
  b.n_tiles = seg_tile_start.back();
  if (!dev_seq) CU(c, ascii.ensure(((n_bases + 63) & ~63ull) + 64));
  CU(c, b.packed.ensure(((n_bases + 15) / 16 + 8) * 4));
  CU(c, b.inv.ensure(((n_bases + 31) / 32 + 8) * 4));
  CU(c, b.off.ensure((uint64_t(n) + 1) * 8));
  CU(c, b.segs.ensure(seg_tile_start.size() * 4));
  CU(c, b.flags.ensure((uint64_t(n) + 1) * 4));
  if (want_rank) {
    const uint64_t nw = (n_bases >> 6) + 2ull * n + 4;
    CU(c, b.vs.ensure(nw * 8));
    CU(c, b.vrank.ensure(nw * 4));
  }
  {
    Span sp(c, F_H2D, s_copy);
    if (n_bases && !dev_seq) CU(c, cudaMemcpyAsync(ascii.p, seq + base0, n_bases, cudaMemcpyHostToDevice, s_copy));
    CU(c, cudaMemcpyAsync(b.off.p, off, (uint64_t(n) + 1) * 8, cudaMemcpyHostToDevice, s_copy));
    CU(c, cudaMemcpyAsync(b.segs.p, seg_tile_start.data(), seg_tile_start.size() * 4, cudaMemcpyHostToDevice, s_copy));
  }
  CU(c, cudaEventRecord(ev_copied, s_copy));
  CU(c, cudaStreamWaitEvent(s_comp, ev_copied, 0));
  {
    Span sp(c, F_PACK, s_comp);
    CU(c, cudaMemsetAsync(b.flags.p, 0, (uint64_t(n) + 1) * 4, s_comp));
    // padding words the k-mer window loads may touch
    CU(c, cudaMemsetAsync(b.packed.as<uint32_t>() + (n_bases + 15) / 16, 0, 8 * 4, s_comp));
    CU(c, cudaMemsetAsync(b.inv.as<uint32_t>() + (n_bases + 31) / 32, 0xFF, 8 * 4, s_comp));
    if (n_bases) {
      const uint64_t threads = (n_bases + 31) / 32;
      const uint8_t* src = dev_seq ? dev_seq + base0 : ascii.as<uint8_t>();
      pack_kernel<<<unsigned((threads + kCtaThreads - 1) / kCtaThreads), kCtaThreads, 0, s_comp>>>(
        src, n_bases, b.packed.as<uint32_t>(), b.inv.as<uint32_t>(), b.off.as<uint64_t>(), n, base0, b.flags.as<uint32_t>());
      c->launches++;
      if (want_rank) {
        rank_build_kernel<<<n, kCtaThreads, 0, s_comp>>>(b.inv.as<uint32_t>(), b.off.as<uint64_t>(), base0,
                                                         b.flags.as<uint32_t>(), int(c->prm.k),
                                                         b.vs.as<unsigned long long>(), b.vrank.as<uint32_t>());
        c->launches++;
      }
    }
  }
  CU(c, cudaGetLastError());
  return LINKS_OK;
}

// Latest completed occupancy snapshot -> (known_unique, known_cand).
void poll_snapshots(links_ctx* c)
{
  for (int i = 0; i < links_ctx::kSnaps; ++i) {
    if (!c->snap_live[i]) continue;
    if (cudaEventQuery(c->ev_snap[i]) != cudaSuccess) { cudaGetLastError(); continue; }
    c->snap_live[i] = false;
    if (c->snap_cand[i] >= c->known_cand) { c->known_cand = c->snap_cand[i]; c->known_unique = c->h_snap[i].n_unique; }
  }
}

int enqueue_snapshot(links_ctx* c)
{
  int i = c->snap_next;
  c->snap_next = (i + 1) % links_ctx::kSnaps;
  if (c->snap_live[i]) { // ring wrapped: wait for the old one
    CU(c, cudaEventSynchronize(c->ev_snap[i]));
    poll_snapshots(c);
  }
  CU(c, cudaMemcpyAsync(&c->h_snap[i], c->d_ctr.p, sizeof(Counters), cudaMemcpyDeviceToHost, c->s_compute));
  CU(c, cudaEventRecord(c->ev_snap[i], c->s_compute));
  c->snap_cand[i] = c->cand_total;
  c->snap_live[i] = true;
  return LINKS_OK;
}

// Pair-set capacity.  Invariant: occupied + in-flight inserts <= 0.7 * cap, so probing terminates.
// The occupied count is only known with a lag (async snapshots), so the host bound is
//   known_unique + (candidates enqueued after that snapshot) + incoming.
// `later` = candidates of the rest of this push after `incoming` (used to size a growth once).
int ensure_pair_capacity(links_ctx* c, uint64_t incoming, uint64_t later)
{
  auto bound = [&]() { return c->known_unique + (c->cand_total - c->known_cand) + incoming; };
  auto fits = [&](uint64_t cap) { return cap && bound() * 10 <= cap * 7; };
  poll_snapshots(c);
  if (fits(c->pair_cap)) return LINKS_OK;
  // tighten with the exact count
  static const bool dbg = getenv("LINKS_DEBUG") != nullptr;
  if (dbg) fprintf(stderr, "[links] pair-set capacity check blocks: cap=%llu known_unique=%llu in-flight cand=%llu incoming=%llu\n",
                   (unsigned long long)c->pair_cap, (unsigned long long)c->known_unique,
                   (unsigned long long)(c->cand_total - c->known_cand), (unsigned long long)incoming);
  CU(c, cudaStreamSynchronize(c->s_compute));
  poll_snapshots(c);
  Counters h;
  CU(c, cudaMemcpy(&h, c->d_ctr.p, sizeof h, cudaMemcpyDeviceToHost));
  c->known_unique = h.n_unique; c->known_cand = c->cand_total;
  if (fits(c->pair_cap)) return LINKS_OK;
  // grow once to the projected final size: observed unique/candidate rate (x1.25) applied to the rest of the push,
  // plus two chunks of in-flight slack
  double rate = c->cand_total ? std::min(1.0, 1.25 * double(c->known_unique) / double(c->cand_total)) : 0.5;
  const double projected = double(c->known_unique) + rate * double(later) + 2.0 * double(incoming);
  const uint64_t want = uint64_t(std::max(projected, double(bound())) / 0.7) + 1024;
  const uint64_t cap = 1ull << log2_ceil(std::max<uint64_t>(want, 1ull << 16));
  DevBuf nb, nl;
  const bool multi_d = c->prm.n_dist > 1;
  CU(c, nb.ensure(cap * pair_slot_bytes(c)));
  CU(c, nl.ensure(pair_list_capacity(cap) * sizeof(ulonglong2)));
  {
    Span sp(c, F_GROW, c->s_compute);
    CU(c, cudaMemsetAsync(nb.p, 0xFF, cap * pair_slot_bytes(c), c->s_compute));
    if (c->pair_cap && c->known_unique) {
      pair_rehash_kernel<<<grid_for(c, c->pair_cap, 256, 8), 256, 0, c->s_compute>>>(
        pair_view(c), c->pair_cap, pair_view(nb.as<PairSlot>(), cap, nl.as<ulonglong2>(), multi_d), c->d_ctr.as<Counters>(), 0);
      c->launches++;
      CU(c, cudaMemcpyAsync(nl.p, c->d_pair_list.p, c->known_unique * sizeof(ulonglong2), cudaMemcpyDeviceToDevice, c->s_compute));
    }
  }
  CU(c, cudaStreamSynchronize(c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_route));  // a routing kernel may still read the old list
  CU(c, cudaStreamSynchronize(c->s_cls));    // and so may an incremental classify; its row buffers were sized for the old table
  c->cls_valid = false;
  c->d_pairs.release(); c->d_pair_list.release();
  c->d_pairs = nb; c->d_pair_list = nl;
  c->pair_cap = cap;
  return LINKS_OK;
}

// Member set sizing (kernels.cuh MemberView): two slots per contig k-mer position (load <= 0.5).
// Called with the new batch already counted in contig_positions and listed last in contig_batches.
// First sizing uses the filter's own size as the estimate of the assembly (make_bf gives it
// ~1.8 bytes per base, src/LINKS.hpp:297-302) so that a multi-batch push rarely regrows.
int ensure_member_capacity(links_ctx* c)
{
  // LINKS_MEMBER_SET=0 / 1 forces the set off / on; unset = the measured rule below
  static const int env_mode = []() { const char* e = getenv("LINKS_MEMBER_SET"); return !e ? -1 : e[0] == '0' ? 0 : 1; }();
  if (env_mode == 0 || c->member_off) { c->member_buckets = 0; c->member_off = true; return LINKS_OK; }
  const uint64_t need_buckets = c->contig_positions / 2 + 1;  // 4 slots per bucket, load 0.5
  if (c->member_buckets >= need_buckets) return LINKS_OK;
  const uint64_t estimate = c->member_buckets ? need_buckets + need_buckets / 2 : std::max(need_buckets, c->bf_bytes * 9 / 32);
  if (env_mode < 0 && !c->member_buckets) {
    // A set lookup is one sector that replaces hash_num filter sectors for a true member -- a win when those
    // sectors come from DRAM (filter larger than L2: 80 Mbp assembly, extract 131 -> 98 ms), a loss when the filter
    // sits in L2 but the 9x larger set does not (20 Mbp: 14.4 -> 18.0 ms), neutral when both nearly fit (5 Mbp).
    const bool filter_in_l2 = c->bf_bytes < (64ull << 20);
    const bool set_near_l2 = estimate * sizeof(MemberBucket) <= (96ull << 20);
    if (filter_in_l2 && !set_near_l2) { c->member_buckets = 0; c->member_off = true; return LINKS_OK; }
  }
  const size_t total_b = c->dev_total_mem;  // (cudaMemGetInfo is a slow driver call; the total does not change)
  if (estimate * sizeof(MemberBucket) > total_b / 8 && estimate * sizeof(MemberBucket) > c->d_members.cap) {
    c->member_buckets = 0; c->member_off = true;  // large assemblies: every k-mer takes the probe path
    return LINKS_OK;
  }
  CU(c, c->d_members.ensure(estimate * sizeof(MemberBucket)));
  c->member_buckets = estimate;
  CU(c, cudaMemsetAsync(c->d_members.p, 0xFF, estimate * sizeof(MemberBucket), c->s_compute));
  for (size_t i = 0; i + 1 < c->contig_batches.size(); ++i) {  // earlier batches of this job
    SeqBatchDev* b = c->contig_batches[i];
    if (!b->n_tiles) continue;
    ContigBatch cb{b->packed.as<uint32_t>(), b->inv.as<uint32_t>(), b->off.as<uint64_t>(), b->base0,
                   b->segs.as<uint32_t>(), b->n_seq, b->n_tiles, b->first_id};
    contig_kernel<2><<<grid_for(c, b->n_tiles, 1, 8), kCtaThreads, c->prm.k * 4 * sizeof(ulonglong2), c->s_compute>>>(
      cb, int(c->prm.k), c->kseed, c->d_tab.as<ulonglong2>(), bloom_view(c), TableView{nullptr, 0, 0}, member_view(c), c->d_ctr.as<Counters>(),
      BloomPass{0, ~0ull, 1});
    c->launches++;
  }
  CU(c, cudaGetLastError());
  return LINKS_OK;
}

// ---- peer-memory exchange (exchange.cuh)

// how long a rank waits for its peers before it fails the job (LINKS_EXCHANGE_TIMEOUT_MS, default 60 s)
unsigned long long exchange_timeout_ns()
{
  static const unsigned long long ns = []() {
    const char* e = getenv("LINKS_EXCHANGE_TIMEOUT_MS");
    const unsigned long long ms = e ? strtoull(e, nullptr, 10) : 60000ull;
    return (ms ? ms : 1ull) * 1000000ull;
  }();
  return ns;
}

PairView owned_view(links_ctx* c)
{
  return pair_view(c->d_owned.as<PairSlot>(), c->owned_cap, c->d_owned_list.as<ulonglong2>(), c->prm.n_dist > 1);
}

int ensure_owned(links_ctx* c, uint64_t min_cap)
{
  // before a job's first import the table may be replaced by a larger one; later only exchange_settle's
  // re-import (min_cap != 0 there) does that
  uint64_t want = std::max<uint64_t>(std::max(min_cap, c->pair_cap), 1ull << 16);
  if (!c->x_job_started) want = std::max<uint64_t>(want, c->x_cand_hint / 2);  // ~35 % of the candidates survive at most, load 0.7
  const bool may_replace = min_cap != 0 || !c->x_job_started;
  c->x_job_started = true;
  if (c->owned_cap >= want || (c->owned_cap && !may_replace)) return LINKS_OK;
  const uint64_t cap = 1ull << log2_ceil(want);
  CU(c, cudaStreamSynchronize(c->s_route));
  CU(c, cudaStreamSynchronize(c->s_cls));
  CU(c, cudaStreamSynchronize(c->s_import));
  CU(c, cudaStreamSynchronize(c->s_aux));  // links_reset may still be clearing the old table
  c->d_owned.release(); c->d_owned_list.release();
  CU(c, c->d_owned.ensure(cap * pair_slot_bytes(c)));
  CU(c, c->d_owned_list.ensure(pair_list_capacity(cap) * sizeof(ulonglong2)));
  c->owned_cap = cap;
  Span sp(c, F_GROW, c->s_import);
  CU(c, cudaMemsetAsync(c->d_owned.p, 0xFF, cap * pair_slot_bytes(c), c->s_import));
  return LINKS_OK;
}

int classify_increment_owned(links_ctx* c);

// s_import, after ev_routed: insert what has arrived (with `finish`: wait for every rank's "done" first)
int exchange_import_round(links_ctx* c, bool finish)
{
  ExchangeDev* x = c->d_xdev.as<ExchangeDev>();
  CU(c, cudaStreamWaitEvent(c->s_import, c->ev_routed, 0));
  CU(c, cudaStreamWaitEvent(c->s_import, c->ev_cleared, 0));
  int rc = ensure_owned(c, 0);
  if (rc) return rc;
  {
    Span sp(c, F_EXCHANGE, c->s_import);
    if (finish) { inbox_wait_kernel<<<1, kMaxP2PRanks, 0, c->s_import>>>(x, c->x_lay, c->x_epoch, exchange_timeout_ns()); c->launches++; }
    inbox_snapshot_kernel<<<1, kMaxP2PRanks, 0, c->s_import>>>(x, c->x_lay, c->x_epoch, c->owned_cap / 10 * 7);
    inbox_import_kernel<<<c->sm_count * 2, kCtaThreads, 0, c->s_import>>>(x, c->x_lay, c->x_epoch, owned_view(c), c->prm.n_dist > 1 ? 1 : 0);
    c->launches += 2;
  }
  CU(c, cudaGetLastError());
  return finish ? LINKS_OK : classify_increment_owned(c);
}

// s_compute: route the new part of the dense pair list + publish; s_import: insert what has arrived
int exchange_route(links_ctx* c, bool finish)
{
  ExchangeDev* x = c->d_xdev.as<ExchangeDev>();
  // the list length as of now (stream order: after the extract launch just enqueued); the routing itself runs on
  // s_route beside the next extract launch
  CU(c, cudaStreamWaitEvent(c->s_compute, c->ev_routed, 0));  // the previous routing round still reads route_hi
  CU(c, cudaMemcpyAsync(&x->route_hi, &c->d_ctr.as<Counters>()->n_unique, 8, cudaMemcpyDeviceToDevice, c->s_compute));
  CU(c, cudaEventRecord(c->ev_extracted, c->s_compute));
  CU(c, cudaStreamWaitEvent(c->s_route, c->ev_extracted, 0));
  {
    Span sp(c, F_EXCHANGE, c->s_route);
    if (c->pair_cap) {
      route_kernel<<<c->sm_count * 2, kCtaThreads, 0, c->s_route>>>(pair_view(c), x, c->x_lay, c->x_epoch, finish ? 1 : 0);
    } else {
      commit_kernel<<<1, kMaxP2PRanks, 0, c->s_route>>>(x, c->x_lay, c->x_epoch, finish ? 1 : 0);
    }
    c->launches++;
  }
  CU(c, cudaGetLastError());
  CU(c, cudaEventRecord(c->ev_routed, c->s_route));
  return exchange_import_round(c, finish);
}

// fused mode: extract_kernel<.., true> has written its survivors into the owners' inboxes and published the counts
int exchange_after_routed_extract(links_ctx* c, bool finish)
{
  ExchangeDev* x = c->d_xdev.as<ExchangeDev>();
  if (finish) {
    Span sp(c, F_EXCHANGE, c->s_compute);
    commit_kernel<<<1, kMaxP2PRanks, 0, c->s_compute>>>(x, c->x_lay, c->x_epoch, 1);
    c->launches++;
    CU(c, cudaGetLastError());
  }
  CU(c, cudaEventRecord(c->ev_routed, c->s_compute));
  return exchange_import_round(c, finish);
}

// Before stage 4: everything this rank owns is in the owned set (or re-import it into a larger one), then the
// owned set takes the place of the local one so classify and the inspection calls see it.
int exchange_settle(links_ctx* c)
{
  if (c->x_settled) return LINKS_OK;
  NvtxRange nvtx_("links:exchange_settle (wait for peers, last import)");
  if (!c->x_finished) return fail(c, LINKS_E_INVAL, "links_finalize: links_exchange_finish has not been called on this rank");
  ExchangeDev* x = c->d_xdev.as<ExchangeDev>();
  for (int attempt = 0;; ++attempt) {
    CU(c, cudaStreamSynchronize(c->s_compute));
    CU(c, cudaStreamSynchronize(c->s_route));
  CU(c, cudaStreamSynchronize(c->s_cls));
  CU(c, cudaStreamSynchronize(c->s_import));
    ExchangeDev h;
    CU(c, cudaMemcpy(&h, x, sizeof h, cudaMemcpyDeviceToHost));
    if (h.timed_out) return fail(c, LINKS_E_CUDA, "pair exchange: timed out waiting for the other ranks to finish routing (LINKS_EXCHANGE_TIMEOUT_MS)");
    if (h.route_overflow)
      return fail(c, LINKS_E_LIMIT, "pair exchange: a peer inbox region is full (%llu records per source rank); attach a larger inbox%s",
                  (unsigned long long)c->x_lay.pair_cap,
                  c->x_fused ? " or, for reads that repeat the same k-mer pairs many times, set LINKS_EXCHANGE_DEDUP=local (de-duplicate before sending)" : "");
    if (!h.owned_overflow) {
      CU(c, cudaMemcpyAsync(&c->d_ctr.as<Counters>()->n_unique, &x->n_owned, 8, cudaMemcpyDeviceToDevice, c->s_compute));
      break;
    }
    if (attempt) return fail(c, LINKS_E_CUDA, "pair exchange: owned set overflow after re-import");
    // the inbox still holds every record of this job: start over in a table sized for all of them
    c->cls_valid = false;  // rows of the abandoned table are dropped: stage 4 runs once, from scratch, in links_finalize
    InboxHeader hdr;
    CU(c, cudaMemcpy(&hdr, inbox_half(h.inbox[c->x_lay.rank], c->x_lay, c->x_epoch), sizeof hdr, cudaMemcpyDeviceToHost));
    uint64_t total = 0;
    for (uint32_t r = 0; r < c->x_lay.n_ranks; ++r) total += hdr.committed[r] & kCountMask;
    int rc = ensure_owned(c, total / 7 * 10 + 1024);
    if (rc) return rc;
    CU(c, cudaMemsetAsync(reinterpret_cast<char*>(x) + offsetof(ExchangeDev, imported), 0,
                          offsetof(ExchangeDev, routed) - offsetof(ExchangeDev, imported), c->s_import));
    CU(c, cudaMemsetAsync(&x->n_owned, 0, 8, c->s_import));
    CU(c, cudaMemsetAsync(&x->owned_overflow, 0, 8, c->s_import));
    CU(c, cudaMemsetAsync(c->d_owned.p, 0xFF, c->owned_cap * pair_slot_bytes(c), c->s_import));
    Span sp(c, F_EXCHANGE, c->s_import);
    inbox_snapshot_kernel<<<1, kMaxP2PRanks, 0, c->s_import>>>(x, c->x_lay, c->x_epoch, c->owned_cap / 10 * 7);
    inbox_import_kernel<<<c->sm_count * 2, kCtaThreads, 0, c->s_import>>>(x, c->x_lay, c->x_epoch, owned_view(c), c->prm.n_dist > 1 ? 1 : 0);
    c->launches += 2;
    CU(c, cudaGetLastError());
  }
  std::swap(c->d_pairs, c->d_owned); std::swap(c->d_pair_list, c->d_owned_list);
  std::swap(c->pair_cap, c->owned_cap);
  c->x_settled = true;
  return LINKS_OK;
}

// Multi-GPU: every rank's reduce-by-key output {key, links, gaps}[n] goes to rank 0's inbox; rank 0 sums
// rows with equal keys (sort + segmented sum) into f_okeys/f_olinks/f_ogaps.  *n_out: rows now in
// those arrays on rank 0, 0 elsewhere (the job's rows are delivered on rank 0).
int exchange_rows(links_ctx* c, cudaStream_t s, uint32_t n, int key_bits, uint32_t* n_out)
{
  NvtxRange nvtx_("links:exchange_rows (partial link rows to rank 0)");
  ExchangeDev* x = c->d_xdev.as<ExchangeDev>();
  const InboxLayout& L = c->x_lay;
  if (n > L.row_cap) return fail(c, LINKS_E_LIMIT, "row exchange: %u rows, inbox has room for %llu per rank", n, (unsigned long long)L.row_cap);
  {
    Span sp(c, F_EXCHANGE, s);
    rows_send_kernel<<<std::max(1u, std::min<unsigned>((n + 255) / 256, 1024u)), 256, 0, s>>>(
      c->f_okeys.as<uint64_t>(), c->f_olinks.as<unsigned long long>(), c->f_ogaps.as<unsigned long long>(), n, x, L, c->x_epoch);
    c->launches++;
  }
  CU(c, cudaGetLastError());
  *n_out = 0;
  if (L.rank != 0) return LINKS_OK;
  CU(c, c->f_rowcnt.ensure(size_t(kMaxP2PRanks) * 2 * 8));
  unsigned long long* d_cnt = c->f_rowcnt.as<unsigned long long>();
  unsigned long long* d_start = d_cnt + kMaxP2PRanks;
  rows_wait_kernel<<<1, kMaxP2PRanks, 0, s>>>(x, L, c->x_epoch, d_cnt, exchange_timeout_ns());
  c->launches++;
  unsigned long long hc[kMaxP2PRanks], hs[kMaxP2PRanks];
  CU(c, cudaMemcpyAsync(hc, d_cnt, size_t(L.n_ranks) * 8, cudaMemcpyDeviceToHost, s));
  CU(c, cudaStreamSynchronize(s));
  {
    unsigned long long timed_out = 0;
    CU(c, cudaMemcpy(&timed_out, &x->timed_out, 8, cudaMemcpyDeviceToHost));
    if (timed_out) return fail(c, LINKS_E_CUDA, "row exchange: timed out waiting for the other ranks' link rows (LINKS_EXCHANGE_TIMEOUT_MS)");
  }
  unsigned long long total = 0;
  for (uint32_t r = 0; r < L.n_ranks; ++r) { hs[r] = total; total += hc[r]; }
  if (!total) return LINKS_OK;
  if (total >= (1ull << 31)) return fail(c, LINKS_E_LIMIT, "row exchange: more than 2^31 partial rows");
  CU(c, cudaMemcpy(d_start, hs, size_t(L.n_ranks) * 8, cudaMemcpyHostToDevice));
  DevBuf (&d_keys)[2] = c->f_keys; DevBuf (&d_vals)[2] = c->f_vals;
  CU(c, d_keys[0].ensure(total * 8)); CU(c, d_vals[0].ensure(total * 4)); CU(c, c->f_il.ensure(total * 8)); CU(c, c->f_ig.ensure(total * 8));
  CU(c, c->f_flags.ensure(total * 4)); CU(c, c->f_total.ensure(64));
  uint32_t n_seg = 0;
  {
    Span sp(c, F_SORT, s);
    rows_gather_kernel<<<unsigned((total + 255) / 256), 256, 0, s>>>(x, L, c->x_epoch, d_start, uint32_t(total), d_keys[0].as<uint64_t>(),
                                                                    d_vals[0].as<int32_t>(), c->f_il.as<unsigned long long>(),
                                                                    c->f_ig.as<unsigned long long>());
    c->launches++;
    int cur = 0;
    int rc = radix_sort_pairs(c, s, d_keys, d_vals, total, key_bits, cur);
    if (rc) return rc;
    head_flags_kernel<<<unsigned((total + 255) / 256), 256, 0, s>>>(d_keys[cur].as<uint64_t>(), total, c->f_flags.as<uint32_t>());
    c->launches++;
    rc = device_scan(c, s, c->f_flags.as<uint32_t>(), total, c->f_flags.as<uint32_t>(), c->f_scan_tmp, c->f_total.as<uint32_t>());
    if (rc) return rc;
    CU(c, cudaMemcpyAsync(&n_seg, c->f_total.p, 4, cudaMemcpyDeviceToHost, s));
    CU(c, cudaStreamSynchronize(s));
    CU(c, c->f_okeys.ensure(uint64_t(n_seg) * 8)); CU(c, c->f_olinks.ensure(uint64_t(n_seg) * 8)); CU(c, c->f_ogaps.ensure(uint64_t(n_seg) * 8));
    CU(c, cudaMemsetAsync(c->f_olinks.p, 0, uint64_t(n_seg) * 8, s));
    CU(c, cudaMemsetAsync(c->f_ogaps.p, 0, uint64_t(n_seg) * 8, s));
    rows_segment_sum_kernel<<<unsigned((total + 255) / 256), 256, 0, s>>>(
      d_keys[cur].as<uint64_t>(), d_vals[cur].as<int32_t>(), total, c->f_flags.as<uint32_t>(), c->f_il.as<unsigned long long>(),
      c->f_ig.as<unsigned long long>(), c->f_okeys.as<uint64_t>(), c->f_olinks.as<unsigned long long>(), c->f_ogaps.as<unsigned long long>());
    c->launches++;
    CU(c, cudaGetLastError());
  }
  *n_out = n_seg;
  return LINKS_OK;
}

// ---- stage 4 alongside the extraction (single GPU).  Valid while: the contig table is complete and being built or
// built (no contigs after the first reads), the pair table has not been re-allocated, reads arrive in increasing
// global ordinal when there are several distances (so a classified pair's winning distance cannot change later),
// and the issue buffer does not overflow.  Otherwise links_finalize classifies everything from scratch as before.
int classify_increment(links_ctx* c)
{
  if (!c->cls_valid || c->x_attached || !c->pair_cap) return LINKS_OK;
  if (!(c->table_started && c->table_batches_built == c->contig_batches.size())) return LINKS_OK;
  if (!c->cls_started) {
    c->cls_row_cap = pair_list_capacity(c->pair_cap);
    c->cls_pair_cap = c->pair_cap;
    c->cls_issue_cap = c->issue_cap_hint;
    CU(c, c->f_keys[0].ensure(c->cls_row_cap * 8)); CU(c, c->f_vals[0].ensure(c->cls_row_cap * 4));
    CU(c, c->f_issues.ensure(c->cls_issue_cap * sizeof(IssueRec)));
    CU(c, c->f_tiglen.ensure(c->tig_len.size() * 4));
    CU(c, cudaMemcpyAsync(c->f_tiglen.p, c->tig_len.data(), c->tig_len.size() * 4, cudaMemcpyHostToDevice, c->s_cls));
    c->cls_started = true;
  }
  if (c->pair_cap != c->cls_pair_cap) { c->cls_valid = false; return LINKS_OK; }
  if (c->cls_round + 2 >= links_ctx::kClsRounds) return LINKS_OK;  // out of bounds slots: the rest is classified at finalize
  unsigned long long* range = c->d_cls.as<unsigned long long>() + c->cls_round++;
  CU(c, cudaMemcpyAsync(range + 1, &c->d_ctr.as<Counters>()->n_unique, 8, cudaMemcpyDeviceToDevice, c->s_compute));
  CU(c, cudaEventRecord(c->ev_cls_in, c->s_compute));
  CU(c, cudaStreamWaitEvent(c->s_cls, c->ev_cls_in, 0));
  CU(c, cudaStreamWaitEvent(c->s_cls, c->ev_table, 0));
  {
    Span sp(c, F_CLASSIFY, c->s_cls);
    const int tb = log2_ceil(uint64_t(c->n_tigs_seen) * 2 + 2);
    ClassifyParams cp{c->d_dist.as<uint32_t>(), c->d_minallowed.as<int32_t>(), c->prm.n_dist, c->f_tiglen.as<uint32_t>(), tb,
                      c->f_keys[0].as<uint64_t>(), c->f_vals[0].as<int32_t>(), c->cls_row_cap, c->f_issues.as<IssueRec>(), c->cls_issue_cap};
    classify_kernel<<<c->sm_count * 4, kCtaThreads, 0, c->s_cls>>>(pair_view(c), 0, table_view(c), cp, int(c->prm.k),
                                                                   c->d_ctr.as<Counters>(), range);
    c->launches++;
  }
  CU(c, cudaGetLastError());
  CU(c, cudaEventRecord(c->ev_cls_done, c->s_cls));
  return LINKS_OK;
}

// The same beside the imports of an attached exchange: after every import round the pairs it appended to the OWNED list are
// classified on s_cls, so links_finalize is left with the last round's share instead of the whole owned set (the tail every
// rank waits for).  Single distance only: with several, a pair's winning distance may still be lowered by a record of
// another rank that arrives later.  Needs the full contig table (the one restricted to mates is seeded from the final list).
int classify_increment_owned(links_ctx* c)
{
  if (!c->cls_valid || c->prm.n_dist > 1 || !c->owned_cap) return LINKS_OK;
  if (!(c->table_started && c->table_batches_built == c->contig_batches.size())) return LINKS_OK;
  if (!c->cls_started) {
    c->cls_row_cap = pair_list_capacity(c->owned_cap);
    c->cls_pair_cap = c->owned_cap;
    c->cls_issue_cap = c->issue_cap_hint;
    CU(c, c->f_keys[0].ensure(c->cls_row_cap * 8)); CU(c, c->f_vals[0].ensure(c->cls_row_cap * 4));
    CU(c, c->f_issues.ensure(c->cls_issue_cap * sizeof(IssueRec)));
    CU(c, c->f_tiglen.ensure(c->tig_len.size() * 4));
    CU(c, cudaMemcpyAsync(c->f_tiglen.p, c->tig_len.data(), c->tig_len.size() * 4, cudaMemcpyHostToDevice, c->s_cls));
    c->cls_started = true;
  }
  if (c->owned_cap != c->cls_pair_cap) { c->cls_valid = false; return LINKS_OK; }
  if (c->cls_round + 2 >= links_ctx::kClsRounds) return LINKS_OK;
  unsigned long long* range = c->d_cls.as<unsigned long long>() + c->cls_round++;
  ExchangeDev* x = c->d_xdev.as<ExchangeDev>();
  CU(c, cudaMemcpyAsync(range + 1, &x->n_owned, 8, cudaMemcpyDeviceToDevice, c->s_import));
  CU(c, cudaEventRecord(c->ev_imported, c->s_import));
  CU(c, cudaStreamWaitEvent(c->s_cls, c->ev_imported, 0));
  CU(c, cudaStreamWaitEvent(c->s_cls, c->ev_table, 0));
  {
    Span sp(c, F_CLASSIFY, c->s_cls);
    const int tb = log2_ceil(uint64_t(c->n_tigs_seen) * 2 + 2);
    ClassifyParams cp{c->d_dist.as<uint32_t>(), c->d_minallowed.as<int32_t>(), c->prm.n_dist, c->f_tiglen.as<uint32_t>(), tb,
                      c->f_keys[0].as<uint64_t>(), c->f_vals[0].as<int32_t>(), c->cls_row_cap, c->f_issues.as<IssueRec>(), c->cls_issue_cap};
    classify_kernel<<<c->sm_count * 2, kCtaThreads, 0, c->s_cls>>>(owned_view(c), 0, table_view(c), cp, int(c->prm.k),
                                                                   c->d_ctr.as<Counters>(), range);
    c->launches++;
  }
  CU(c, cudaGetLastError());
  CU(c, cudaEventRecord(c->ev_cls_done, c->s_cls));
  return LINKS_OK;
}

} // namespace

// ====================================================================== fixed scratch for links_kmer_hashes_static
//
// No counterpart in the reference: added for callers that may not allocate device
// memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  links_kmer_hashes
// stages the sequence, its 2-bit packing, the invalid-base mask, the record offsets / flags pack_kernel wants and the
// three result arrays in device buffers it allocates and frees on every call; the variant below runs the same two
// kernels over ONE array in the module's data segment.  Every GPU that loads liblinks_gpu.so holds these
// kStaticScratchBytes (~2.4 MB) until the process exits.  In host code `g_static_scratch` is only the symbol's host-side
// stand-in: the device address comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr uint64_t kStaticMaxBases = LINKS_STATIC_MAX_BASES;
constexpr size_t static_al(size_t x) { return (x + 255) & ~size_t(255); }
constexpr size_t kSsAscii = 0;                                                          // len bytes (+ slack for pack_kernel's 16-byte loads)
constexpr size_t kSsPacked = kSsAscii + static_al(kStaticMaxBases + 128);               // 2 bits per base + 8 padding words
constexpr size_t kSsInv = kSsPacked + static_al((kStaticMaxBases / 16 + 8) * 4);        // 1 bit per base + 8 padding words
constexpr size_t kSsOff = kSsInv + static_al((kStaticMaxBases / 32 + 8) * 4);           // record offsets {0, len}
constexpr size_t kSsFlags = kSsOff + 256;                                               // per-record "has an invalid base" (must start at zero)
constexpr size_t kSsFwd = kSsFlags + 256;
constexpr size_t kSsRev = kSsFwd + static_al(kStaticMaxBases * 8);
constexpr size_t kSsValid = kSsRev + static_al(kStaticMaxBases * 8);
constexpr size_t kStaticScratchBytes = kSsValid + static_al(kStaticMaxBases);
static_assert(kStaticScratchBytes <= (4u << 20), "the static scratch is meant to stay small: every GPU holds it until exit");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace {

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;                              // one call at a time in this process: every call shares the array
unsigned char* g_static_ptr[kStaticMaxDevices] = {};  // device address of g_static_scratch on device i (null: not asked yet)

// caller holds g_static_mu and has made c->device current
int static_scratch_base(links_ctx* c, unsigned char** out)
{
  if (c->device < 0 || c->device >= kStaticMaxDevices) return fail(c, LINKS_E_LIMIT, "static scratch: device ordinal %d >= %d", c->device, kStaticMaxDevices);
  if (!g_static_ptr[c->device]) {
    void* p = nullptr;
    CU(c, cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) return fail(c, LINKS_E_CUDA, "static scratch: cudaGetSymbolAddress returned a null address");
    g_static_ptr[c->device] = static_cast<unsigned char*>(p);
  }
  *out = g_static_ptr[c->device];
  return LINKS_OK;
}

} // namespace

// ====================================================================== C ABI

extern "C" {

const char* links_last_error(const links_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

void* links_host_alloc(size_t bytes)
{
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}
void links_host_free(void* p) { if (p) cudaFreeHost(p); }

int links_bf_size(uint64_t file_bytes, float fpr, uint64_t* m_bits, uint64_t* bytes, uint32_t* hash_num)
{
  if (!file_bytes) return LINKS_E_INVAL;
  // LINKS::make_bf, reference src/LINKS.hpp:297-302 -- same expression shapes, double libm.
  uint64_t m = uint64_t(std::ceil((-1 * double(file_bytes) * std::log(double(fpr))) / (std::log(2.0) * std::log(2.0))));
  m = (uint64_t(m / 8) + 1) * 8;
  const unsigned hash_fct = unsigned(std::floor((double(m) / double(file_bytes)) * std::log(2.0)));
  uint64_t b = m / 8;
  b = uint64_t(std::ceil(double(b) / 8.0) * 8.0); // BloomFilter ctor, btllib bloom_filter.hpp:617-621
  if (m_bits) *m_bits = m;
  if (bytes) *bytes = b;
  if (hash_num) *hash_num = hash_fct;
  return LINKS_OK;
}

size_t links_bf_header(char* dst, size_t cap, uint32_t k, uint32_t hash_num, uint64_t bytes)
{
  // KmerBloomFilter::save -> BloomFilter::save, btllib bloom_filter.hpp:766-783,866-888 (cpptoml table dump)
  int n = snprintf(dst, cap,
                   "[BTLKmerBloomFilter_v6]\n\tk = %u\n\thash_fn = \"ntHash_v1\"\n\thash_num = %u\n\tbytes = %llu\n"
                   "[HeaderEnd]\n\n  <binary data>\n",
                   k, hash_num, (unsigned long long)bytes);
  if (n < 0 || size_t(n) + 49 > cap) return 0;
  for (int i = 0; i < 48; ++i) dst[n++] = '\n';
  dst[n] = 0;
  return size_t(n);
}

int links_create(links_ctx** out, const links_params* p)
{
  if (!out || !p) return fail(nullptr, LINKS_E_INVAL, "links_create: null argument");
  *out = nullptr;
  if (p->k == 0 || p->k > LINKS_MAX_K) return fail(nullptr, LINKS_E_LIMIT, "links_create: k=%u outside 1..%u", p->k, LINKS_MAX_K);
  if (p->n_dist == 0 || p->n_dist > LINKS_MAX_DIST || !p->dist || !p->step)
    return fail(nullptr, LINKS_E_LIMIT, "links_create: need 1..%u distances with matching steps", LINKS_MAX_DIST);
  for (uint32_t i = 0; i < p->n_dist; ++i)
    if (p->step[i] == 0) return fail(nullptr, LINKS_E_INVAL, "links_create: step size 0");
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
    cudaGetLastError();
    return fail(nullptr, LINKS_E_NODEVICE, "links_create: no CUDA device (this library has no CPU path)");
  }
  if (p->device < 0 || p->device >= n_dev) return fail(nullptr, LINKS_E_INVAL, "links_create: device %d of %d", p->device, n_dev);
  links_ctx* c = new (std::nothrow) links_ctx();
  if (!c) return fail(nullptr, LINKS_E_NOMEM, "links_create: out of host memory");
  c->device = p->device;
  c->prm = *p;
  c->dist.assign(p->dist, p->dist + p->n_dist);
  c->step.assign(p->step, p->step + p->n_dist);
  c->prm.dist = c->dist.data();
  c->prm.step = nullptr;
  c->min_allowed.resize(p->n_dist);
  for (uint32_t i = 0; i < p->n_dist; ++i) {
    const int64_t insert_size = c->dist[i];
    c->min_allowed[i] = int(-1 * (p->insert_stdev * insert_size)); // reference src/LINKS.hpp:838 (float product, truncation)
  }
  c->kseed = uint64_t(p->k) * kMultiSeed;
  c->tig_len.assign(1, 0);
#define CUC(call)                                                                                          \
  do {                                                                                                     \
    cudaError_t e__ = (call);                                                                              \
    if (e__ != cudaSuccess) {                                                                              \
      cudaGetLastError();                                                                                  \
      fail(nullptr, LINKS_E_CUDA, "links_create: %s: %s", #call, cudaGetErrorString(e__));                 \
      links_destroy(c);                                                                                    \
      return LINKS_E_CUDA;                                                                                 \
    }                                                                                                      \
  } while (0)
  host_mark(c, "create: device count known (driver initialised)");
  CUC(cudaSetDevice(c->device));
  CUC(cudaFree(nullptr));  // force the primary context now so that the marks below separate it from the module load
  host_mark(c, "create: primary context ready");
  cudaDeviceProp prop;
  CUC(cudaGetDeviceProperties(&prop, c->device));
  c->sm_count = prop.multiProcessorCount;
  c->dev_total_mem = prop.totalGlobalMem;
  CUC(cudaStreamCreateWithFlags(&c->s_compute, cudaStreamNonBlocking));
  CUC(cudaStreamCreateWithFlags(&c->s_copy, cudaStreamNonBlocking));
  CUC(cudaStreamCreateWithFlags(&c->s_aux, cudaStreamNonBlocking));
  CUC(cudaStreamCreateWithFlags(&c->s_pack, cudaStreamNonBlocking));
  CUC(cudaStreamCreateWithFlags(&c->s_import, cudaStreamNonBlocking));
  CUC(cudaStreamCreateWithFlags(&c->s_route, cudaStreamNonBlocking));
  CUC(cudaStreamCreateWithFlags(&c->s_cls, cudaStreamNonBlocking));
  CUC(cudaEventCreateWithFlags(&c->ev_cls_in, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&c->ev_cls_done, cudaEventDisableTiming));
  CUC(c->d_cls.ensure(size_t(links_ctx::kClsRounds) * 8));
  CUC(cudaMemset(c->d_cls.p, 0, size_t(links_ctx::kClsRounds) * 8));
  { const char* e = getenv("LINKS_INCREMENTAL_CLASSIFY"); c->cls_on = !(e && e[0] == '0'); c->cls_valid = c->cls_on; }
  { const char* e = getenv("LINKS_TABLE_MODE"); c->table_mode = !e ? LINKS_TABLE_AUTO : !strcmp(e, "full") ? LINKS_TABLE_FULL : !strcmp(e, "mates") ? LINKS_TABLE_MATES : LINKS_TABLE_AUTO; }
  CUC(cudaEventCreateWithFlags(&c->ev_extracted, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&c->ev_routed, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&c->ev_imported, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&c->ev_table, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&c->ev_contig, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&c->ev_contig_packed, cudaEventDisableTiming));
  CUC(cudaEventCreateWithFlags(&c->ev_cleared, cudaEventDisableTiming));
  for (int i = 0; i < links_ctx::kSlots; ++i) {
    CUC(cudaEventCreateWithFlags(&c->ev_copied[i], cudaEventDisableTiming));
    CUC(cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming));
    CUC(cudaEventCreateWithFlags(&c->ev_packed[i], cudaEventDisableTiming));
  }
  {
    std::vector<ulonglong2> tab(size_t(p->k) * 4);
    build_nthash_table(p->k, tab.data());
    CUC(c->d_tab.ensure(tab.size() * sizeof(ulonglong2)));
    CUC(cudaMemcpy(c->d_tab.p, tab.data(), tab.size() * sizeof(ulonglong2), cudaMemcpyHostToDevice));
    CUC(c->d_dist.ensure(p->n_dist * 4));
    CUC(cudaMemcpy(c->d_dist.p, c->dist.data(), p->n_dist * 4, cudaMemcpyHostToDevice));
    CUC(c->d_step.ensure(p->n_dist * 4));
    CUC(cudaMemcpy(c->d_step.p, c->step.data(), p->n_dist * 4, cudaMemcpyHostToDevice));
    CUC(c->d_minallowed.ensure(p->n_dist * 4));
    CUC(cudaMemcpy(c->d_minallowed.p, c->min_allowed.data(), p->n_dist * 4, cudaMemcpyHostToDevice));
    CUC(c->d_ctr.ensure(sizeof(Counters)));
    CUC(cudaMemset(c->d_ctr.p, 0, sizeof(Counters)));
    std::vector<ulonglong2> tab2(size_t(nthash_pair_count(p->k)) * 16);
    build_nthash_pair_table(p->k, tab2.data());
    CUC(c->d_tab2.ensure(tab2.size() * sizeof(ulonglong2)));
    CUC(cudaMemcpy(c->d_tab2.p, tab2.data(), tab2.size() * sizeof(ulonglong2), cudaMemcpyHostToDevice));
    CUC(c->d_cursor.ensure(64));
    {
      // member map (member_map.cuh): x is needed iff (x+1) % t_i == 0 [an A k-mer] or (x+1-d_i) % t_i == 0 [a B k-mer] for some i
      const char* e = getenv("LINKS_MEMBER_MAP");
      c->map_mode = !e ? -1 : e[0] == '0' ? 0 : 1;
      uint64_t T = 1;
      for (uint32_t i = 0; i < p->n_dist && T <= kMapMaxPeriod; ++i) { uint64_t a = T, b = c->step[i]; while (b) { const uint64_t r = a % b; a = b; b = r; } T = T / a * c->step[i]; }
      if (T <= kMapMaxPeriod && c->map_mode != 0) {
        c->need_idx.assign(size_t(T), int16_t(-1));
        std::vector<uint16_t> res;
        for (uint64_t x = 0; x < T; ++x) {
          bool need = false;
          for (uint32_t i = 0; i < p->n_dist && !need; ++i) {
            const uint64_t t = c->step[i], d = c->dist[i];
            need = (x + 1) % t == 0 || (x + 1 + t - d % t) % t == 0;
          }
          if (need) { c->need_idx[size_t(x)] = int16_t(res.size()); res.push_back(uint16_t(x)); }
        }
        c->map_T = uint32_t(T); c->map_c = uint32_t(res.size());
        CUC(c->d_need_idx.ensure(c->need_idx.size() * 2));
        CUC(cudaMemcpy(c->d_need_idx.p, c->need_idx.data(), c->need_idx.size() * 2, cudaMemcpyHostToDevice));
        CUC(c->d_need_res.ensure(res.size() * 2));
        CUC(cudaMemcpy(c->d_need_res.p, res.data(), res.size() * 2, cudaMemcpyHostToDevice));
      }
    }
    host_mark(c, "create: streams, events, hash tables on device");
    int occ = 0;
    for (uint32_t h : kExtractVariants)
      for (int v = 0; v < 12; ++v) {
        const ExtractFn fn = extract_fn_for(h, (v & 1) != 0, (v & 2) != 0, v >> 2);
        CUC(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(extract_smem_bytes(p->k))));
        CUC(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kCtaThreads, extract_smem_bytes(p->k)));
        c->extract_ctas_per_sm = std::max(occ, 1);
      }
    for (uint32_t h : kExtractVariants)
      for (int v = 0; v < 2; ++v) {
        cudaFuncAttributes fa;
        CUC(cudaFuncGetAttributes(&fa, map_fn_for(h, v != 0)));
      }
    {
      // load every kernel now: with CUDA's lazy module loading a first launch may have to wait for the
      // device to go idle, which must not happen while a peer-wait kernel (exchange.cuh) is resident
      cudaFuncAttributes fa;
      const void* fns[] = {(const void*)pack_kernel, (const void*)rank_build_kernel, (const void*)table_init_kernel,
                           (const void*)contig_kernel<0>, (const void*)contig_kernel<1>, (const void*)contig_kernel<2>,
                           (const void*)contig_kernel<3>, (const void*)table_seed_kernel, (const void*)bloom_or_reduce_kernel, (const void*)bloom_partition_kernel<true>, (const void*)bloom_partition_kernel<false>, (const void*)bloom_apply_kernel,
                           (const void*)pair_rehash_kernel, (const void*)classify_kernel, (const void*)mirror_expand_kernel,
                           (const void*)gather_rows_kernel, (const void*)pair_bucket_count_kernel, (const void*)pair_bucket_scatter_kernel,
                           (const void*)pair_import_kernel, (const void*)pair_export_kernel, (const void*)route_kernel,
                           (const void*)commit_kernel, (const void*)inbox_wait_kernel, (const void*)inbox_snapshot_kernel,
                           (const void*)inbox_import_kernel, (const void*)rows_send_kernel,
                           (const void*)rows_wait_kernel, (const void*)rows_gather_kernel, (const void*)scan_sums_kernel,
                           (const void*)scan_apply_kernel, (const void*)scan_small_kernel, (const void*)radix_hist_kernel,
                           (const void*)radix_scatter_kernel, (const void*)head_flags_kernel, (const void*)reduce_by_key_kernel,
                           (const void*)rows_segment_sum_kernel};
      host_mark(c, "create: extract variants configured");
      for (const void* f : fns) CUC(cudaFuncGetAttributes(&fa, f));
      host_mark(c, "create: all kernels loaded");
    }
    CUC(cudaHostAlloc(reinterpret_cast<void**>(&c->h_snap), sizeof(Counters) * links_ctx::kSnaps, cudaHostAllocDefault));
    for (int i = 0; i < links_ctx::kSnaps; ++i) CUC(cudaEventCreateWithFlags(&c->ev_snap[i], cudaEventDisableTiming));
  }
#undef CUC
  *out = c;
  return LINKS_OK;
}

void links_destroy(links_ctx* c)
{
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  resolve_spans(c);
  for (auto e : c->ev_pool) cudaEventDestroy(e);
  for (auto* b : c->contig_batches) { b->release(); delete b; }
  for (auto* b : c->contig_pool) { b->release(); delete b; }
  if (c->ev_contig) cudaEventDestroy(c->ev_contig);
  if (c->ev_contig_packed) cudaEventDestroy(c->ev_contig_packed);
  if (c->ev_cleared) cudaEventDestroy(c->ev_cleared);
  for (int i = 0; i < links_ctx::kSlots; ++i) {
    c->slot[i].release(); c->slot_ascii[i].release();
    if (c->ev_copied[i]) cudaEventDestroy(c->ev_copied[i]);
    if (c->ev_done[i]) cudaEventDestroy(c->ev_done[i]);
    if (c->ev_packed[i]) cudaEventDestroy(c->ev_packed[i]);
  }
  c->contig_ascii.release();
  for (int i = 0; i < links_ctx::kSnaps; ++i) if (c->ev_snap[i]) cudaEventDestroy(c->ev_snap[i]);
  if (c->h_snap) cudaFreeHost(c->h_snap);
  if (c->h_rows) cudaFreeHost(c->h_rows);
  for (DevBuf* b : {&c->f_keys[0], &c->f_keys[1], &c->f_vals[0], &c->f_vals[1], &c->f_issues, &c->f_tiglen, &c->f_hist,
                    &c->f_scan_tmp, &c->f_flags, &c->f_total, &c->f_okeys, &c->f_olinks, &c->f_ogaps, &c->f_final, &c->f_exch, &c->f_il, &c->f_ig}) b->release();
  c->d_tab.release(); c->d_tab2.release(); c->d_cursor.release(); c->d_dist.release(); c->d_step.release(); c->d_minallowed.release();
  c->d_table_occ.release(); c->d_bf_stage.release(); c->d_bf_cursor.release(); c->d_bloom.release(); c->d_pairs.release(); c->d_pair_list.release(); c->d_ctr.release(); c->d_table.release();
  if (c->s_compute && c->own_compute_stream) cudaStreamDestroy(c->s_compute);
  if (c->s_copy) cudaStreamDestroy(c->s_copy);
  if (c->s_aux) cudaStreamDestroy(c->s_aux);
  if (c->s_pack) cudaStreamDestroy(c->s_pack);
  if (c->s_import) cudaStreamDestroy(c->s_import);
  if (c->s_route) cudaStreamDestroy(c->s_route);
  if (c->s_cls) cudaStreamDestroy(c->s_cls);
  if (c->ev_cls_in) cudaEventDestroy(c->ev_cls_in);
  if (c->ev_cls_done) cudaEventDestroy(c->ev_cls_done);
  c->d_cls.release();
  if (c->ev_extracted) cudaEventDestroy(c->ev_extracted);
  if (c->ev_routed) cudaEventDestroy(c->ev_routed);
  if (c->ev_imported) cudaEventDestroy(c->ev_imported);
  c->d_xdev.release(); c->d_owned.release(); c->d_owned_list.release(); c->f_rowcnt.release();
  c->d_members.release(); c->d_need_idx.release(); c->d_need_res.release();
  if (c->ev_table) cudaEventDestroy(c->ev_table);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  cudaGetLastError();
  delete c;
}

int links_sync(links_ctx* c)
{
  if (!c) return LINKS_E_INVAL;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->s_copy));
  CU(c, cudaStreamSynchronize(c->s_aux));
  CU(c, cudaStreamSynchronize(c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_route));
  CU(c, cudaStreamSynchronize(c->s_cls));
  CU(c, cudaStreamSynchronize(c->s_import));
  return LINKS_OK;
}

int links_bf_alloc(links_ctx* c, uint64_t bytes, uint32_t hash_num)
{
  if (!c) return LINKS_E_INVAL;
  if (bytes == 0 || (bytes & 7) || hash_num == 0 || hash_num > LINKS_MAX_HASH_NUM)
    return fail(c, LINKS_E_INVAL, "links_bf_alloc: bytes must be a positive multiple of 8 and hash_num in 1..%u", LINKS_MAX_HASH_NUM);
  CU(c, cudaSetDevice(c->device));
  CU(c, c->d_bloom.ensure(bytes));
  CU(c, cudaMemsetAsync(c->d_bloom.p, 0, bytes, c->s_compute));
  c->bloom_ext = nullptr; c->bf_share_n = 1; c->bf_share_rank = 0; c->bf_reduced = true;
  c->bf_bytes = bytes; c->hash_num = hash_num; c->bf_ready = true; c->bf_insert = true;
  return LINKS_OK;
}

int links_bf_alloc_shared(links_ctx* c, uint64_t bytes, uint32_t hash_num, uint32_t rank, uint32_t n_ranks, void* const* bloom)
{
  if (!c || !bloom) return LINKS_E_INVAL;
  if (bytes == 0 || (bytes & 7) || hash_num == 0 || hash_num > LINKS_MAX_HASH_NUM)
    return fail(c, LINKS_E_INVAL, "links_bf_alloc_shared: bytes must be a positive multiple of 8 and hash_num in 1..%u", LINKS_MAX_HASH_NUM);
  if (n_ranks == 0 || n_ranks > uint32_t(kMaxP2PRanks) || rank >= n_ranks)
    return fail(c, LINKS_E_LIMIT, "links_bf_alloc_shared: rank %u of %u (at most %d ranks)", rank, n_ranks, kMaxP2PRanks);
  for (uint32_t r = 0; r < n_ranks; ++r)
    if (!bloom[r] || (reinterpret_cast<uintptr_t>(bloom[r]) & 15)) return fail(c, LINKS_E_INVAL, "links_bf_alloc_shared: bloom[%u] is null or not 16-byte aligned", r);
  if (!c->contig_batches.empty()) return fail(c, LINKS_E_INVAL, "links_bf_alloc_shared: call before the job's first links_contigs_push");
  CU(c, cudaSetDevice(c->device));
  for (uint32_t r = 0; r < n_ranks; ++r) c->bloom_peers[r] = bloom[r];
  c->bloom_ext = bloom[rank];
  c->bf_share_rank = rank; c->bf_share_n = n_ranks; c->bf_reduced = n_ranks == 1;
  CU(c, cudaMemsetAsync(c->bloom_ext, 0, bytes, c->s_compute));
  c->bf_bytes = bytes; c->hash_num = hash_num; c->bf_ready = true; c->bf_insert = true;
  c->member_off = n_ranks > 1;  // the exact member set would hold this rank's share only: every k-mer takes the filter's own bits
  return LINKS_OK;
}

int links_bf_allreduce(links_ctx* c)
{
  if (!c) return LINKS_E_INVAL;
  if (!c->bloom_ext) return fail(c, LINKS_E_INVAL, "links_bf_allreduce: the filter was not made by links_bf_alloc_shared");
  if (c->bf_reduced) return LINKS_OK;
  NvtxRange nvtx_("links_bf_allreduce");
  CU(c, cudaSetDevice(c->device));
  BloomPeers bp;
  for (uint32_t r = 0; r < c->bf_share_n; ++r) bp.p[r] = static_cast<uint4*>(c->bloom_peers[r]);
  const uint64_t n16 = c->bf_bytes / 16, tail8 = (c->bf_bytes % 16) / 8;  // bytes is a multiple of 8
  const uint64_t lo = n16 * c->bf_share_rank / c->bf_share_n, hi = n16 * (c->bf_share_rank + 1) / c->bf_share_n;
  {
    Span sp(c, F_BF, c->s_compute);
    bloom_or_reduce_kernel<<<c->sm_count * 8, kCtaThreads, 0, c->s_compute>>>(bp, c->bf_share_n, lo, hi,
                                                                             c->bf_share_rank == c->bf_share_n - 1 ? tail8 : 0, n16);
    c->launches++;
  }
  CU(c, cudaGetLastError());
  CU(c, cudaStreamSynchronize(c->s_compute));  // the caller's barrier that follows must mean "my slice is written everywhere"
  c->bf_reduced = true;
  return LINKS_OK;
}

int links_bf_load(links_ctx* c, const uint8_t* bits, uint64_t bytes, uint32_t hash_num)
{
  if (!c || !bits) return LINKS_E_INVAL;
  if (bytes == 0 || (bytes & 3) || hash_num == 0 || hash_num > LINKS_MAX_HASH_NUM)
    return fail(c, LINKS_E_INVAL, "links_bf_load: bytes must be a positive multiple of 4 and hash_num in 1..%u", LINKS_MAX_HASH_NUM);
  CU(c, cudaSetDevice(c->device));
  CU(c, c->d_bloom.ensure(bytes));
  c->bloom_ext = nullptr; c->bf_share_n = 1; c->bf_share_rank = 0; c->bf_reduced = true;
  CU(c, cudaMemcpyAsync(c->d_bloom.p, bits, bytes, cudaMemcpyHostToDevice, c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_compute));
  c->bf_bytes = bytes; c->hash_num = hash_num; c->bf_ready = true; c->bf_insert = false;
  return LINKS_OK;
}

int links_bf_covers_contigs(links_ctx* c, int yes)
{
  if (!c) return LINKS_E_INVAL;
  if (!c->bf_ready) return fail(c, LINKS_E_INVAL, "links_bf_covers_contigs: no Bloom filter");
  if (!c->contig_batches.empty()) return fail(c, LINKS_E_INVAL, "links_bf_covers_contigs: call before the first links_contigs_push");
  c->bf_covers_contigs = yes != 0;
  return LINKS_OK;
}

int links_bf_download(links_ctx* c, uint8_t* dst)
{
  if (!c || !dst) return LINKS_E_INVAL;
  if (!c->bf_ready) return fail(c, LINKS_E_INVAL, "links_bf_download: no Bloom filter");
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaMemcpyAsync(dst, bloom_view(c).words, c->bf_bytes, cudaMemcpyDeviceToHost, c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_compute));
  return LINKS_OK;
}

static int contigs_push_impl(links_ctx* c, const char* seq, const uint8_t* dev_seq, const uint64_t* off, uint32_t n,
                             uint32_t first_tig_id)
{
  NvtxRange nvtx_("links_contigs_push");
  if (!c || !off || (!seq && !dev_seq && n && off[n] != off[0])) return LINKS_E_INVAL;
  if (!c->bf_ready) return fail(c, LINKS_E_INVAL, "links_contigs_push: call links_bf_alloc or links_bf_load first");
  if (first_tig_id == 0) return fail(c, LINKS_E_INVAL, "links_contigs_push: contig ids are 1-based");
  if (n == 0) return LINKS_OK;
  if (uint64_t(first_tig_id) + n - 1 > LINKS_MAX_TIGS) return fail(c, LINKS_E_LIMIT, "links_contigs_push: more than %u contigs", LINKS_MAX_TIGS);
  CU(c, cudaSetDevice(c->device));
  const uint32_t k = c->prm.k;
  std::vector<uint32_t> segs(size_t(n) + 1);
  uint64_t tiles = 0, positions = 0;
  if (c->cls_started) { CU(c, cudaStreamSynchronize(c->s_cls)); c->cls_valid = false; }  // contigs after the first reads: stage 4 starts over at finalize
  if (c->tig_len.size() < size_t(first_tig_id) + n) c->tig_len.resize(size_t(first_tig_id) + n, 0);
  for (uint32_t i = 0; i < n; ++i) {
    if (off[i + 1] < off[i]) return fail(c, LINKS_E_INVAL, "links_contigs_push: offsets not monotone");
    const uint64_t len = off[i + 1] - off[i];
    if (len >= (1ull << 32)) return fail(c, LINKS_E_LIMIT, "links_contigs_push: contig longer than 2^32-1");
    segs[i] = uint32_t(tiles);
    if (len >= k) {
      c->tig_len[first_tig_id + i] = uint32_t(len); // reference src/LINKS.hpp:655-658
      tiles += (len - k + 1 + kTile - 1) / kTile;
      positions += len - k + 1;
    }
    if (tiles >= (1ull << 31)) return fail(c, LINKS_E_LIMIT, "links_contigs_push: batch too large, push fewer contigs per call");
  }
  segs[n] = uint32_t(tiles);
  // reuse a released batch object (links_reset keeps their device buffers)
  SeqBatchDev* b = nullptr;
  if (!c->contig_pool.empty()) { b = c->contig_pool.back(); c->contig_pool.pop_back(); }
  else b = new (std::nothrow) SeqBatchDev();
  if (!b) return fail(c, LINKS_E_NOMEM, "links_contigs_push: out of host memory");
  b->first_id = first_tig_id;
  cudaEvent_t ev = c->ev_contig;
  // the single contig staging buffer is reused: wait until the previous pack consumed it
  if (!dev_seq && c->contig_pack_pending) CU(c, cudaEventSynchronize(c->ev_contig_packed));
  int rc = upload_and_pack(c, *b, c->contig_ascii, seq, dev_seq, off, n, segs, false, c->s_copy, c->s_compute, ev);
  if (rc) { b->release(); delete b; return rc; }
  CU(c, cudaEventRecord(c->ev_contig_packed, c->s_compute));
  c->contig_pack_pending = true;
  c->contig_batches.push_back(b);
  c->contig_positions += positions;
  c->n_tigs_seen = std::max(c->n_tigs_seen, first_tig_id + n - 1);
  if ((c->bf_insert || c->bf_covers_contigs) && b->n_tiles) {
    Span sp(c, F_BF, c->s_compute);
    rc = ensure_member_capacity(c);  // may re-insert the earlier batches into a larger set
    if (rc) return rc;
    ContigBatch cb{b->packed.as<uint32_t>(), b->inv.as<uint32_t>(), b->off.as<uint64_t>(), b->base0,
                   b->segs.as<uint32_t>(), b->n_seq, b->n_tiles, b->first_id};
    if (c->bf_insert) {
      // a filter beyond L2 is filled block by block (BloomPass): LINKS_BF_BLOCK_MB sets the block size, 0 = always one pass
      static const uint64_t block_bytes = []() {
        const char* e = getenv("LINKS_BF_BLOCK_MB");
        return uint64_t((e ? strtod(e, nullptr) : 64.0) * double(1 << 20));  // fractions allowed (tests: 0.25)
      }();
      const uint64_t n_words = (c->bf_bytes + 3) / 4;
      // LINKS_BF_PARTITION=0 keeps the older block passes / single DRAM pass below; default: filters beyond L2 are built by
      // partition + apply (every k-mer hashed once, its updates streamed into one bucket per L2-sized block, then replayed)
      static const int part_env = []() { const char* e = getenv("LINKS_BF_PARTITION"); return !e ? -1 : e[0] == '0' ? 0 : 1; }();
      static const uint64_t stage_bytes = []() { const char* e = getenv("LINKS_BF_STAGE_MB"); return uint64_t(e ? strtod(e, nullptr) : 4096.0) << 20; }();
      bool partitioned = false;
      // (every block of the older form is another full hashing pass over the contigs.  With the replay launched once per
      //  bucket: C3 / C5, 177 MB, 12.6 ms against 14.5-14.8 ms in three passes; C4 at a tenth, 560 MB, 42 ms against 136 ms
      //  in one DRAM pass.  Two passes are left as they are: not measured against the partition.)
      if (block_bytes && (part_env == 1 || (part_env < 0 && c->bf_bytes > 2 * block_bytes + block_bytes / 2))) {
        uint32_t block_shift = uint32_t(log2_ceil(std::max<uint64_t>(block_bytes / 4, 64)));
        while (((n_words >> block_shift) + 1 > uint64_t(kMaxBloomBuckets)) && block_shift < 27) ++block_shift;
        const uint32_t n_buckets = uint32_t((n_words + (1ull << block_shift) - 1) >> block_shift);
        if (n_buckets <= uint32_t(kMaxBloomBuckets)) {
          partitioned = true;
          const uint32_t share_n = c->bf_share_n;
          // tiles per batch so that the entries of one batch fit the staging budget
          const uint64_t budget_entries = std::max<uint64_t>(stage_bytes / 4, 1ull << 20);
          uint64_t tiles_per_batch = std::max<uint64_t>(budget_entries * 10 / 11 / (uint64_t(kTile) * c->hash_num), 1) * share_n;
          tiles_per_batch = std::min<uint64_t>(tiles_per_batch, b->n_tiles);
          const double own_positions = double(tiles_per_batch) * kTile / share_n;
          const double per_bucket = own_positions * c->hash_num / (double(n_words) / double(1ull << block_shift));
          const uint64_t cap = uint64_t(per_bucket * 1.1) + (1ull << 16);
          CU(c, c->d_bf_stage.ensure(uint64_t(n_buckets) * cap * 4));
          CU(c, c->d_bf_cursor.ensure(size_t(kMaxBloomBuckets) * 8 + 64));
          unsigned int* d_over = reinterpret_cast<unsigned int*>(c->d_bf_cursor.as<unsigned long long>() + kMaxBloomBuckets);
          CU(c, cudaMemsetAsync(d_over, 0, 4, c->s_compute));
          BloomBuckets bk{c->d_bf_stage.as<uint32_t>(), c->d_bf_cursor.as<unsigned long long>(), cap, n_buckets, block_shift, d_over};
          const bool with_set = c->member_buckets != 0;
          for (uint64_t t0 = 0; t0 < b->n_tiles; t0 += tiles_per_batch) {
            const uint32_t t1 = uint32_t(std::min<uint64_t>(t0 + tiles_per_batch, b->n_tiles));
            CU(c, cudaMemsetAsync(c->d_bf_cursor.p, 0, size_t(kMaxBloomBuckets) * 8, c->s_compute));
            const int grid = grid_for(c, (t1 - t0 + share_n - 1) / share_n, 1, 8);
            if (with_set)
              bloom_partition_kernel<true><<<grid, kCtaThreads, k * 4 * sizeof(ulonglong2), c->s_compute>>>(
                cb, uint32_t(t0), t1, int(k), c->kseed, c->d_tab.as<ulonglong2>(), bloom_view(c), member_view(c), bk, c->bf_share_rank, share_n);
            else
              bloom_partition_kernel<false><<<grid, kCtaThreads, k * 4 * sizeof(ulonglong2), c->s_compute>>>(
                cb, uint32_t(t0), t1, int(k), c->kseed, c->d_tab.as<ulonglong2>(), bloom_view(c), member_view(c), bk, c->bf_share_rank, share_n);
            for (uint32_t bb = 0; bb < n_buckets; ++bb) bloom_apply_kernel<<<c->sm_count * 8, kCtaThreads, 0, c->s_compute>>>(bloom_view(c), bk, bb);
            c->launches += 1 + n_buckets;
          }
          CU(c, cudaGetLastError());
          unsigned int h_over = 0;  // a bucket that overflowed dropped updates: redo the batch with direct atomics (idempotent)
          CU(c, cudaMemcpyAsync(&h_over, d_over, 4, cudaMemcpyDeviceToHost, c->s_compute));
          CU(c, cudaStreamSynchronize(c->s_compute));
          if (h_over) {
            contig_kernel<0><<<grid_for(c, b->n_tiles, 1, 8), kCtaThreads, k * 4 * sizeof(ulonglong2), c->s_compute>>>(
              cb, int(k), c->kseed, c->d_tab.as<ulonglong2>(), bloom_view(c), TableView{nullptr, 0, 0}, member_view(c), c->d_ctr.as<Counters>(),
              BloomPass{0, ~0ull, 0, c->bf_share_rank, c->bf_share_n});
            c->launches++;
          }
        }
      }
      uint64_t passes = block_bytes && c->bf_bytes > block_bytes + block_bytes / 2 ? (c->bf_bytes + block_bytes - 1) / block_bytes : 1;
      // every pass hashes all k-mers again (~20 G positions/s with the atomics in L2) against ~3.2 G positions/s for one
      // pass with DRAM-resident atomics: beyond six blocks the single pass is the cheaper one
      static const uint64_t max_passes = []() { const char* e = getenv("LINKS_BF_MAX_PASSES"); return e ? strtoull(e, nullptr, 10) : 6ull; }();  // (only reached with LINKS_BF_PARTITION=0 beyond four blocks)
      if (passes > max_passes) passes = 1;
      for (uint64_t ps = 0; ps < passes && !partitioned; ++ps) {
        const BloomPass bp{n_words * ps / passes, ps + 1 == passes ? ~0ull : n_words * (ps + 1) / passes, ps == 0, c->bf_share_rank, c->bf_share_n};
        contig_kernel<0><<<grid_for(c, b->n_tiles, 1, 8), kCtaThreads, k * 4 * sizeof(ulonglong2), c->s_compute>>>(
          cb, int(k), c->kseed, c->d_tab.as<ulonglong2>(), bloom_view(c), TableView{nullptr, 0, 0}, member_view(c), c->d_ctr.as<Counters>(), bp);
        c->launches++;
      }
    } else if (c->member_buckets) {
      contig_kernel<2><<<grid_for(c, b->n_tiles, 1, 8), kCtaThreads, k * 4 * sizeof(ulonglong2), c->s_compute>>>(
        cb, int(k), c->kseed, c->d_tab.as<ulonglong2>(), bloom_view(c), TableView{nullptr, 0, 0}, member_view(c), c->d_ctr.as<Counters>(),
        BloomPass{0, ~0ull, 1});
      c->launches++;
    }
    c->member_positions = c->contig_positions;
    CU(c, cudaGetLastError());
  }
  CU(c, cudaStreamSynchronize(c->s_copy)); // host buffers reusable
  return LINKS_OK;
}

int links_contigs_push(links_ctx* c, const char* seq, const uint64_t* off, uint32_t n, uint32_t first_tig_id)
{
  return contigs_push_impl(c, seq, nullptr, off, n, first_tig_id);
}
int links_contigs_push_dev(links_ctx* c, const void* dev_seq, const uint64_t* off, uint32_t n, uint32_t first_tig_id)
{
  if (!dev_seq) return LINKS_E_INVAL;
  return contigs_push_impl(c, nullptr, static_cast<const uint8_t*>(dev_seq), off, n, first_tig_id);
}

// map_base / map_segs: empty = this chunk walks the filter per candidate; else (n+1 entries each) the member map's layout
static int push_read_chunk(links_ctx* c, const char* seq, const uint8_t* dev_seq, const uint64_t* off, uint32_t n,
                           uint64_t first_read_idx, const std::vector<uint32_t>& segs, uint64_t cand, uint64_t later,
                           const std::vector<unsigned long long>& map_base, const std::vector<uint32_t>& map_segs)
{
  const int si = c->next_slot;
  c->next_slot = (si + 1) % links_ctx::kSlots;
  if (c->slot_used[si]) {
    // the slot's previous chunk must have been consumed by its kernels before we overwrite it
    CU(c, cudaEventSynchronize(c->ev_done[si]));
  }
  CU(c, cudaStreamWaitEvent(c->s_compute, c->ev_cleared, 0));  // links_reset clears the pair set on s_aux
  const bool fused = c->x_attached && c->x_fused;  // no local pair set: survivors go straight to their owners
  int rc = fused ? LINKS_OK : ensure_pair_capacity(c, cand, later);
  if (rc) return rc;
  SeqBatchDev& b = c->slot[si];
  b.first_read_idx = first_read_idx;
  const bool mapped = !map_base.empty() && map_base.back() > 0;
  if (mapped) {  // (on s_copy in front of upload_and_pack's copies: covered by the same ev_copied)
    CU(c, b.mm_bytes.ensure(map_base.back()));
    CU(c, b.mm_base.ensure(map_base.size() * 8));
    CU(c, b.mm_segs.ensure(map_segs.size() * 4));
    CU(c, cudaMemcpyAsync(b.mm_base.p, map_base.data(), map_base.size() * 8, cudaMemcpyHostToDevice, c->s_copy));
    CU(c, cudaMemcpyAsync(b.mm_segs.p, map_segs.data(), map_segs.size() * 4, cudaMemcpyHostToDevice, c->s_copy));
  }
  // copy on s_copy, pack on s_pack, extract on s_compute: chunk i+1 is packed while chunk i is being extracted
  rc = upload_and_pack(c, b, c->slot_ascii[si], seq, dev_seq, off, n, segs, true, c->s_copy, c->s_pack, c->ev_copied[si]);
  if (rc) return rc;
  CU(c, cudaEventRecord(c->ev_packed[si], c->s_pack));
  CU(c, cudaStreamWaitEvent(c->s_compute, c->ev_packed[si], 0));
  c->cand_total += cand;
  if (b.n_tiles) {
    {
      Span sp(c, F_EXTRACT, c->s_compute);
      ReadBatch rb{b.packed.as<uint32_t>(), b.inv.as<uint32_t>(), b.off.as<uint64_t>(), b.base0, b.segs.as<uint32_t>(),
                   b.n_seq, b.n_tiles, b.flags.as<uint32_t>(), b.vs.as<unsigned long long>(), b.vrank.as<uint32_t>(),
                   first_read_idx, c->d_cursor.as<unsigned int>()};
      DistParams dp{c->d_dist.as<uint32_t>(), c->d_step.as<uint32_t>(), c->prm.n_dist, c->prm.offset};
      CU(c, cudaMemsetAsync(c->d_cursor.p, 0, sizeof(unsigned int), c->s_compute));
      MemberMap mm{nullptr, nullptr, nullptr, nullptr, 0, 0, 0, 0};
      if (mapped) {
        mm = MemberMap{b.mm_bytes.as<uint8_t>(), b.mm_base.as<unsigned long long>(), c->d_need_idx.as<int16_t>(), c->d_need_res.as<uint16_t>(),
                       c->map_T, c->map_c, map_magic(c->map_T), map_magic(c->map_c)};
        MapBatch mb{b.packed.as<uint32_t>(), b.off.as<uint64_t>(), b.base0, b.mm_segs.as<uint32_t>(), b.n_seq, map_segs.back(), b.flags.as<uint32_t>()};
        map_fn_for(c->hash_num, bloom_view(c).bits32 != 0)<<<grid_for(c, map_segs.back(), 1, 8), kCtaThreads,
                                                           size_t((c->prm.k + 1) / 2) * 16 * sizeof(ulonglong2), c->s_compute>>>(
          mb, mm, int(c->prm.k), c->prm.offset, c->kseed, c->d_tab2.as<ulonglong2>(), bloom_view(c), member_view(c), c->d_ctr.as<Counters>());
        c->launches++;
      }
      // probes one at a time (fewest sectors) once the filter lives in DRAM; grouped (fewest round trips) while it sits in L2
      static const int seq_env = []() { const char* e = getenv("LINKS_SEQ_PROBES"); return !e ? -1 : e[0] == '0' ? 0 : 1; }();
      const bool seq = seq_env >= 0 ? seq_env == 1 : c->bf_bytes > (64ull << 20);
      extract_fn_for(c->hash_num, fused, bloom_view(c).bits32 != 0, mapped ? 1 : seq ? 2 : 0)<<<grid_for(c, b.n_tiles, kWarpsPerCta, c->extract_ctas_per_sm), kCtaThreads,
                                           extract_smem_bytes(c->prm.k), c->s_compute>>>(
        rb, dp, int(c->prm.k), c->kseed, c->d_tab2.as<ulonglong2>(), bloom_view(c), member_view(c),
        pair_view(c), c->d_ctr.as<Counters>(), c->d_xdev.as<ExchangeDev>(), c->x_lay, c->x_epoch, mm);
      c->launches++;
      c->extract_launches++;
      CU(c, cudaGetLastError());
    }
    rc = enqueue_snapshot(c);
    if (rc) return rc;
    rc = classify_increment(c);
    if (rc) return rc;
    if (c->x_attached) {
      // the owners insert this launch's pairs while the next chunk is extracted
      rc = fused ? exchange_after_routed_extract(c, false) : exchange_route(c, false);
      if (rc) return rc;
    }
  }
  CU(c, cudaEventRecord(c->ev_done[si], c->s_compute));
  c->slot_used[si] = true;
  return LINKS_OK;
}

static int start_table_build(links_ctx* c);

static int reads_push_impl(links_ctx* c, const char* seq, const uint8_t* dev_seq, const uint64_t* off, uint32_t n,
                           uint64_t first_read_idx)
{
  NvtxRange nvtx_("links_reads_push");
  if (!c || !off || (!seq && !dev_seq && n && off[n] != off[0])) return LINKS_E_INVAL;
  if (!c->bf_ready) return fail(c, LINKS_E_INVAL, "links_reads_push: no Bloom filter yet");
  if (c->finalized) return fail(c, LINKS_E_INVAL, "links_reads_push: context already finalized (links_reset first)");
  if (!c->bf_reduced) return fail(c, LINKS_E_INVAL, "links_reads_push: links_bf_allreduce has not been called on the shared Bloom filter");
  if (n == 0) return LINKS_OK;
  // the 64-bit order key (28 bits of read ordinal, 8 of distance index, 28 of k-mer rank) only exists with several distances
  const bool okey_limits = c->prm.n_dist > 1;
  if (okey_limits && first_read_idx + n >= (1ull << 28)) return fail(c, LINKS_E_LIMIT, "links_reads_push: more than 2^28 reads with several -d values");
  CU(c, cudaSetDevice(c->device));
  {
    const int rc = start_table_build(c);
    if (rc) return rc;
  }
  const uint32_t k = c->prm.k, nd = c->prm.n_dist;
  const uint64_t o = c->prm.offset;
  // chunk the push at record boundaries so copies overlap kernels
  // Chunk size in bases.  Host pushes ramp up (8 MB, 16, 32, then 64 MB chunks) so the first copy that
  // nothing can overlap is short while later launches are long enough to amortise their tails;
  // device-resident pushes have no copy to overlap and take 64 MB chunks from the start.
  static const uint64_t kChunkMax = []() {
    const char* e = getenv("LINKS_CHUNK_BASES");
    return e ? strtoull(e, nullptr, 10) : (64ull << 20);
  }();
  // Device-resident pushes go in equal chunks of the maximum size.  LINKS_CHUNK_SHAPE=taper makes them ramp up and taper off
  // like host pushes (a short first pack, a short last classify); measured on C2: 5.51 ms per job against 5.37 flat -- the
  // extra launches cost more than the shorter tail saves -- so it is not the default.
  static const bool kFlat = []() { const char* e = getenv("LINKS_CHUNK_SHAPE"); return !(e && !strcmp(e, "taper")); }();
  uint64_t kChunkBases = dev_seq ? (kFlat ? kChunkMax : std::min<uint64_t>(kChunkMax, 32ull << 20)) : std::min<uint64_t>(kChunkMax, 8ull << 20);
  const uint64_t kChunkCand = 1ull << 27;
  // upper bound on candidates per (read, distance): exact when the read has no invalid base (reference src/LINKS.hpp:525-543)
  auto seg_count = [&](uint64_t len, uint32_t di) -> uint64_t {
    if (len < k) return 0;
    const uint64_t npos = len - k + 1, d = c->dist[di];
    return npos > o + d ? (npos - o - d) / c->step[di] : 0;
  };
  uint64_t total_cand = 0;
  for (uint32_t i = 0; i < n; ++i) {
    if (off[i + 1] < off[i]) return fail(c, LINKS_E_INVAL, "links_reads_push: offsets not monotone");
    const uint64_t len = off[i + 1] - off[i];
    if (okey_limits && len >= (1ull << 28)) return fail(c, LINKS_E_LIMIT, "links_reads_push: read longer than 2^28-1 with several -d values");
    if (len >= (1ull << 32)) return fail(c, LINKS_E_LIMIT, "links_reads_push: read longer than 2^32-1");
    for (uint32_t di = 0; di < nd; ++di) total_cand += seg_count(len, di);
  }
  c->x_cand_hint = std::max(c->x_cand_hint, total_cand);
  if (nd > 1 && first_read_idx < c->max_read_end) c->cls_valid = false;  // an earlier ordinal may still take a classified pair's distance
  c->max_read_end = std::max<uint64_t>(c->max_read_end, first_read_idx + n);
  uint32_t i0 = 0;
  uint64_t done_cand = 0;
  std::vector<uint32_t> segs, map_segs;
  std::vector<unsigned long long> map_base;
  const bool map_possible = c->map_T && c->hash_num <= 127 && c->map_mode != 0;
  while (i0 < n) {
    segs.clear(); map_segs.clear(); map_base.clear();
    uint64_t tiles = 0, cand = 0, map_entries = 0, map_tiles = 0;
    uint32_t i = i0;
    // host pushes also taper off: the last chunk's kernels run after its copy with nothing left to overlap
    uint64_t chunk_limit = dev_seq && kFlat ? kChunkBases : std::min(kChunkBases, std::max<uint64_t>((off[n] - off[i0]) / 2, dev_seq ? 16ull << 20 : 8ull << 20));
    // with an exchange attached the records of the LAST launch are imported, and their pairs classified, after every rank has
    // finished: a short last chunk is a short tail for all ranks (LINKS_EXCHANGE_LAST_CHUNK bases; 0 = off)
    static const uint64_t kLastChunk = []() { const char* e = getenv("LINKS_EXCHANGE_LAST_CHUNK"); return e ? strtoull(e, nullptr, 10) : 0ull; }();
    if (c->x_attached && kLastChunk) {
      const uint64_t remaining = off[n] - off[i0];
      if (remaining > kLastChunk + (kLastChunk >> 1) && remaining <= chunk_limit) chunk_limit = remaining - kLastChunk;
    }
    for (; i < n; ++i) {
      const uint64_t len = off[i + 1] - off[i];
      if (i > i0 && (off[i + 1] - off[i0] > chunk_limit || cand > kChunkCand || uint64_t(i - i0) * nd > (1ull << 26))) break;
      for (uint32_t di = 0; di < nd; ++di) {
        segs.push_back(uint32_t(tiles));
        const uint64_t cnt = seg_count(len, di);
        tiles += (cnt + kExtractTile - 1) / kExtractTile;
        cand += cnt;
      }
      if (tiles >= (1ull << 31)) return fail(c, LINKS_E_LIMIT, "links_reads_push: chunk too large");
      if (map_possible) {  // member map: c entries per period of T positions over x = position - offset in [0, npos - offset)
        map_base.push_back(map_entries);
        map_segs.push_back(uint32_t(map_tiles));
        const uint64_t X = len >= k && len - k + 1 > o ? len - k + 1 - o : 0;
        const uint64_t e = (X + c->map_T - 1) / c->map_T * c->map_c;
        map_entries += e;
        map_tiles += (e + kMapTile - 1) / kMapTile;
      }
    }
    segs.push_back(uint32_t(tiles));
    if (map_possible) {
      map_base.push_back(map_entries);
      map_segs.push_back(uint32_t(map_tiles));
      // One filter walk per map entry against ~1.45 per candidate (B is only asked when A passes, ~45 % on ONT-like reads):
      // the map pays with several distances that share positions, not for one distance on reads barely longer than it
      const bool pays = c->map_mode == 1 || double(map_entries) < 0.8 * 1.45 * double(cand);
      if (!pays || map_tiles >= (1ull << 31)) { map_base.clear(); map_segs.clear(); }
    }
    done_cand += cand;
    int rc = push_read_chunk(c, seq, dev_seq, off + i0, i - i0, first_read_idx + i0, segs, cand, total_cand - done_cand, map_base, map_segs);
    if (rc) return rc;
    i0 = i;
    kChunkBases = std::min(kChunkMax, kChunkBases * 2);
  }
  CU(c, cudaStreamSynchronize(c->s_copy)); // host buffers reusable
  return LINKS_OK;
}

int links_reads_push(links_ctx* c, const char* seq, const uint64_t* off, uint32_t n, uint64_t first_read_idx)
{
  return reads_push_impl(c, seq, nullptr, off, n, first_read_idx);
}
int links_reads_push_dev(links_ctx* c, const void* dev_seq, const uint64_t* off, uint32_t n, uint64_t first_read_idx)
{
  if (!dev_seq) return LINKS_E_INVAL;
  return reads_push_impl(c, nullptr, static_cast<const uint8_t*>(dev_seq), off, n, first_read_idx);
}

int links_table_mode(links_ctx* c, int mode)
{
  if (!c) return LINKS_E_INVAL;
  if (mode != LINKS_TABLE_AUTO && mode != LINKS_TABLE_FULL && mode != LINKS_TABLE_MATES) return fail(c, LINKS_E_INVAL, "links_table_mode: mode %d", mode);
  if (c->table_started || c->table_mates) return fail(c, LINKS_E_INVAL, "links_table_mode: call before the job's first links_reads_push");
  c->table_mode = mode;
  return LINKS_OK;
}

int links_set_stream(links_ctx* c, void* cuda_stream)
{
  if (!c) return LINKS_E_INVAL;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->s_compute));
  if (c->own_compute_stream && c->s_compute) cudaStreamDestroy(c->s_compute);
  c->s_compute = static_cast<cudaStream_t>(cuda_stream);
  c->own_compute_stream = false;
  return LINKS_OK;
}

// Forget the job (Bloom bits, contigs, pair set, counters) but keep every device allocation.
int links_reset(links_ctx* c)
{
  if (!c) return LINKS_E_INVAL;
  NvtxRange nvtx_("links_reset");
  CU(c, cudaSetDevice(c->device));
  host_mark(c, "reset: enter");
  CU(c, cudaStreamSynchronize(c->s_copy));
  CU(c, cudaStreamSynchronize(c->s_aux));
  CU(c, cudaStreamSynchronize(c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_route));
  CU(c, cudaStreamSynchronize(c->s_cls));
  CU(c, cudaStreamSynchronize(c->s_import));
  host_mark(c, "reset: streams idle");
  c->table_started = false; c->table_batches_built = 0; c->table_mates = false;
  poll_snapshots(c);
  for (auto* b : c->contig_batches) c->contig_pool.push_back(b);
  c->contig_batches.clear();
  c->tig_len.assign(1, 0);
  c->contig_positions = 0; c->n_tigs_seen = 0;
  c->bf_ready = false; c->bf_insert = false; c->bf_covers_contigs = false;
  c->bloom_ext = nullptr; c->bf_share_n = 1; c->bf_share_rank = 0; c->bf_reduced = true;
  c->member_buckets = 0; c->member_positions = 0; c->member_off = false;
  c->cand_total = 0; c->known_unique = 0; c->known_cand = 0;
  for (int i = 0; i < links_ctx::kSnaps; ++i) c->snap_live[i] = false;
  for (int i = 0; i < links_ctx::kSlots; ++i) c->slot_used[i] = false;
  c->next_slot = 0;  // same chunk -> slot assignment every job: slot buffers stop growing after the first one
  CU(c, cudaMemsetAsync(c->d_ctr.p, 0, sizeof(Counters), c->s_compute));
  CU(c, cudaMemsetAsync(c->d_cls.p, 0, 64, c->s_compute));  // bounds[0] = 0; later bounds are written before they are read
  c->cls_round = 0;
  c->cls_valid = c->cls_on; c->cls_started = false; c->max_read_end = 0;
  if (c->x_attached && c->x_settled) {  // the previous job swapped the owned set in: undo, so each buffer keeps its role (and size)
    std::swap(c->d_pairs, c->d_owned); std::swap(c->d_pair_list, c->d_owned_list);
    std::swap(c->pair_cap, c->owned_cap);
  }
  if (c->pair_cap) {
    // off the compute stream: the Bloom build and the first copies do not have to wait for it
    Span sp(c, F_GROW, c->s_aux);
    CU(c, cudaMemsetAsync(c->d_pairs.p, 0xFF, c->pair_cap * pair_slot_bytes(c), c->s_aux));
  }
  if (c->x_attached) {
    c->x_epoch++;
    c->x_finished = false; c->x_settled = false; c->x_job_started = false;
    CU(c, cudaMemsetAsync(reinterpret_cast<char*>(c->d_xdev.p) + offsetof(ExchangeDev, cursor), 0,
                          sizeof(ExchangeDev) - offsetof(ExchangeDev, cursor), c->s_compute));
    if (c->owned_cap) {
      Span sp(c, F_GROW, c->s_aux);
      CU(c, cudaMemsetAsync(c->d_owned.p, 0xFF, c->owned_cap * pair_slot_bytes(c), c->s_aux));
    }
  }
  CU(c, cudaEventRecord(c->ev_cleared, c->s_aux));
  c->contig_pack_pending = false;
  c->finalized = false;
  host_mark(c, "reset: return");
  return LINKS_OK;
}

int links_get_timing(links_ctx* c, links_timing* t, int reset)
{
  if (!c || !t) return LINKS_E_INVAL;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->s_copy));
  CU(c, cudaStreamSynchronize(c->s_aux));
  CU(c, cudaStreamSynchronize(c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_route));
  CU(c, cudaStreamSynchronize(c->s_cls));
  CU(c, cudaStreamSynchronize(c->s_import));
  resolve_spans(c);
  t->ms_h2d = c->fam_ms[F_H2D]; t->ms_pack = c->fam_ms[F_PACK]; t->ms_bf_build = c->fam_ms[F_BF];
  t->ms_table_build = c->fam_ms[F_TABLE]; t->ms_extract = c->fam_ms[F_EXTRACT]; t->ms_classify = c->fam_ms[F_CLASSIFY];
  t->ms_sort_reduce = c->fam_ms[F_SORT]; t->ms_d2h = c->fam_ms[F_D2H]; t->ms_grow = c->fam_ms[F_GROW];
  t->ms_exchange = c->fam_ms[F_EXCHANGE];
  t->launches = c->launches; t->extract_launches = c->extract_launches;
  if (reset) { for (auto& m : c->fam_ms) m = 0; c->launches = 0; c->extract_launches = 0; }
  return LINKS_OK;
}

static uint64_t full_table_cap(const links_ctx* c);

static int build_table(links_ctx* c, cudaStream_t st)
{
  const uint32_t k = c->prm.k;
  const uint64_t cap = full_table_cap(c);
  CU(c, c->d_table.ensure((cap + 1) * sizeof(TableSlot)));
  c->table_cap = cap;
  c->table_has_occ = false;
  Span sp(c, F_TABLE, st);
  CU(c, cudaMemsetAsync(reinterpret_cast<char*>(c->d_ctr.p) + offsetof(Counters, n_contig_kmers), 0, sizeof(unsigned long long), st));
  // keys = EMPTY (all ones), vals = 0
  table_init_kernel<<<grid_for(c, cap + 1, 256 * 4, 16), 256, 0, st>>>(c->d_table.as<TableSlot>(), cap + 1);
  c->launches++;
  for (auto* b : c->contig_batches) {
    if (!b->n_tiles) continue;
    ContigBatch cb{b->packed.as<uint32_t>(), b->inv.as<uint32_t>(), b->off.as<uint64_t>(), b->base0,
                   b->segs.as<uint32_t>(), b->n_seq, b->n_tiles, b->first_id};
    contig_kernel<1><<<grid_for(c, b->n_tiles, 1, 8), kCtaThreads, k * 4 * sizeof(ulonglong2), st>>>(
      cb, int(k), c->kseed, c->d_tab.as<ulonglong2>(), bloom_view(c), table_view(c), MemberView{nullptr, 0}, c->d_ctr.as<Counters>(),
      BloomPass{0, ~0ull, 1});
    c->launches++;
  }
  CU(c, cudaGetLastError());
  c->table_batches_built = c->contig_batches.size();
  return LINKS_OK;
}

// The contig k-mer table needs only the contigs, which are resident before the first read
// arrives (call order of the reference: Bloom filter -> reads -> contigs; here the contigs come
// with the filter).  Start it on the auxiliary stream at the first reads push so it runs under the
// reads' H2D copies and extract launches instead of in front of stage 4.
// The table restricted to `mates` (kernels.cuh table_seed / table_update): slots for the hashes of the n_pairs listed
// pairs, then one pass over the resident contigs that fills in those slots only.
static int build_table_mates(links_ctx* c, cudaStream_t st, uint64_t n_pairs)
{
  const uint32_t k = c->prm.k;
  // <= 2 hashes per pair; load <= 0.4 (a miss that gets past the occupancy bitmap should end in its first sector), but never
  // more than a quarter of the device for the table
  uint64_t cap = 1ull << log2_ceil(std::max<uint64_t>(n_pairs * 2 * 10 / 4, 1ull << 10));
  while (cap > (1ull << 10) && (cap + 1) * sizeof(TableSlot) > c->dev_total_mem / 4 && (cap >> 1) * 6 / 10 >= n_pairs * 2) cap >>= 1;
  CU(c, c->d_table.ensure((cap + 1) * sizeof(TableSlot)));
  CU(c, c->d_table_occ.ensure(cap / 16 + 64));
  c->table_cap = cap;
  c->table_has_occ = true;
  Span sp(c, F_TABLE, st);
  CU(c, cudaMemsetAsync(c->d_table_occ.p, 0, cap / 16 + 64, st));
  CU(c, cudaMemsetAsync(reinterpret_cast<char*>(c->d_ctr.p) + offsetof(Counters, n_contig_kmers), 0, sizeof(unsigned long long), st));
  table_init_kernel<<<grid_for(c, cap + 1, 256 * 4, 16), 256, 0, st>>>(c->d_table.as<TableSlot>(), cap + 1);
  c->launches++;
  if (n_pairs) {
    table_seed_kernel<<<grid_for(c, n_pairs, kCtaThreads, 8), kCtaThreads, 0, st>>>(c->d_pair_list.as<ulonglong2>(), n_pairs, table_view(c));
    c->launches++;
  }
  // (Pinning the occupancy bitmap in L2 with an access-policy window for the duration of the scan was tried and is slower:
  // the carve-out is taken from the L2 every other kernel uses, profiles/r02w_table_l2_persist_ab.txt.)
  for (auto* b : c->contig_batches) {
    if (!b->n_tiles) continue;
    ContigBatch cb{b->packed.as<uint32_t>(), b->inv.as<uint32_t>(), b->off.as<uint64_t>(), b->base0,
                   b->segs.as<uint32_t>(), b->n_seq, b->n_tiles, b->first_id};
    contig_kernel<3><<<grid_for(c, b->n_tiles, 1, 8), kCtaThreads, k * 4 * sizeof(ulonglong2), st>>>(
      cb, int(k), c->kseed, c->d_tab.as<ulonglong2>(), bloom_view(c), table_view(c), MemberView{nullptr, 0}, c->d_ctr.as<Counters>(),
      BloomPass{0, ~0ull, 1});
    c->launches++;
  }
  CU(c, cudaGetLastError());
  return LINKS_OK;
}

static uint64_t full_table_cap(const links_ctx* c)
{
  return 1ull << log2_ceil(std::max<uint64_t>(c->contig_positions * 4, 1ull << 10));  // 2 entries per position, load <= 0.5
}

// FULL costs 64 bytes per contig position whatever the reads are and is built while they stream in; MATES costs at most
// ~50 bytes per distinct pair and is built after the last read.  AUTO keeps FULL while it is a small part of the device.
static bool want_mates_table(const links_ctx* c)
{
  if (c->table_mode == LINKS_TABLE_FULL) return false;
  if (c->table_mode == LINKS_TABLE_MATES) return true;
  return (full_table_cap(c) + 1) * sizeof(TableSlot) > c->dev_total_mem / 8;
}

static int start_table_build(links_ctx* c)
{
  if (c->table_started || c->table_mates || c->contig_batches.empty()) return LINKS_OK;
  if (want_mates_table(c)) { c->table_mates = true; return LINKS_OK; }
  CU(c, cudaEventRecord(c->ev_fork, c->s_compute));  // contig packs (and the counter reset) are ordered before
  CU(c, cudaStreamWaitEvent(c->s_aux, c->ev_fork, 0));
  int rc = build_table(c, c->s_aux);
  if (rc) return rc;
  CU(c, cudaEventRecord(c->ev_table, c->s_aux));
  c->table_started = true;
  return LINKS_OK;
}

int links_finalize(links_ctx* c, links_row** rows_out, uint64_t* n_rows_out, links_issue** issues_out,
                   uint64_t* n_issues_out, links_stats* stats)
{
  if (!c || !rows_out || !n_rows_out) return LINKS_E_INVAL;
  *rows_out = nullptr; *n_rows_out = 0;
  if (issues_out) *issues_out = nullptr;
  if (n_issues_out) *n_issues_out = 0;
  NvtxRange nvtx_("links_finalize");
  CU(c, cudaSetDevice(c->device));
  host_mark(c, "finalize: enter");
  cudaStream_t s = c->s_compute;
  CU(c, cudaStreamSynchronize(c->s_copy));
  CU(c, cudaStreamWaitEvent(s, c->ev_cleared, 0));
  int rc = LINKS_OK;
  if (!c->table_started && !c->table_mates && want_mates_table(c)) c->table_mates = true;  // no read was pushed
  if (c->table_mates) {
    // built below, once the (owned) pairs are all here
  } else if (c->table_started && c->table_batches_built == c->contig_batches.size()) {
    CU(c, cudaStreamWaitEvent(s, c->ev_table, 0));
  } else {  // no read was pushed, or contigs arrived after the early build: (re)build here
    if (c->table_started) CU(c, cudaStreamSynchronize(c->s_aux));
    rc = build_table(c, s);
    if (rc) return rc;
  }
  if (c->x_attached) {  // multi-GPU: stage 4 runs over the pairs this rank OWNS
    rc = exchange_settle(c);
    if (rc) return rc;
  }
  if (c->cls_started) {  // incremental stage 4 in flight on s_cls
    CU(c, cudaEventRecord(c->ev_cls_done, c->s_cls));
    CU(c, cudaStreamWaitEvent(s, c->ev_cls_done, 0));
  }
  host_mark(c, "finalize: settled");
  // unique pair count sizes the row buffers
  CU(c, cudaStreamSynchronize(s));
  host_mark(c, "finalize: extraction drained");
  Counters h;
  CU(c, cudaMemcpy(&h, c->d_ctr.p, sizeof h, cudaMemcpyDeviceToHost));
  const uint64_t n_unique = h.n_unique;
  if (n_unique >= (1ull << 32) - kSortTile) return fail(c, LINKS_E_LIMIT, "links_finalize: more than 2^32 unique pairs on one GPU");
  if (c->table_mates) {
    rc = build_table_mates(c, s, c->pair_cap ? n_unique : 0);
    if (rc) return rc;
  }

  const int tb = log2_ceil(uint64_t(c->n_tigs_seen) * 2 + 2);
  DevBuf (&d_keys)[2] = c->f_keys; DevBuf (&d_vals)[2] = c->f_vals;
  DevBuf &d_issues = c->f_issues, &d_tiglen = c->f_tiglen, &d_scan_tmp = c->f_scan_tmp, &d_flags = c->f_flags,
         &d_total = c->f_total, &d_okeys = c->f_okeys, &d_olinks = c->f_olinks, &d_ogaps = c->f_ogaps;
  const bool incremental = c->cls_valid && c->cls_started && c->pair_cap == c->cls_pair_cap;
  const uint64_t row_cap = incremental ? c->cls_row_cap : std::max<uint64_t>(n_unique, 1);
  uint64_t issue_cap = incremental ? c->cls_issue_cap : c->issue_cap_hint;
  CU(c, d_keys[0].ensure(row_cap * 8)); CU(c, d_vals[0].ensure(row_cap * 4));
  CU(c, d_tiglen.ensure(c->tig_len.size() * 4));
  CU(c, cudaMemcpyAsync(d_tiglen.p, c->tig_len.data(), c->tig_len.size() * 4, cudaMemcpyHostToDevice, s));
  CU(c, d_total.ensure(64));

  for (int attempt = 0; attempt < 2; ++attempt) {
    CU(c, d_issues.ensure(issue_cap * sizeof(IssueRec)));
    const bool rest_only = incremental && attempt == 0;  // pairs up to the last extract launch but one are already classified
    unsigned long long* range = c->d_cls.as<unsigned long long>() + c->cls_round;  // [end of the last round, all pairs)
    if (rest_only) {
      CU(c, cudaMemcpyAsync(range + 1, &c->d_ctr.as<Counters>()->n_unique, 8, cudaMemcpyDeviceToDevice, s));
    } else {
      // reset the stage-4 counters (keeps the stage-2/3 ones)
      CU(c, cudaMemsetAsync(reinterpret_cast<char*>(c->d_ctr.p) + offsetof(Counters, ct_ok_pairs), 0,
                            sizeof(Counters) - offsetof(Counters, ct_ok_pairs), s));
    }
    if (c->pair_cap && n_unique) {
      Span sp(c, F_CLASSIFY, s);
      ClassifyParams cp{c->d_dist.as<uint32_t>(), c->d_minallowed.as<int32_t>(), c->prm.n_dist, d_tiglen.as<uint32_t>(), tb,
                        d_keys[0].as<uint64_t>(), d_vals[0].as<int32_t>(), row_cap, d_issues.as<IssueRec>(), issue_cap};
      classify_kernel<<<rest_only ? c->sm_count * 4 : grid_for(c, n_unique, kCtaThreads, 8), kCtaThreads, 0, s>>>(
        pair_view(c), n_unique, table_view(c), cp,
        int(c->prm.k), c->d_ctr.as<Counters>(), rest_only ? range : nullptr);
      c->launches++;
      CU(c, cudaGetLastError());
    }
    CU(c, cudaStreamSynchronize(s));
    CU(c, cudaMemcpy(&h, c->d_ctr.p, sizeof h, cudaMemcpyDeviceToHost));
    if (h.n_issues <= issue_cap) break;
    issue_cap = h.n_issues; // rerun with an exact-size buffer
    c->issue_cap_hint = issue_cap;
  }
  host_mark(c, "finalize: classified");
  const uint64_t n_rows_raw = h.n_rows;
  if (n_rows_raw > row_cap) return fail(c, LINKS_E_CUDA, "links_finalize: internal row overflow");

  // ---- K5b: sort + reduce-by-key on the packed row key, then mirror expansion + a final small sort
  FinalRow* hrows = nullptr;
  uint32_t n_seg = 0, n_out32 = 0;
  const int key_bits = 3 + 2 * tb;
  if (n_rows_raw) {
    Span sp(c, F_SORT, s);
    int cur = 0;
    rc = radix_sort_pairs(c, s, d_keys, d_vals, n_rows_raw, key_bits, cur);
    if (rc) return rc;
    CU(c, d_flags.ensure(n_rows_raw * 4));
    head_flags_kernel<<<unsigned((n_rows_raw + 255) / 256), 256, 0, s>>>(d_keys[cur].as<uint64_t>(), n_rows_raw, d_flags.as<uint32_t>());
    c->launches++;
    rc = device_scan(c, s, d_flags.as<uint32_t>(), n_rows_raw, d_flags.as<uint32_t>(), d_scan_tmp, d_total.as<uint32_t>());
    if (rc) return rc;
    CU(c, cudaMemcpyAsync(&n_seg, d_total.p, 4, cudaMemcpyDeviceToHost, s));
    CU(c, cudaStreamSynchronize(s));
    CU(c, d_okeys.ensure(uint64_t(n_seg) * 8)); CU(c, d_olinks.ensure(uint64_t(n_seg) * 8)); CU(c, d_ogaps.ensure(uint64_t(n_seg) * 8));
    CU(c, cudaMemsetAsync(d_olinks.p, 0, uint64_t(n_seg) * 8, s));
    CU(c, cudaMemsetAsync(d_ogaps.p, 0, uint64_t(n_seg) * 8, s));
    const uint64_t thr = (n_rows_raw + kReduceItems - 1) / kReduceItems;
    reduce_by_key_kernel<<<unsigned((thr + 255) / 256), 256, 0, s>>>(d_keys[cur].as<uint64_t>(), d_vals[cur].as<int32_t>(), n_rows_raw,
                                                                   d_flags.as<uint32_t>(), d_okeys.as<uint64_t>(),
                                                                   d_olinks.as<unsigned long long>(), d_ogaps.as<unsigned long long>());
    c->launches++;
  }
  if (c->x_attached && c->x_lay.n_ranks > 1) {  // partial link tables of all ranks are summed on rank 0, on the device
    rc = exchange_rows(c, s, n_seg, key_bits, &n_seg);
    if (rc) return rc;
  }
  if (n_seg) {
    Span sp(c, F_SORT, s);
    int cur = 0;
    // every canonical row stands for two tsv rows (add_to_pair_map, reference :717-761): expand, sort, gather
    n_out32 = n_seg * 2;
    CU(c, d_keys[0].ensure(uint64_t(n_out32) * 8)); CU(c, d_vals[0].ensure(uint64_t(n_out32) * 4));
    mirror_expand_kernel<<<(n_seg + 255) / 256, 256, 0, s>>>(d_okeys.as<uint64_t>(), n_seg, tb, d_keys[0].as<uint64_t>(), d_vals[0].as<int32_t>());
    c->launches++;
    cur = 0;
    rc = radix_sort_pairs(c, s, d_keys, d_vals, n_out32, key_bits, cur);
    if (rc) return rc;
    CU(c, c->f_final.ensure(uint64_t(n_out32) * sizeof(FinalRow)));
    gather_rows_kernel<<<(n_out32 + 255) / 256, 256, 0, s>>>(d_keys[cur].as<uint64_t>(), d_vals[cur].as<int32_t>(), n_out32, tb,
                                                           d_olinks.as<unsigned long long>(), d_ogaps.as<unsigned long long>(),
                                                           c->f_final.as<FinalRow>());
    c->launches++;
    CU(c, cudaGetLastError());
    if (c->h_rows_cap < n_out32) {
      if (c->h_rows) cudaFreeHost(c->h_rows);
      c->h_rows = nullptr; c->h_rows_cap = 0;
      const size_t want = size_t(n_out32) + n_out32 / 4 + 1024;
      CU(c, cudaHostAlloc(reinterpret_cast<void**>(&c->h_rows), want * sizeof(FinalRow), cudaHostAllocDefault));
      c->h_rows_cap = want;
    }
    hrows = c->h_rows;
  }
  std::vector<IssueRec> hi(h.n_issues);
  {
    Span sp(c, F_D2H, s);
    if (n_out32) CU(c, cudaMemcpyAsync(hrows, c->f_final.p, uint64_t(n_out32) * sizeof(FinalRow), cudaMemcpyDeviceToHost, s));
    if (h.n_issues) CU(c, cudaMemcpyAsync(hi.data(), d_issues.p, h.n_issues * sizeof(IssueRec), cudaMemcpyDeviceToHost, s));
  }
  CU(c, cudaStreamSynchronize(s));
  host_mark(c, "finalize: rows on host");

  // ---- host: the rows arrive as links_row records in (bin, tig_a, rev_a, tig_b, rev_b) order (the packed key is monotone in it)
  static_assert(sizeof(FinalRow) == sizeof(links_row), "row record layout");
  const uint64_t n_out = n_out32;
  links_row* rows = static_cast<links_row*>(result_alloc(std::max<uint64_t>(n_out, 1) * sizeof(links_row)));
  if (!rows) return fail(c, LINKS_E_NOMEM, "links_finalize: out of host memory");
  if (n_out) memcpy(rows, hrows, n_out * sizeof(links_row));
  host_mark(c, "finalize: rows unpacked");
  links_issue* issues = nullptr;
  if (issues_out) {
    issues = static_cast<links_issue*>(result_alloc(std::max<uint64_t>(hi.size(), 1) * sizeof(links_issue)));
    if (!issues) { result_free(rows); return fail(c, LINKS_E_NOMEM, "links_finalize: out of host memory"); }
    static_assert(sizeof(links_issue) == sizeof(IssueRec), "issue record layout");
    if (!hi.empty()) memcpy(issues, hi.data(), hi.size() * sizeof(links_issue));
    std::sort(issues, issues + hi.size(), [](const links_issue& a, const links_issue& b) {
#define LK(f) if (a.f != b.f) return a.f < b.f
      LK(tig); LK(s1); LK(s2); LK(e1); LK(e2); LK(o1); LK(o2); LK(pet_size); LK(insert_size); LK(tig_len);
#undef LK
      return false;
    });
    *issues_out = issues;
    if (n_issues_out) *n_issues_out = hi.size();
  }
  *rows_out = rows; *n_rows_out = n_out;
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->n_candidates = h.n_candidates; stats->n_survivors = h.n_survivors; stats->n_unique_pairs = h.n_unique;
    stats->n_bf_probes = h.n_bf_probes; stats->n_contig_kmers = h.n_contig_kmers;
    stats->n_bf_loads = h.n_bf_loads; stats->n_member_sectors = h.n_member_sectors;
    stats->ct_ok_pairs = h.ct_ok_pairs; stats->ct_problem_pairs = h.ct_problem_pairs; stats->ct_ok_contig = h.ct_ok_contig;
    stats->ct_iz_issues = h.ct_iz_issues; stats->ct_illogical = h.ct_illogical; stats->ct_gate_fail = h.ct_gate_fail;
    stats->n_rows = n_out;
  }
  c->finalized = true;
  host_mark(c, "finalize: return");
  return LINKS_OK;
}

void links_free_result(links_row* rows, links_issue* issues) { result_free(rows); result_free(issues); }

int links_format_row(const links_row* r, char* dst, size_t cap)
{
  char a[16] = "", b[16] = "";
  if (r->tig_a) snprintf(a, sizeof a, "%u", r->tig_a);
  if (r->tig_b) snprintf(b, sizeof b, "%u", r->tig_b);
  return snprintf(dst, cap, "%d\t%c%s\t%c%s\t%llu\t%lld\n", r->bin, r->rev_a ? 'r' : 'f', a, r->rev_b ? 'r' : 'f', b,
                  (unsigned long long)r->links, (long long)r->gaps);
}

int links_format_issue(const links_issue* r, char* dst, size_t cap)
{
  char t[16] = "";
  if (r->tig) snprintf(t, sizeof t, "%u", r->tig);
  return snprintf(dst, cap,
                  "Pairs unsatisfied in distance within a contig.  Pairs on contig id: %s (length: %llu)\n"
                  "kmer start: %llu kmer end: %llu orientation: %d\n"
                  "kmer start: %llu kmer end: %llu orientation: %d\n"
                  "Calculated distance apart on contig: %llu - Distance on read: %lld\n",
                  t, (unsigned long long)r->tig_len, (unsigned long long)r->s1, (unsigned long long)r->e1, int(r->o1),
                  (unsigned long long)r->s2, (unsigned long long)r->e2, int(r->o2), (unsigned long long)r->pet_size,
                  (long long)r->insert_size);
}

// ---------------------------------------------------------------- inspection utilities

int links_kmer_hashes(links_ctx* c, const char* seq, uint64_t len, uint64_t* fwd, uint64_t* rev, uint8_t* valid)
{
  if (!c || !seq || !fwd || !rev || !valid) return LINKS_E_INVAL;
  const uint32_t k = c->prm.k;
  if (len < k) return LINKS_OK;
  CU(c, cudaSetDevice(c->device));
  const uint64_t npos = len - k + 1;
  SeqBatchDev b; DevBuf ascii, d_f, d_r, d_v;
  struct Cleanup { SeqBatchDev& b; DevBuf &a, &f, &r, &v; ~Cleanup() { b.release(); a.release(); f.release(); r.release(); v.release(); } } cl{b, ascii, d_f, d_r, d_v};
  const uint64_t off[2] = {0, len};
  std::vector<uint32_t> segs = {0, 0};
  CU(c, cudaStreamSynchronize(c->s_compute));
  int rc = upload_and_pack(c, b, ascii, seq, nullptr, off, 1, segs, false, c->s_copy, c->s_compute, c->ev_contig);
  if (rc) return rc;
  CU(c, d_f.ensure(npos * 8)); CU(c, d_r.ensure(npos * 8)); CU(c, d_v.ensure(npos));
  kmer_hashes_kernel<<<grid_for(c, npos, 256, 16), 256, k * 4 * sizeof(ulonglong2), c->s_compute>>>(
    b.packed.as<uint32_t>(), b.inv.as<uint32_t>(), npos, int(k), c->d_tab.as<ulonglong2>(), d_f.as<uint64_t>(), d_r.as<uint64_t>(),
    d_v.as<uint8_t>());
  c->launches++;
  CU(c, cudaGetLastError());
  CU(c, cudaMemcpyAsync(fwd, d_f.p, npos * 8, cudaMemcpyDeviceToHost, c->s_compute));
  CU(c, cudaMemcpyAsync(rev, d_r.p, npos * 8, cudaMemcpyDeviceToHost, c->s_compute));
  CU(c, cudaMemcpyAsync(valid, d_v.p, npos, cudaMemcpyDeviceToHost, c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_compute));
  return LINKS_OK;
}

// links_kmer_hashes without a single device allocation (see g_static_scratch above): same pack_kernel + kmer_hashes_kernel,
// same launch shapes, on the fixed array.  The array keeps the previous call's contents, so what the kernels expect to find
// zeroed / preset is set here on every call (the record flags pack_kernel ORs into, the padding words behind the packed
// bases and behind the mask).  Calls through this entry point are serialised per process; work that another process or
// another entry point queues on the same GPU does not touch the array.
int links_kmer_hashes_static(links_ctx* c, const char* seq, uint64_t len, uint64_t* fwd, uint64_t* rev, uint8_t* valid)
{
  if (!c || !seq || !fwd || !rev || !valid) return LINKS_E_INVAL;
  if (len > kStaticMaxBases)
    return fail(c, LINKS_E_LIMIT, "links_kmer_hashes_static: %llu bases, the static scratch holds at most %llu (LINKS_STATIC_MAX_BASES); use links_kmer_hashes",
                (unsigned long long)len, (unsigned long long)kStaticMaxBases);
  const uint32_t k = c->prm.k;
  if (len < k) return LINKS_OK;
  CU(c, cudaSetDevice(c->device));
  std::lock_guard<std::mutex> lk(g_static_mu);
  unsigned char* base = nullptr;
  int rc = static_scratch_base(c, &base);
  if (rc) return rc;
  uint8_t* d_ascii = base + kSsAscii;
  uint32_t* d_packed = reinterpret_cast<uint32_t*>(base + kSsPacked);
  uint32_t* d_inv = reinterpret_cast<uint32_t*>(base + kSsInv);
  uint64_t* d_off = reinterpret_cast<uint64_t*>(base + kSsOff);
  uint32_t* d_flags = reinterpret_cast<uint32_t*>(base + kSsFlags);
  uint64_t* d_f = reinterpret_cast<uint64_t*>(base + kSsFwd);
  uint64_t* d_r = reinterpret_cast<uint64_t*>(base + kSsRev);
  uint8_t* d_v = base + kSsValid;
  const uint64_t npos = len - k + 1;
  const uint64_t off[2] = {0, len};
  cudaStream_t s = c->s_compute;
  CU(c, cudaStreamSynchronize(s));
  CU(c, cudaMemcpyAsync(d_ascii, seq, len, cudaMemcpyHostToDevice, s));
  CU(c, cudaMemcpyAsync(d_off, off, sizeof off, cudaMemcpyHostToDevice, s));
  CU(c, cudaMemsetAsync(d_flags, 0, 2 * 4, s));
  CU(c, cudaMemsetAsync(d_packed + (len + 15) / 16, 0, 8 * 4, s));
  CU(c, cudaMemsetAsync(d_inv + (len + 31) / 32, 0xFF, 8 * 4, s));
  const uint64_t threads = (len + 31) / 32;
  pack_kernel<<<unsigned((threads + kCtaThreads - 1) / kCtaThreads), kCtaThreads, 0, s>>>(d_ascii, len, d_packed, d_inv, d_off, 1, 0, d_flags);
  c->launches++;
  kmer_hashes_kernel<<<grid_for(c, npos, 256, 16), 256, k * 4 * sizeof(ulonglong2), s>>>(d_packed, d_inv, npos, int(k), c->d_tab.as<ulonglong2>(),
                                                                                        d_f, d_r, d_v);
  c->launches++;
  CU(c, cudaGetLastError());
  CU(c, cudaMemcpyAsync(fwd, d_f, npos * 8, cudaMemcpyDeviceToHost, s));
  CU(c, cudaMemcpyAsync(rev, d_r, npos * 8, cudaMemcpyDeviceToHost, s));
  CU(c, cudaMemcpyAsync(valid, d_v, npos, cudaMemcpyDeviceToHost, s));
  CU(c, cudaStreamSynchronize(s));
  return LINKS_OK;
}

int links_bf_contains(links_ctx* c, const uint64_t* fwd, const uint64_t* rev, uint64_t n, uint8_t* out)
{
  if (!c || !fwd || !rev || !out) return LINKS_E_INVAL;
  if (!c->bf_ready) return fail(c, LINKS_E_INVAL, "links_bf_contains: no Bloom filter");
  if (!n) return LINKS_OK;
  CU(c, cudaSetDevice(c->device));
  DevBuf d_f, d_r, d_o;
  struct Cleanup { DevBuf &f, &r, &o; ~Cleanup() { f.release(); r.release(); o.release(); } } cl{d_f, d_r, d_o};
  CU(c, d_f.ensure(n * 8)); CU(c, d_r.ensure(n * 8)); CU(c, d_o.ensure(n));
  cudaStream_t s = c->s_compute;
  CU(c, cudaMemcpyAsync(d_f.p, fwd, n * 8, cudaMemcpyHostToDevice, s));
  CU(c, cudaMemcpyAsync(d_r.p, rev, n * 8, cudaMemcpyHostToDevice, s));
  bf_contains_kernel<<<unsigned((n + 255) / 256), 256, 0, s>>>(bloom_view(c), c->kseed, d_f.as<uint64_t>(), d_r.as<uint64_t>(), n, d_o.as<uint8_t>());
  c->launches++;
  CU(c, cudaGetLastError());
  CU(c, cudaMemcpyAsync(out, d_o.p, n, cudaMemcpyDeviceToHost, s));
  CU(c, cudaStreamSynchronize(s));
  return LINKS_OK;
}

int links_pairs_download(links_ctx* c, uint64_t* h_a, uint64_t* h_b, uint32_t* d_idx, uint64_t cap, uint64_t* n)
{
  if (!c || !n) return LINKS_E_INVAL;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->s_compute));
  Counters h;
  CU(c, cudaMemcpy(&h, c->d_ctr.p, sizeof h, cudaMemcpyDeviceToHost));
  *n = h.n_unique;
  if (!h_a || !h_b || !d_idx) return LINKS_OK;
  if (cap < h.n_unique) return fail(c, LINKS_E_INVAL, "links_pairs_download: cap %llu < %llu", (unsigned long long)cap, (unsigned long long)h.n_unique);
  if (!h.n_unique) return LINKS_OK;
  DevBuf& d_out = c->f_exch;  // scratch kept by the context (no cudaFree, which would wait for every kernel on the device)
  CU(c, d_out.ensure(std::max<uint64_t>(h.n_unique * 24, size_t(kMaxRanks) * 2 * 8)));
  pair_export_kernel<<<grid_for(c, h.n_unique, 256, 8), 256, 0, c->s_compute>>>(
    pair_view(c), h.n_unique, d_out.as<uint64_t>());
  c->launches++;
  CU(c, cudaGetLastError());
  std::vector<uint64_t> tmp(h.n_unique * 3);
  CU(c, cudaMemcpyAsync(tmp.data(), d_out.p, h.n_unique * 24, cudaMemcpyDeviceToHost, c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_compute));
  for (uint64_t i = 0; i < h.n_unique; ++i) {
    h_a[i] = tmp[3 * i]; h_b[i] = tmp[3 * i + 1];
    d_idx[i] = c->prm.n_dist > 1 ? okey_d_idx(tmp[3 * i + 2]) : 0;
  }
  return LINKS_OK;
}

// ---------------------------------------------------------------- multi-GPU exchange hooks

// Sum rows with equal (bin, tig_a, rev_a, tig_b, rev_b): the per-rank partial link tables of the
// multi-GPU path, gathered on one rank.  Same sort + segmented reduction as stage 4, on the device.
int links_rows_merge(links_ctx* c, const links_row* rows, uint64_t n, links_row** out, uint64_t* n_out)
{
  if (!c || !out || !n_out || (!rows && n)) return LINKS_E_INVAL;
  *out = nullptr; *n_out = 0;
  if (n >= (1ull << 31)) return fail(c, LINKS_E_LIMIT, "links_rows_merge: more than 2^31 rows");
  links_row* res = static_cast<links_row*>(result_alloc(std::max<uint64_t>(n, 1) * sizeof(links_row)));
  if (!res) return fail(c, LINKS_E_NOMEM, "links_rows_merge: out of host memory");
  *out = res;
  if (!n) return LINKS_OK;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t s = c->s_compute;
  uint32_t max_tig = 0;
  for (uint64_t i = 0; i < n; ++i) max_tig = std::max(max_tig, std::max(rows[i].tig_a, rows[i].tig_b));
  const int tb = log2_ceil(uint64_t(max_tig) * 2 + 2);
  std::vector<uint64_t> hk(n), hl(n), hg(n);
  std::vector<int32_t> hidx(n);
  for (uint64_t i = 0; i < n; ++i) {
    const links_row& r = rows[i];
    const uint32_t bc = r.bin < 0 ? 0 : r.bin == 10 ? 1 : r.bin == 500 ? 2 : r.bin == 5000 ? 3 : 4;
    hk[i] = pack_row_key(bc, (r.tig_a << 1) | (r.rev_a & 1u), (r.tig_b << 1) | (r.rev_b & 1u), tb);
    hl[i] = r.links; hg[i] = uint64_t(r.gaps); hidx[i] = int32_t(i);
  }
  DevBuf (&d_keys)[2] = c->f_keys; DevBuf (&d_vals)[2] = c->f_vals;
  DevBuf &d_il = c->f_il, &d_ig = c->f_ig;
  CU(c, d_keys[0].ensure(n * 8)); CU(c, d_vals[0].ensure(n * 4)); CU(c, d_il.ensure(n * 8)); CU(c, d_ig.ensure(n * 8));
  CU(c, c->f_flags.ensure(n * 4)); CU(c, c->f_total.ensure(64));
  uint32_t n_seg = 0;
  int cur = 0;
  {
    Span sp(c, F_SORT, s);
    CU(c, cudaMemcpyAsync(d_keys[0].p, hk.data(), n * 8, cudaMemcpyHostToDevice, s));
    CU(c, cudaMemcpyAsync(d_vals[0].p, hidx.data(), n * 4, cudaMemcpyHostToDevice, s));
    CU(c, cudaMemcpyAsync(d_il.p, hl.data(), n * 8, cudaMemcpyHostToDevice, s));
    CU(c, cudaMemcpyAsync(d_ig.p, hg.data(), n * 8, cudaMemcpyHostToDevice, s));
    int rc = radix_sort_pairs(c, s, d_keys, d_vals, n, 3 + 2 * tb, cur);
    if (rc) return rc;
    head_flags_kernel<<<unsigned((n + 255) / 256), 256, 0, s>>>(d_keys[cur].as<uint64_t>(), n, c->f_flags.as<uint32_t>());
    c->launches++;
    rc = device_scan(c, s, c->f_flags.as<uint32_t>(), n, c->f_flags.as<uint32_t>(), c->f_scan_tmp, c->f_total.as<uint32_t>());
    if (rc) return rc;
    CU(c, cudaMemcpyAsync(&n_seg, c->f_total.p, 4, cudaMemcpyDeviceToHost, s));
    CU(c, cudaStreamSynchronize(s));
    CU(c, c->f_okeys.ensure(uint64_t(n_seg) * 8)); CU(c, c->f_olinks.ensure(uint64_t(n_seg) * 8)); CU(c, c->f_ogaps.ensure(uint64_t(n_seg) * 8));
    CU(c, cudaMemsetAsync(c->f_olinks.p, 0, uint64_t(n_seg) * 8, s));
    CU(c, cudaMemsetAsync(c->f_ogaps.p, 0, uint64_t(n_seg) * 8, s));
    rows_segment_sum_kernel<<<unsigned((n + 255) / 256), 256, 0, s>>>(
      d_keys[cur].as<uint64_t>(), d_vals[cur].as<int32_t>(), n, c->f_flags.as<uint32_t>(), d_il.as<unsigned long long>(),
      d_ig.as<unsigned long long>(), c->f_okeys.as<uint64_t>(), c->f_olinks.as<unsigned long long>(), c->f_ogaps.as<unsigned long long>());
    c->launches++;
    CU(c, cudaGetLastError());
    hk.resize(n_seg); hl.resize(n_seg); hg.resize(n_seg);
    CU(c, cudaMemcpyAsync(hk.data(), c->f_okeys.p, uint64_t(n_seg) * 8, cudaMemcpyDeviceToHost, s));
    CU(c, cudaMemcpyAsync(hl.data(), c->f_olinks.p, uint64_t(n_seg) * 8, cudaMemcpyDeviceToHost, s));
    CU(c, cudaMemcpyAsync(hg.data(), c->f_ogaps.p, uint64_t(n_seg) * 8, cudaMemcpyDeviceToHost, s));
    CU(c, cudaStreamSynchronize(s));
  }
  static const int32_t bin_of_code[5] = {-1, 10, 500, 5000, 10000};
  const uint64_t tmask = (1ull << tb) - 1;
  for (uint32_t i = 0; i < n_seg; ++i) {
    const uint64_t key = hk[i];
    const uint32_t B = uint32_t(key & tmask), A = uint32_t((key >> tb) & tmask), bc = uint32_t(key >> (2 * tb));
    links_row r;
    memset(&r, 0, sizeof r);
    r.bin = bin_of_code[bc]; r.links = hl[i]; r.gaps = int64_t(hg[i]);
    r.tig_a = A >> 1; r.rev_a = A & 1; r.tig_b = B >> 1; r.rev_b = B & 1;
    res[i] = r;
  }
  *n_out = n_seg;
  return LINKS_OK;
}

uint32_t links_pair_owner(uint64_t h_a, uint64_t h_b, uint32_t n_ranks)
{
  uint64_t x = h_a * 0x9E3779B97F4A7C15ULL + h_b; // slot_hash2, host copy
  x ^= x >> 32;
  x *= 0xD6E8FEB86659FD93ULL;
  return dest_rank_of_hash(x, n_ranks ? n_ranks : 1);
}

int links_pairs_export(links_ctx* c, uint32_t n_ranks, void* dev_dst, uint64_t cap_records, uint64_t* counts)
{
  if (!c || !counts) return LINKS_E_INVAL;
  if (n_ranks == 0 || n_ranks > LINKS_MAX_RANKS) return fail(c, LINKS_E_LIMIT, "links_pairs_export: n_ranks must be 1..%u", LINKS_MAX_RANKS);
  CU(c, cudaSetDevice(c->device));
  cudaStream_t s = c->s_compute;
  CU(c, cudaStreamSynchronize(c->s_copy));
  CU(c, cudaStreamSynchronize(s));
  Counters h;
  CU(c, cudaMemcpy(&h, c->d_ctr.p, sizeof h, cudaMemcpyDeviceToHost));
  const uint64_t n = h.n_unique;
  for (uint32_t r = 0; r < n_ranks; ++r) counts[r] = 0;
  if (!n) return LINKS_OK;
  CU(c, c->f_exch.ensure(size_t(kMaxRanks) * 2 * 8));
  unsigned long long* d_counts = c->f_exch.as<unsigned long long>();
  unsigned long long* d_cursors = d_counts + kMaxRanks;
  const PairView pv = pair_view(c);
  {
    Span sp(c, F_EXCHANGE, s);
    CU(c, cudaMemsetAsync(d_counts, 0, size_t(kMaxRanks) * 8, s));
    pair_bucket_count_kernel<<<grid_for(c, n, kCtaThreads * 8, 8), kCtaThreads, 0, s>>>(pv.list, n, n_ranks, d_counts);
    c->launches++;
    CU(c, cudaGetLastError());
  }
  std::vector<unsigned long long> hc(n_ranks), start(n_ranks);
  CU(c, cudaMemcpyAsync(hc.data(), d_counts, size_t(n_ranks) * 8, cudaMemcpyDeviceToHost, s));
  CU(c, cudaStreamSynchronize(s));
  unsigned long long run = 0;
  for (uint32_t r = 0; r < n_ranks; ++r) { counts[r] = hc[r]; start[r] = run; run += hc[r]; }
  if (!dev_dst) return LINKS_OK;
  if (cap_records < n) return fail(c, LINKS_E_INVAL, "links_pairs_export: room for %llu records, %llu needed", (unsigned long long)cap_records, (unsigned long long)n);
  {
    Span sp(c, F_EXCHANGE, s);
    CU(c, cudaMemcpyAsync(d_cursors, start.data(), size_t(n_ranks) * 8, cudaMemcpyHostToDevice, s));
    pair_bucket_scatter_kernel<<<grid_for(c, n, kCtaThreads, 8), kCtaThreads, 0, s>>>(pv, n, n_ranks, c->prm.n_dist > 1 ? 1 : 0, d_cursors,
                                                                                     static_cast<uint64_t*>(dev_dst));
    c->launches++;
    CU(c, cudaGetLastError());
  }
  CU(c, cudaStreamSynchronize(s)); // `start` is pageable host memory; also: records are complete on return
  return LINKS_OK;
}

int links_pairs_clear(links_ctx* c)
{
  if (!c) return LINKS_E_INVAL;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->s_copy));
  CU(c, cudaStreamSynchronize(c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_cls));
  c->cls_valid = false;  // the set is re-filled by imports: stage 4 runs once, at finalize
  poll_snapshots(c);
  for (int i = 0; i < links_ctx::kSnaps; ++i) c->snap_live[i] = false;
  c->known_unique = 0; c->known_cand = c->cand_total;
  Span sp(c, F_EXCHANGE, c->s_compute);
  CU(c, cudaMemsetAsync(reinterpret_cast<char*>(c->d_ctr.p) + offsetof(Counters, n_unique), 0, sizeof(unsigned long long), c->s_compute));
  if (c->pair_cap) CU(c, cudaMemsetAsync(c->d_pairs.p, 0xFF, c->pair_cap * pair_slot_bytes(c), c->s_compute));
  return LINKS_OK;
}

int links_pairs_import(links_ctx* c, const void* dev_records, uint64_t n_records)
{
  if (!c || (!dev_records && n_records)) return LINKS_E_INVAL;
  if (c->finalized) return fail(c, LINKS_E_INVAL, "links_pairs_import: context already finalized (links_reset first)");
  if (!n_records) return LINKS_OK;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamWaitEvent(c->s_compute, c->ev_cleared, 0));
  CU(c, cudaStreamSynchronize(c->s_cls));
  c->cls_valid = false;
  int rc = ensure_pair_capacity(c, n_records, 0);
  if (rc) return rc;
  {
    Span sp(c, F_EXCHANGE, c->s_compute);
    pair_import_kernel<<<grid_for(c, n_records, kCtaThreads, 8), kCtaThreads, 0, c->s_compute>>>(
      static_cast<const uint64_t*>(dev_records), n_records,
      pair_view(c), c->prm.n_dist > 1 ? 1 : 0, c->d_ctr.as<Counters>());
    c->launches++;
    CU(c, cudaGetLastError());
  }
  // the records may be freed by the caller on return; the next capacity check needs the exact occupancy
  CU(c, cudaStreamSynchronize(c->s_compute));
  Counters h;
  CU(c, cudaMemcpy(&h, c->d_ctr.p, sizeof h, cudaMemcpyDeviceToHost));
  c->known_unique = h.n_unique; c->known_cand = c->cand_total;
  return LINKS_OK;
}

int links_exchange_attach(links_ctx* c, uint32_t rank, uint32_t n_ranks, void* const* inbox, uint64_t inbox_bytes)
{
  if (!c || !inbox) return LINKS_E_INVAL;
  if (n_ranks == 0 || n_ranks > uint32_t(kMaxP2PRanks) || rank >= n_ranks)
    return fail(c, LINKS_E_LIMIT, "links_exchange_attach: rank %u of %u (at most %d ranks)", rank, n_ranks, kMaxP2PRanks);
  if (c->x_attached) return fail(c, LINKS_E_INVAL, "links_exchange_attach: already attached");
  if (!c->contig_batches.empty() || c->cand_total) return fail(c, LINKS_E_INVAL, "links_exchange_attach: attach before the first job");
  for (uint32_t r = 0; r < n_ranks; ++r)
    if (!inbox[r] || (reinterpret_cast<uintptr_t>(inbox[r]) & 255)) return fail(c, LINKS_E_INVAL, "links_exchange_attach: inbox[%u] is null or not 256-byte aligned", r);
  CU(c, cudaSetDevice(c->device));
  InboxLayout L{};
  L.n_ranks = n_ranks; L.rank = rank;
  L.rec_words = c->prm.n_dist > 1 ? 3 : 2;
  L.half_bytes = (inbox_bytes / 2) & ~uint64_t(4095);
  L.pair_off = (sizeof(InboxHeader) + 4095) & ~uint64_t(4095);
  if (L.half_bytes <= L.pair_off + (1ull << 16)) return fail(c, LINKS_E_INVAL, "links_exchange_attach: inbox of %llu bytes is too small", (unsigned long long)inbox_bytes);
  const uint64_t body = L.half_bytes - L.pair_off;
  const uint64_t row_bytes = (body / 16) & ~uint64_t(255);  // 1/16 of a half for the partial link rows
  L.row_cap = row_bytes / n_ranks / sizeof(RowRec);
  L.pair_cap = ((body - row_bytes) / n_ranks / (L.rec_words * 8)) & ~uint64_t(15);
  L.row_off = L.pair_off + uint64_t(n_ranks) * L.pair_cap * L.rec_words * 8;
  L.row_off = (L.row_off + 255) & ~uint64_t(255);
  if (!L.pair_cap || !L.row_cap) return fail(c, LINKS_E_INVAL, "links_exchange_attach: inbox too small for %u ranks", n_ranks);
  CU(c, c->d_xdev.ensure(sizeof(ExchangeDev)));
  ExchangeDev h;
  memset(&h, 0, sizeof h);
  for (uint32_t r = 0; r < n_ranks; ++r) h.inbox[r] = static_cast<char*>(inbox[r]);
  CU(c, cudaMemcpy(c->d_xdev.p, &h, sizeof h, cudaMemcpyHostToDevice));
  // my own headers start at zero; peers only write them after the caller's barrier that follows the attach
  CU(c, cudaMemset(inbox[rank], 0, sizeof(InboxHeader)));
  CU(c, cudaMemset(static_cast<char*>(inbox[rank]) + L.half_bytes, 0, sizeof(InboxHeader)));
  CU(c, cudaDeviceSynchronize());
  c->x_lay = L;
  c->x_epoch = 1;
  c->x_attached = true; c->x_finished = false; c->x_settled = false; c->x_job_started = false;
  {
    const char* e = getenv("LINKS_EXCHANGE_DEDUP");
    c->x_fused = !(e && strcmp(e, "local") == 0);
  }
  return LINKS_OK;
}

int links_exchange_finish(links_ctx* c)
{
  if (!c) return LINKS_E_INVAL;
  if (!c->x_attached) return fail(c, LINKS_E_INVAL, "links_exchange_finish: no inbox attached");
  if (c->x_finished) return LINKS_OK;
  NvtxRange nvtx_("links_exchange_finish");
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamWaitEvent(c->s_compute, c->ev_cleared, 0));
  int rc = c->x_fused ? exchange_after_routed_extract(c, true) : exchange_route(c, true);
  if (rc) return rc;
  c->x_finished = true;
  return LINKS_OK;
}

int links_bf_download_dev(links_ctx* c, void* dev_dst)
{
  if (!c || !dev_dst) return LINKS_E_INVAL;
  if (!c->bf_ready) return fail(c, LINKS_E_INVAL, "links_bf_download_dev: no Bloom filter");
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaMemcpyAsync(dev_dst, bloom_view(c).words, c->bf_bytes, cudaMemcpyDeviceToDevice, c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_compute));
  return LINKS_OK;
}

int links_bf_load_dev(links_ctx* c, const void* dev_bits, uint64_t bytes, uint32_t hash_num)
{
  if (!c || !dev_bits) return LINKS_E_INVAL;
  if (bytes == 0 || (bytes & 3) || hash_num == 0 || hash_num > LINKS_MAX_HASH_NUM)
    return fail(c, LINKS_E_INVAL, "links_bf_load_dev: bytes must be a positive multiple of 4 and hash_num in 1..%u", LINKS_MAX_HASH_NUM);
  CU(c, cudaSetDevice(c->device));
  CU(c, c->d_bloom.ensure(bytes));
  c->bloom_ext = nullptr; c->bf_share_n = 1; c->bf_share_rank = 0; c->bf_reduced = true;
  CU(c, cudaMemcpyAsync(c->d_bloom.p, dev_bits, bytes, cudaMemcpyDeviceToDevice, c->s_compute));
  CU(c, cudaStreamSynchronize(c->s_compute));
  c->bf_bytes = bytes; c->hash_num = hash_num; c->bf_ready = true; c->bf_insert = false;
  return LINKS_OK;
}

} // extern "C"

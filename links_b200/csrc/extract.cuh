// links_b200/csrc/extract.cuh -- extract_kernel: read k-mer pairs -> Bloom probe -> pair-set insert
// (reference src/LINKS.hpp:515-567 extract_mate_pair + :354-378 write_from_block_to_map).
//
// The reference's per-candidate sequence is
//     hash A -> Bloom(A) -> hash B -> Bloom(B) -> pair key -> first-wins insert
// with an early exit after every step; on ONT-like reads ~40 % of the candidates leave after
// A's first two probes, ~54 % never reach B, ~20 % reach the insert.  Run naively (one thread =
// one candidate to completion) the warps thin out to a few live lanes.
//
// Here every WARP is an independent five-stage pipeline over one tile of candidates:
//     S1  hash A, probes 0-1          -> Q1 {q, h0A}
//     S2  probes 2..H-1 of A          -> Q2 {q}
//     S3  hash B, probes 0-1          -> Q3 {q, fB, rB}
//     S4  probes 2..H-1 of B          -> Q4 {q, fB, rB}
//     S5  re-hash A, pair key, insert -> Q5 {h_a, h_b}   (fresh pairs for the dense list)
// The queues are 64-entry per-warp stacks in shared memory.  A stage only runs when its input
// queue holds >= 32 entries (all lanes busy), deepest ready stage first; S1 refills from the
// tile; at the end of a tile the pipeline drains front to back.  There is no block-wide barrier
// after the table load and no lane idles except in the <= 5 drain rounds per tile.  Within a
// stage all probes of a k-mer are issued together (independent loads in flight).
#pragma once
#include "exchange.cuh"
#include "kernels.cuh"
#include "member_map.cuh"

namespace links {

#ifndef LINKS_EXTRACT_TILE
#define LINKS_EXTRACT_TILE 512
#endif
#ifndef LINKS_PROBE_WITH_SET
#define LINKS_PROBE_WITH_SET 0
#endif
#ifndef LINKS_EXTRACT_MIN_CTAS
#define LINKS_EXTRACT_MIN_CTAS 4
#endif
#ifndef LINKS_SEQ_PROBES
#define LINKS_SEQ_PROBES 0   // 1: one Bloom bit per visit of S2 / S4, survivors re-queued (the reference's early exit: fewest sectors)
#endif
#ifndef LINKS_REST_FIRST
#define LINKS_REST_FIRST 0   // > 0: S2 / S4 test that many of the remaining Bloom bits first and the others only if those are all set
#endif
constexpr int kExtractTile = LINKS_EXTRACT_TILE;  // candidates of one (read, distance) per warp tile
#ifndef LINKS_HASH_SPLIT
#define LINKS_HASH_SPLIT 1
#endif
constexpr int kQueueCap = 64;
constexpr int kWarpsPerCta = kCtaThreads / 32;

struct ReadBatch {
  const uint32_t* packed; const uint32_t* inv; const uint64_t* off; uint64_t base0;
  const uint32_t* seg_tile_start; uint32_t n_seq, n_tiles;
  const uint32_t* flags; const unsigned long long* vs; const uint32_t* vrank;
  uint64_t first_read_idx;
  unsigned int* tile_cursor;  // zeroed before every launch: tiles are handed out dynamically
};

struct DistParams { const uint32_t* dist; const uint32_t* step; uint32_t n_dist; uint64_t offset; };

struct __align__(16) WarpQueues {
  uint64_t q1_h[kQueueCap];
  uint64_t q3_f[kQueueCap], q3_r[kQueueCap];
  uint64_t q4_f[kQueueCap], q4_r[kQueueCap];
  uint64_t q5_a[kQueueCap], q5_b[kQueueCap];
  uint16_t q1_i[kQueueCap], q2_i[kQueueCap], q3_i[kQueueCap], q4_i[kQueueCap];
};

__host__ __device__ inline size_t extract_smem_bytes(uint32_t k)
{
  return size_t((k + 1) / 2) * 16 * sizeof(ulonglong2) + size_t(kWarpsPerCta) * sizeof(WarpQueues);
}

// probes kFirst .. H-1 of a k-mer whose first kFirst bits were set (stages S2 and S4).  All together (one memory round
// trip, H - kFirst sectors), or -- LINKS_REST_FIRST = g -- g of them first and the rest only for the lanes that still
// pass: a k-mer that is not in the filter got here by chance (2^-kFirst) and fails each further bit with probability
// ~1/2, so the second group is rarely needed and its sectors are saved at the price of a second round trip for few lanes.
template <int H, bool SMALL>
__device__ __forceinline__ bool bloom_probe_rest(const BloomView& bf, uint64_t h0, uint64_t kseed, uint32_t h_first, uint32_t hn,
                                                 uint32_t& n_probe, uint32_t& n_load)
{
  constexpr int kFirst = H >= 2 ? 2 : H;
  constexpr int kRest = H >= 2 ? H - 2 : 0;
  if (H == 0) { n_load += hn - h_first; return bloom_probe_loop<SMALL>(bf, h0, kseed, h_first, hn, n_probe); }
  constexpr int g = LINKS_REST_FIRST > 0 && LINKS_REST_FIRST < kRest ? LINKS_REST_FIRST : kRest;
  n_load += g;
  if (!bloom_probe_fixed<g, SMALL>(bf, h0, kseed, kFirst, n_probe)) return false;
  if (g == kRest) return true;
  n_load += kRest - g;
  return bloom_probe_fixed<kRest - g, SMALL>(bf, h0, kseed, kFirst + g, n_probe);
}

// lanes with `pass` get consecutive slots starting at `n`; returns the new count (warp-uniform)
__device__ __forceinline__ uint32_t warp_push_slot(bool pass, uint32_t n, uint32_t& slot)
{
  const uint32_t m = __ballot_sync(0xffffffffu, pass);
  slot = n + __popc(m & ((1u << (threadIdx.x & 31)) - 1u));
  return n + __popc(m);
}

// H = Bloom hash count known at compile time (0: runtime bf.hash_num)
// ROUTED (multi-GPU, exchange.cuh): S5 does not insert into a local pair set; it writes the surviving pair as a
// record straight into its OWNER rank's inbox over NVLink peer memory (one cursor atomic per warp and owner), and
// the kernel's last block publishes the new record counts after system-scope fences.  The transfer is part of
// the extraction, tile by tile; the owner de-duplicates while this rank keeps extracting.
// SMALL: the filter has fewer than 2^31 bits (bf.bits32 set): 32-bit remainder in bloom_mod
// MODE 0: S1 tests two Bloom bits, S2 the rest together (fewest round trips: right while the filter and the sets sit in L2)
// MODE 1 (MAP): the launch comes with a member map (member_map.cuh); reads without an invalid base take their answers from it
// MODE 2 (SEQ): one Bloom bit per visit of S2 / S4, survivors re-queued -- the reference's early exit, fewest sectors:
//               right when the filter lives in DRAM, where random sectors are the scarce resource (gather_peak.cu)
template <int H, bool ROUTED, bool SMALL, int MODE>
__global__ void __launch_bounds__(kCtaThreads, LINKS_EXTRACT_MIN_CTAS) extract_kernel(ReadBatch rb, DistParams dp, int k, uint64_t kseed,
                                                                 const ulonglong2* __restrict__ g_tab2, BloomView bf,
                                                                 MemberView mv, PairView pv, Counters* __restrict__ ctr,
                                                                 ExchangeDev* __restrict__ x, InboxLayout L, uint64_t epoch, MemberMap mm)
{
  extern __shared__ ulonglong2 s_dyn[];
  const int n_tab = ((k + 1) >> 1) * 16;
  WarpQueues& Q = reinterpret_cast<WarpQueues*>(s_dyn + n_tab)[threadIdx.x >> 5];
#if LINKS_HASH_SPLIT
  uint64_t* s_tf = reinterpret_cast<uint64_t*>(s_dyn);  // forward halves, then reverse halves (hash_kmer2_split)
  uint64_t* s_tr = s_tf + n_tab;
  for (int i = threadIdx.x; i < n_tab; i += blockDim.x) { const ulonglong2 e = g_tab2[i]; s_tf[i] = e.x; s_tr[i] = e.y; }
#define LINKS_HASH(pos, f, r) hash_kmer2_split(rb.packed, pos, k, s_tf, s_tr, f, r)
#else
  ulonglong2* s_tab = s_dyn;
  for (int i = threadIdx.x; i < n_tab; i += blockDim.x) s_tab[i] = g_tab2[i];
#define LINKS_HASH(pos, f, r) hash_kmer2(rb.packed, pos, k, s_tab, f, r)
#endif
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const bool multi_d = dp.n_dist > 1;
  const uint32_t hn = H ? uint32_t(H) : bf.hash_num;
  const uint32_t h_first = hn < 2u ? hn : 2u;
  // one Bloom bit per visit of S2 / S4 (queue entries carry the bit index above the 9 bits of the candidate index)
  constexpr bool MAP = MODE == 1;
  constexpr bool kSeq = (MODE == 2 || LINKS_SEQ_PROBES != 0) && H != 0 && kExtractTile <= 512;
  constexpr int kFirst = H >= 2 ? 2 : H;        // compile-time split when H is known
  const uint32_t n_seg = rb.n_seq * dp.n_dist;
  unsigned long long n_cand = 0, n_surv = 0;
  uint32_t n_probe = 0, n_load = 0, n_msec = 0;  // reference-order bit tests, bit tests issued, member-set sectors
  unsigned long long n_probe_acc = 0, n_load_acc = 0, n_msec_acc = 0;
  const bool has_set = mv.buckets != nullptr;

  for (;;) {
    uint32_t tile = 0;
    if (lane == 0) tile = atomicAdd(rb.tile_cursor, 1u);
    tile = __shfl_sync(0xffffffffu, tile, 0);
    if (tile >= rb.n_tiles) break;

    // ---- tile header (warp-uniform; every lane computes the same values)
    const uint32_t s = find_segment_warp(rb.seg_tile_start, n_seg, tile);
    const uint32_t r = s / dp.n_dist, di = s - r * dp.n_dist;
    const uint64_t g0 = rb.off[r] - rb.base0;
    const uint64_t len = rb.off[r + 1] - rb.off[r];
    const uint64_t npos = len - k + 1;  // reads shorter than k own zero tiles
    const uint64_t d = dp.dist[di], t = dp.step[di], o = dp.offset;
    const bool flagged = rb.flags[r] != 0;
    RankView rv;
    uint32_t va0 = 0, vb0 = 0;
    unsigned long long cnt = 0;
    if (!flagged) {
      if (npos > o + d) cnt = (npos - o - d) / t;  // a[n] = o+n, b[n] = o+d+n
    } else {
      rv.vs = rb.vs; rv.vrank = rb.vrank; rv.wb = (g0 >> 6) + 2ull * r; rv.npos = npos;
      rv.nwords = (npos + 63) >> 6; rv.total = rb.vrank[rv.wb + rv.nwords];
      va0 = rank_of(rv, o); vb0 = rank_of(rv, o + d);
      const uint32_t na = rv.total - va0, nb = rv.total - vb0;
      cnt = (na < nb ? na : nb) / t;
    }
    // member_map.cuh: the filter's answers for this read's k-mers are already in the byte map (reads without an invalid base)
    const bool use_map = MAP && !flagged;
    const uint8_t* map_r = use_map ? mm.bytes + mm.base[r] : nullptr;
    const uint64_t j_begin = uint64_t(tile - rb.seg_tile_start[s]) * kExtractTile;
    const uint32_t n_tile = j_begin < cnt ? uint32_t(min((unsigned long long)kExtractTile, cnt - j_begin)) : 0u;
    if (lane == 0) n_cand += n_tile;

    // rank among valid k-mers of candidate q of this tile (ref :530-543), and the two positions
    auto rank_n = [&](uint32_t q) -> uint64_t { return (j_begin + q) * t + (t - 1); };
    auto pos_a = [&](uint64_t n) -> uint64_t { return g0 + (flagged ? select_of(rv, va0 + uint32_t(n)) : o + n); };
    auto pos_b = [&](uint64_t n) -> uint64_t { return g0 + (flagged ? select_of(rv, vb0 + uint32_t(n)) : o + n + d); };

    uint32_t next_q = 0, n1 = 0, n2 = 0, n3 = 0, n4 = 0, n5 = 0;
    for (;;) {
      const bool feeding = next_q < n_tile;
      // deepest stage with a full warp of input first; when the tile is exhausted, drain front to back
      int stage;
      if (n5 >= 32) stage = 6;
      else if (n4 >= 32) stage = 5;
      else if (n3 >= 32) stage = 4;
      else if (n2 >= 32) stage = 3;
      else if (n1 >= 32) stage = 2;
      else if (feeding) stage = 1;
      else if (n1) stage = 2;
      else if (n2) stage = 3;
      else if (n3) stage = 4;
      else if (n4) stage = 5;
      else if (n5) stage = 6;
      else break;

      if (stage == 1 && use_map) {  // ---- S1 with the map: two byte loads answer contains(A) && contains(B); survivors go to S5
        const uint32_t q = next_q + lane;
        bool pass = false;
        if (q < n_tile) {
          const uint64_t n = rank_n(q);
          const uint8_t ba = __ldg(map_r + map_entry_of(mm, n));
          n_probe += ba & 0x7Fu;
          if (ba & 0x80u) {  // ref :554: B is only asked when A is in the filter
            const uint8_t bb = __ldg(map_r + map_entry_of(mm, n + d));
            n_probe += bb & 0x7Fu;
            pass = (bb & 0x80u) != 0;
          }
        }
        next_q += 32;
        uint32_t slot;
        n4 = warp_push_slot(pass, n4, slot);
        if (pass) Q.q4_i[slot] = uint16_t(q);
      } else if (stage == 1) {  // ---- S1: hash A, first probes
        const uint32_t q = next_q + lane;
        bool pass = false, member = false;
        uint64_t h0 = 0;
        uint32_t i_next = 0;  // kSeq: the next Bloom bit to test rides in the queue entry's high bits
        if (q < n_tile) {
          uint64_t f, rr;
          LINKS_HASH(pos_a(rank_n(q)), f, rr);
          h0 = rr < f ? rr : f;
#if LINKS_PROBE_WITH_SET
          // set bucket and the first probes are independent loads: issue them together (one latency, not two);
          // a proven member's probe result is simply not used
          uint32_t np = 0;
          pass = H ? bloom_probe_fixed<kFirst, SMALL>(bf, h0, kseed, 0, np) : bloom_probe_loop<SMALL>(bf, h0, kseed, 0, h_first, np);
          n_load += h_first;
          if (has_set) member = member_contains(mv, h0, n_msec);
          if (member) { n_probe += hn; pass = false; } else n_probe += np;
#else
          if (has_set) member = member_contains(mv, h0, n_msec);
          if (member) n_probe += hn;  // the filter holds every set key: all hash_num bits are set, no need to look
          else if (kSeq) {
            // bits are tested one per visit of S2; without a set the first one is tested here
            if (has_set) pass = true;
            else {
              ++n_probe; ++n_load;
              pass = bloom_bit<SMALL>(bf, h0, kseed, 0) != 0;
              i_next = 1;
              if (hn == 1) { member = pass; pass = false; }
            }
          } else {
            pass = H ? bloom_probe_fixed<kFirst, SMALL>(bf, h0, kseed, 0, n_probe) : bloom_probe_loop<SMALL>(bf, h0, kseed, 0, h_first, n_probe);
            n_load += h_first;
          }
#endif
        }
        next_q += 32;
        uint32_t slot;
        n1 = warp_push_slot(pass, n1, slot);
        if (pass) { Q.q1_i[slot] = uint16_t(q | (i_next << 9)); Q.q1_h[slot] = h0; }
        if (has_set || kSeq) {  // proven members skip S2
          n2 = warp_push_slot(member, n2, slot);
          if (member) Q.q2_i[slot] = uint16_t(q);
        }
      } else if (stage == 2) {  // ---- S2: remaining probes of A
        const uint32_t take = n1 < 32u ? n1 : 32u;
        n1 -= take;
        bool pass = false, again = false;
        uint16_t q = 0;
        uint64_t h0 = 0;
        if (uint32_t(lane) < take) {
          q = Q.q1_i[n1 + lane];
          h0 = Q.q1_h[n1 + lane];
          if (kSeq) {
            const uint32_t i = q >> 9;
            ++n_probe; ++n_load;
            const bool bit = bloom_bit<SMALL>(bf, h0, kseed, i) != 0;
            pass = bit && i + 1 == hn;
            again = bit && !pass;
            q = uint16_t(again ? q + 512u : q & 511u);
          } else pass = bloom_probe_rest<H, SMALL>(bf, h0, kseed, h_first, hn, n_probe, n_load);
        }
        __syncwarp();
        uint32_t slot;
        n2 = warp_push_slot(pass, n2, slot);
        if (pass) Q.q2_i[slot] = q;
        if (kSeq) {  // still in the filter after this bit: back into the queue for the next one
          n1 = warp_push_slot(again, n1, slot);
          if (again) { Q.q1_i[slot] = q; Q.q1_h[slot] = h0; }
        }
      } else if (stage == 3) {  // ---- S3: hash B, first probes (ref :554 tests B only if A is in the filter)
        const uint32_t take = n2 < 32u ? n2 : 32u;
        n2 -= take;
        bool pass = false, member = false;
        uint16_t q = 0;
        uint64_t f = 0, rr = 0;
        if (uint32_t(lane) < take) {
          q = Q.q2_i[n2 + lane];
          LINKS_HASH(pos_b(rank_n(q)), f, rr);
          const uint64_t h0 = rr < f ? rr : f;
#if LINKS_PROBE_WITH_SET
          uint32_t np = 0;
          pass = H ? bloom_probe_fixed<kFirst, SMALL>(bf, h0, kseed, 0, np) : bloom_probe_loop<SMALL>(bf, h0, kseed, 0, h_first, np);
          n_load += h_first;
          if (has_set) member = member_contains(mv, h0, n_msec);
          if (member) { n_probe += hn; pass = false; } else n_probe += np;
#else
          if (has_set) member = member_contains(mv, h0, n_msec);
          if (member) n_probe += hn;
          else if (kSeq) {
            if (has_set) pass = true;
            else {
              ++n_probe; ++n_load;
              pass = bloom_bit<SMALL>(bf, h0, kseed, 0) != 0;
              q = uint16_t(q + 512u);
              if (hn == 1) { member = pass; pass = false; q = uint16_t(q & 511u); }
            }
          } else {
            pass = H ? bloom_probe_fixed<kFirst, SMALL>(bf, h0, kseed, 0, n_probe) : bloom_probe_loop<SMALL>(bf, h0, kseed, 0, h_first, n_probe);
            n_load += h_first;
          }
#endif
        }
        __syncwarp();
        uint32_t slot;
        n3 = warp_push_slot(pass, n3, slot);
        if (pass) { Q.q3_i[slot] = q; Q.q3_f[slot] = f; Q.q3_r[slot] = rr; }
        if (has_set || kSeq) {  // proven members skip S4
          n4 = warp_push_slot(member, n4, slot);
          if (member) { Q.q4_i[slot] = q; Q.q4_f[slot] = f; Q.q4_r[slot] = rr; }
        }
      } else if (stage == 4) {  // ---- S4: remaining probes of B
        const uint32_t take = n3 < 32u ? n3 : 32u;
        n3 -= take;
        bool pass = false, again = false;
        uint16_t q = 0;
        uint64_t f = 0, rr = 0;
        if (uint32_t(lane) < take) {
          q = Q.q3_i[n3 + lane]; f = Q.q3_f[n3 + lane]; rr = Q.q3_r[n3 + lane];
          const uint64_t h0 = rr < f ? rr : f;
          if (kSeq) {
            const uint32_t i = q >> 9;
            ++n_probe; ++n_load;
            const bool bit = bloom_bit<SMALL>(bf, h0, kseed, i) != 0;
            pass = bit && i + 1 == hn;
            again = bit && !pass;
            q = uint16_t(again ? q + 512u : q & 511u);
          } else pass = bloom_probe_rest<H, SMALL>(bf, h0, kseed, h_first, hn, n_probe, n_load);
        }
        __syncwarp();
        uint32_t slot;
        n4 = warp_push_slot(pass, n4, slot);
        if (pass) { Q.q4_i[slot] = q; Q.q4_f[slot] = f; Q.q4_r[slot] = rr; }
        if (kSeq) {
          n3 = warp_push_slot(again, n3, slot);
          if (again) { Q.q3_i[slot] = q; Q.q3_f[slot] = f; Q.q3_r[slot] = rr; }
        }
      } else if (stage == 5) {  // ---- S5: pair key (ref :546-552) + first-wins insert (ref :370-373)
        const uint32_t take = n4 < 32u ? n4 : 32u;
        n4 -= take;
        bool fresh = false;
        uint64_t h_a = 0, h_b = 0, okey_r = 0;
        if (uint32_t(lane) < take) {
          const uint32_t q = Q.q4_i[n4 + lane];
          uint64_t fb, rbv;
          const uint64_t n = rank_n(q);
          if (use_map) LINKS_HASH(pos_b(n), fb, rbv);  // the map path carries no hashes
          else { fb = Q.q4_f[n4 + lane]; rbv = Q.q4_r[n4 + lane]; }
          uint64_t fa, ra;
          LINKS_HASH(pos_a(n), fa, ra);
          if (fa < rbv) { h_a = rbv; h_b = ra; } else { h_a = fa; h_b = fb; }
          if (!ROUTED) fresh = pair_insert(pv, h_a, h_b, make_okey(rb.first_read_idx + r, di, n), multi_d);
          else okey_r = make_okey(rb.first_read_idx + r, di, n);
        }
        if (lane == 0) n_surv += take;
        __syncwarp();
        if (ROUTED) {
          const bool live = uint32_t(lane) < take;
          const uint32_t dst = live ? dest_rank(h_a, h_b, L.n_ranks) : 0xFFFFFFFFu;
          const uint32_t peers = __match_any_sync(0xffffffffu, dst);
          const int leader = __ffs(peers) - 1;
          unsigned long long base = 0;
          if (live && lane == leader) base = atomicAdd(&x->cursor[dst], (unsigned long long)__popc(peers));
          base = __shfl_sync(0xffffffffu, base, leader);
          if (live) {
            const unsigned long long slot = base + __popc(peers & ((1u << lane) - 1u));
            if (slot < L.pair_cap) {
              uint64_t* rec = inbox_pair_region(x->inbox[dst], L, epoch, L.rank) + slot * L.rec_words;
              if (L.rec_words == 2) *reinterpret_cast<ulonglong2*>(rec) = make_ulonglong2(h_a, h_b);
              else { rec[0] = h_a; rec[1] = h_b; rec[2] = okey_r; }
            } else {
              x->route_overflow = 1;
            }
          }
        } else {
          uint32_t slot;
          n5 = warp_push_slot(fresh, n5, slot);
          if (fresh) { Q.q5_a[slot] = h_a; Q.q5_b[slot] = h_b; }
        }
      } else {  // ---- append fresh pairs to the dense unique-pair list: one global atomic per 32 pairs
        const uint32_t take = n5 < 32u ? n5 : 32u;
        n5 -= take;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&ctr->n_unique, (unsigned long long)take);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (uint32_t(lane) < take) pv.list[base + lane] = make_ulonglong2(Q.q5_a[n5 + lane], Q.q5_b[n5 + lane]);
      }
      __syncwarp();
    }
    n_probe_acc += n_probe; n_load_acc += n_load; n_msec_acc += n_msec;
    n_probe = 0; n_load = 0; n_msec = 0;
  }
  n_probe_acc = warp_sum_ull(n_probe_acc);
  n_load_acc = warp_sum_ull(n_load_acc);
  n_msec_acc = warp_sum_ull(n_msec_acc);
  if (lane == 0) {
    if (n_cand) atomicAdd(&ctr->n_candidates, n_cand);
    if (n_surv) atomicAdd(&ctr->n_survivors, n_surv);
    if (n_probe_acc) atomicAdd(&ctr->n_bf_probes, n_probe_acc);
    if (n_load_acc) atomicAdd(&ctr->n_bf_loads, n_load_acc);
    if (n_msec_acc) atomicAdd(&ctr->n_member_sectors, n_msec_acc);
  }
  if (ROUTED) {  // every record of this launch is visible to its owner before the count that announces it
    if (last_block_after_fence(&x->blocks_done)) publish_counts(x, L, epoch, 0);
  }
}

#undef LINKS_HASH

}  // namespace links

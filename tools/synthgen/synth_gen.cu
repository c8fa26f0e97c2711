// tools/synthgen/synth_gen.cu -- device-side generator of SURVEY 8(d)'s synthetic inputs for workloads that are too
// large for the numpy generator (links_b200/synth.py): BASELINE.json configs[3], a 3.1 Gbp draft with 20x 20 kb reads
// = 62 Gbases.  BENCH / TEST INFRASTRUCTURE, not part of the product ABI (include/links_gpu.h) and not linked into
// liblinks_gpu.so.  Same shapes and error model as synth.py (i.i.d. genome with copied repeats, contigs cut from it with
// gaps and strand flips, ONT-like reads with 5 % errors sub:ins:del = 1:1:1, a few reads with an N), but from a
// counter-based generator: every output byte is a pure function of (seed, index), so any rank can generate any slice
// of the job and N ranks see exactly the reads one rank would.
#include <cstdint>
#include <cuda_runtime.h>

namespace {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x)
{
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

__device__ __forceinline__ uint8_t base_of(uint32_t c) { return uint8_t(0x54474341u >> (8 * (c & 3u))); }  // "ACGT"
__device__ __forceinline__ uint8_t complement(uint8_t b)
{
  switch (b) { case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A'; default: return b; }
}

// raw genome: 32 bases per 64-bit draw
__global__ void genome_kernel(uint8_t* __restrict__ out, uint64_t n, uint64_t seed)
{
  const uint64_t w = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint64_t g0 = w * 32;
  if (g0 >= n) return;
  uint64_t r = mix64(seed ^ (w * 0xD1B54A32D192ED03ULL));
  const int m = int(n - g0 < 32 ? n - g0 : 32);
  for (int j = 0; j < m; ++j) { out[g0 + j] = base_of(uint32_t(r)); r >>= 2; }
}

// range i: dst[dst_pos[i] .. +len[i]) <- src[src_pos[i] .. +len[i]), reverse-complemented when rc[i]
__global__ void copy_ranges_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, const uint64_t* __restrict__ dst_pos,
                                   const uint64_t* __restrict__ src_pos, const uint64_t* __restrict__ len, const uint8_t* __restrict__ rc,
                                   uint32_t n)
{
  for (uint32_t i = blockIdx.y; i < n; i += gridDim.y) {
    const uint64_t L = len[i];
    const uint8_t* s = src + src_pos[i];
    uint8_t* d = dst + dst_pos[i];
    const bool flip = rc && rc[i] != 0;
    for (uint64_t j = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; j < L; j += uint64_t(gridDim.x) * blockDim.x)
      d[j] = flip ? complement(s[L - 1 - j]) : s[j];
  }
}

__global__ void fill_runs_kernel(uint8_t* __restrict__ buf, const uint64_t* __restrict__ pos, const uint32_t* __restrict__ len, uint32_t n, uint8_t v)
{
  const uint32_t i = blockIdx.x;
  if (i >= n) return;
  for (uint32_t j = threadIdx.x; j < len[i]; j += blockDim.x) buf[pos[i] + j] = v;
}

// error draw of source base i of read `rid`: bytes emitted (0 deletion, 1 none/substitution, 2 insertion after)
struct Draw { uint32_t emit; uint32_t kind; uint64_t u; };
__device__ __forceinline__ Draw draw_of(uint64_t seed, uint64_t rid, uint64_t i, uint32_t err_q20)
{
  Draw d;
  d.u = mix64(seed ^ (rid * 0xA24BAED4963EE407ULL) ^ (i * 0x9FB21C651E98DF25ULL));
  const bool e = uint32_t(d.u & 0xFFFFFu) < err_q20;
  d.kind = e ? 1u + uint32_t((d.u >> 20) % 3u) : 0u;  // 1 substitution, 2 insertion after, 3 deletion
  d.emit = d.kind == 2u ? 2u : d.kind == 3u ? 0u : 1u;
  return d;
}

constexpr int kReadThreads = 256;

__device__ __forceinline__ uint32_t block_sum(uint32_t v, uint32_t* sm)
{
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  uint32_t t = 0;
  for (int w = 0; w < kReadThreads / 32; ++w) t += sm[w];
  __syncthreads();
  return t;
}

// out_len[r] = bytes read (first + r) will have
__global__ void __launch_bounds__(kReadThreads) read_len_kernel(const uint64_t* __restrict__ len, uint32_t n, uint64_t first, uint64_t seed,
                                                                uint32_t err_q20, uint64_t* __restrict__ out_len)
{
  __shared__ uint32_t sm[kReadThreads / 32];
  for (uint32_t r = blockIdx.x; r < n; r += gridDim.x) {
    uint32_t c = 0;
    for (uint64_t i = threadIdx.x; i < len[r]; i += kReadThreads) c += draw_of(seed, first + r, i, err_q20).emit;
    const uint32_t t = block_sum(c, sm);
    if (threadIdx.x == 0) out_len[r] = t;
  }
}

// write read (first + r) at out + off[r]: source genome[start[r] .. +len[r]) with errors applied, then strand flip
__global__ void __launch_bounds__(kReadThreads) read_emit_kernel(const uint8_t* __restrict__ genome, const uint64_t* __restrict__ start,
                                                                 const uint64_t* __restrict__ len, const uint8_t* __restrict__ rc,
                                                                 const uint64_t* __restrict__ off, uint32_t n, uint64_t first, uint64_t seed,
                                                                 uint32_t err_q20, uint32_t n_read_q20, uint8_t* __restrict__ out)
{
  __shared__ uint32_t s_scan[kReadThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t r = blockIdx.x; r < n; r += gridDim.x) {
    const uint64_t rid = first + r, L = len[r], olen = off[r + 1] - off[r];
    const uint8_t* src = genome + start[r];
    uint8_t* dst = out + off[r];
    const bool flip = rc[r] != 0;
    uint64_t base = 0;
    for (uint64_t i0 = 0; i0 < L; i0 += kReadThreads) {
      const uint64_t i = i0 + threadIdx.x;
      Draw d; d.emit = 0; d.kind = 0; d.u = 0;
      if (i < L) d = draw_of(seed, rid, i, err_q20);
      // block exclusive scan of d.emit
      uint32_t v = d.emit;
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
      if (lane == 31) s_scan[warp] = v;
      __syncthreads();
      uint32_t wbase = 0, total = 0;
      for (int w = 0; w < kReadThreads / 32; ++w) { if (w < warp) wbase += s_scan[w]; total += s_scan[w]; }
      __syncthreads();
      if (i < L && d.emit) {
        uint64_t o = base + wbase + v - d.emit;
        uint8_t b = src[i];
        if (d.kind == 1u) {  // a different base: rotate within ACGT by 1..3
          const uint32_t code = b == 'A' ? 0u : b == 'C' ? 1u : b == 'G' ? 2u : 3u;
          b = base_of(code + 1u + uint32_t((d.u >> 24) % 3u));
        }
        const uint64_t p0 = flip ? olen - 1 - o : o;
        dst[p0] = flip ? complement(b) : b;
        if (d.emit == 2u) {
          const uint8_t ins = base_of(uint32_t(d.u >> 28));
          const uint64_t p1 = flip ? olen - 2 - o : o + 1;
          dst[p1] = flip ? complement(ins) : ins;
        }
      }
      base += total;
    }
    // ~0.1 % of the reads carry one N
    __syncthreads();
    if (threadIdx.x == 0 && olen > 10) {
      const uint64_t u = mix64(seed ^ 0x5851F42D4C957F2DULL ^ (rid * 0xDA942042E4DD58B5ULL));
      if (uint32_t(u & 0xFFFFFu) < n_read_q20) dst[(u >> 20) % olen] = 'N';
    }
  }
}

}  // namespace

extern "C" {

// every function returns a cudaError_t value (0 = success) after synchronising the default stream

int synth_genome(uint8_t* dev_out, uint64_t n, uint64_t seed)
{
  if (n) genome_kernel<<<unsigned(((n + 31) / 32 + 255) / 256), 256>>>(dev_out, n, seed);
  return int(cudaDeviceSynchronize());
}

// dst[dst_pos[i] .. +len[i]) <- src[src_pos[i] .. +len[i]) (reverse complement when rc[i]; rc may be NULL).  Used for the
// genome's copied repeats (through a scratch buffer, so overlapping repeats do not chain) and to cut the contigs.
int synth_copy_ranges(const uint8_t* dev_src, uint8_t* dev_dst, const uint64_t* dev_dst_pos, const uint64_t* dev_src_pos,
                      const uint64_t* dev_len, const uint8_t* dev_rc, uint32_t n)
{
  if (n) copy_ranges_kernel<<<dim3(16, n < 8192 ? n : 8192), 256>>>(dev_src, dev_dst, dev_dst_pos, dev_src_pos, dev_len, dev_rc, n);
  return int(cudaDeviceSynchronize());
}

int synth_fill_runs(uint8_t* dev_buf, const uint64_t* dev_pos, const uint32_t* dev_len, uint32_t n, uint8_t value)
{
  if (n) fill_runs_kernel<<<n, 64>>>(dev_buf, dev_pos, dev_len, n, value);
  return int(cudaDeviceSynchronize());
}

int synth_read_lengths(const uint64_t* dev_len, uint32_t n, uint64_t first_read, uint64_t seed, double err, uint64_t* dev_out_len)
{
  if (n) read_len_kernel<<<n < 148 * 16 ? n : 148 * 16, kReadThreads>>>(dev_len, n, first_read, seed, uint32_t(err * 1048576.0), dev_out_len);
  return int(cudaDeviceSynchronize());
}

int synth_reads(const uint8_t* dev_genome, const uint64_t* dev_start, const uint64_t* dev_len, const uint8_t* dev_rc, const uint64_t* dev_off,
                uint32_t n, uint64_t first_read, uint64_t seed, double err, double n_read_frac, uint8_t* dev_out)
{
  if (n) read_emit_kernel<<<n < 148 * 16 ? n : 148 * 16, kReadThreads>>>(dev_genome, dev_start, dev_len, dev_rc, dev_off, n, first_read, seed,
                                                                         uint32_t(err * 1048576.0), uint32_t(n_read_frac * 1048576.0), dev_out);
  return int(cudaDeviceSynchronize());
}

}  // extern "C"

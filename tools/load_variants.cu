// tools/load_variants.cu -- which global-load flavour costs the least DRAM traffic per random 4-byte read on B200?
// MEASUREMENT PROBE, not product code.  ncu showed the table kernels' L2 seeing ~3x the read sectors their loads ask
// for; this runs the same random-gather loop (tools/gather_peak.cu, 4 loads in flight per thread) with one PTX load
// flavour per launch over a 1 GB table and prints G loads/s; run it under
//   ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum
// to get DRAM bytes and L2 sectors per load for each flavour.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix64(uint64_t x)
{
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

template <int V>
__device__ __forceinline__ uint32_t load4(const uint32_t* p)
{
  uint32_t v;
  if (V == 0) asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 1) asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 2) asm volatile("ld.global.cs.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 3) asm volatile("ld.global.cv.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 4) asm volatile("ld.global.lu.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 5) asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 6) asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 7) asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 8) asm volatile("ld.global.nc.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 9) asm volatile("ld.global.nc.L2::128B.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (V == 10) asm volatile("atom.global.or.b32 %0, [%1], 0;" : "=r"(v) : "l"(p));
  else if (V == 11) asm volatile("ld.global.nc.L1::evict_first.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else {  // 32-byte load with an L2 evict_first hint (the hint exists for 256-bit loads only)
    unsigned long long a, b, c, d;
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(reinterpret_cast<uintptr_t>(p) & ~uintptr_t(31)));
    v = uint32_t(a ^ b ^ c ^ d);
  }
  return v;
}

template <int V>
__global__ void __launch_bounds__(256) gather_kernel(const uint8_t* __restrict__ tab, uint64_t n_sectors, uint32_t rounds, uint64_t* __restrict__ sink)
{
  const uint64_t tid = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  uint64_t acc = 0;
  for (uint32_t r = 0; r < rounds; ++r) {
    uint32_t v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint64_t h = mix64((tid * rounds + r) * 4 + u);
      v[u] = load4<V>(reinterpret_cast<const uint32_t*>(tab + __umul64hi(h, n_sectors) * 32 + ((h >> 3) & 28)));
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) acc ^= v[u];
  }
  if (acc == 0x1234567ull) sink[0] = acc;
}

template <int V>
static void run(const char* name, const uint8_t* tab, uint64_t n_sectors, uint64_t* sink, int grid, uint64_t mb)
{
  const uint64_t threads = uint64_t(grid) * 256;
  const uint32_t rounds = uint32_t((1ull << 28) / (threads * 4));
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e30f;
  for (int it = 0; it < 3; ++it) {
    cudaEventRecord(a);
    gather_kernel<V><<<grid, 256>>>(tab, n_sectors, rounds, sink);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (it && ms < best) best = ms;
  }
  const cudaError_t e = cudaGetLastError();
  printf("{\"probe\": \"load_variants\", \"variant\": %d, \"ptx\": \"%s\", \"footprint_mb\": %llu, \"loads\": %.0f, \"ms\": %.4f, \"gloads_per_s\": %.2f, \"cuda\": \"%s\"}\n",
         V, name, (unsigned long long)mb, double(threads) * rounds * 4, best, double(threads) * rounds * 4 / (best * 1e-3) / 1e9, cudaGetErrorString(e));
  fflush(stdout);
}

int main(int argc, char** argv)
{
  const uint64_t mb = argc > 1 ? strtoull(argv[1], nullptr, 10) : 1024;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 1; }
  const int grid = prop.multiProcessorCount * 8;
  uint64_t* sink;
  cudaMalloc(&sink, 8);
  uint8_t* tab;
  if (cudaMalloc(&tab, mb << 20) != cudaSuccess) { fprintf(stderr, "cudaMalloc failed\n"); return 1; }
  cudaMemset(tab, 0x5A, mb << 20);
  const uint64_t ns = (mb << 20) / 32;
  run<0>("ld.global.nc", tab, ns, sink, grid, mb);
  run<1>("ld.global.cg", tab, ns, sink, grid, mb);
  run<2>("ld.global.cs", tab, ns, sink, grid, mb);
  run<3>("ld.global.cv", tab, ns, sink, grid, mb);
  run<4>("ld.global.lu", tab, ns, sink, grid, mb);
  run<5>("ld.global.nc.L1::no_allocate", tab, ns, sink, grid, mb);
  run<6>("ld.global.ca", tab, ns, sink, grid, mb);
  run<7>("ld.relaxed.gpu.global", tab, ns, sink, grid, mb);
  run<8>("ld.global.nc.L2::64B", tab, ns, sink, grid, mb);
  run<9>("ld.global.nc.L2::128B", tab, ns, sink, grid, mb);
  run<10>("atom.global.or.b32 x,0", tab, ns, sink, grid, mb);
  run<11>("ld.global.nc.L1::evict_first", tab, ns, sink, grid, mb);
  run<12>("ld.global.nc.L1::no_allocate.L2::evict_first.v4.u64 (32 B)", tab, ns, sink, grid, mb);
  return 0;
}

// tools/gather_peak.cu -- measured ceiling for the access pattern that bounds the LINKS hot path on B200: independent
// random reads of one 32-byte sector each (a Bloom-filter probe reads 4 bytes of a random sector, a member-set or
// pair-set lookup reads the whole 32-byte bucket).  MEASUREMENT PROBE, not product code.
//
//   tools/gather_peak [--mb 64,1024,8192] [--iters 3] [--fetch 32|64|128]
//
// For every footprint (a table of that many MB of 32-byte sectors), load width (4 or 32 bytes per sector) and number
// of independent loads in flight per thread (1..16) it prints one JSON line: sectors/s, GB/s counted as 32 bytes per
// sector touched, and the same as a fraction of the streaming-copy peak passed with --copy-gbs.  Addresses come from a
// counter hash (no dependent chain), the grid is 148 SMs x 8 CTAs x 256 threads, timing is CUDA events around one
// launch after a warm-up launch.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <string>
#include <vector>

__device__ __forceinline__ uint64_t mix64(uint64_t x)
{
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

template <int U, int WIDTH>
__global__ void __launch_bounds__(256) gather_kernel(const uint8_t* __restrict__ tab, uint64_t n_sectors, uint32_t rounds, uint64_t* __restrict__ sink)
{
  const uint64_t tid = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  uint64_t acc = 0;
  for (uint32_t r = 0; r < rounds; ++r) {
    uint64_t v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t h = mix64((tid * rounds + r) * U + u);
      const uint64_t s = __umul64hi(h, n_sectors);
      if (WIDTH == 4) {
        v[u] = __ldg(reinterpret_cast<const uint32_t*>(tab + s * 32 + ((h >> 3) & 28)));
      } else {
        unsigned long long a, b, c, d;
        asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(tab + s * 32));
        v[u] = a ^ b ^ c ^ d;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) acc ^= v[u];
  }
  if (acc == 0x1234567ull) sink[0] = acc;  // keeps the loads alive
}

template <int U, int WIDTH>
static float run(const uint8_t* tab, uint64_t n_sectors, uint32_t rounds, uint64_t* sink, int grid, int iters)
{
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  gather_kernel<U, WIDTH><<<grid, 256>>>(tab, n_sectors, rounds, sink);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int i = 0; i < iters; ++i) {
    cudaEventRecord(a);
    gather_kernel<U, WIDTH><<<grid, 256>>>(tab, n_sectors, rounds, sink);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (ms < best) best = ms;
  }
  cudaEventDestroy(a); cudaEventDestroy(b);
  return best;
}

int main(int argc, char** argv)
{
  std::vector<uint64_t> mbs = {64, 1024, 8192};
  int iters = 3;
  double copy_gbs = 0;
  int fetch = 0;  // cudaLimitMaxL2FetchGranularity to ask for (32 / 64 / 128); 0 = leave the default
  for (int i = 1; i < argc; ++i) {
    if (!strcmp(argv[i], "--mb") && i + 1 < argc) {
      mbs.clear();
      for (char* tok = strtok(argv[++i], ","); tok; tok = strtok(nullptr, ",")) mbs.push_back(strtoull(tok, nullptr, 10));
    } else if (!strcmp(argv[i], "--iters") && i + 1 < argc) iters = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--copy-gbs") && i + 1 < argc) copy_gbs = atof(argv[++i]);
    else if (!strcmp(argv[i], "--fetch") && i + 1 < argc) fetch = atoi(argv[++i]);
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 1; }
  const int grid = prop.multiProcessorCount * 8;
  if (fetch) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, size_t(fetch));
  size_t fetch_now = 0;
  cudaDeviceGetLimit(&fetch_now, cudaLimitMaxL2FetchGranularity);
  uint64_t* sink;
  cudaMalloc(&sink, 8);
  for (uint64_t mb : mbs) {
    const uint64_t bytes = mb << 20, n_sectors = bytes / 32;
    uint8_t* tab;
    if (cudaMalloc(&tab, bytes) != cudaSuccess) { fprintf(stderr, "cudaMalloc %llu MB failed\n", (unsigned long long)mb); continue; }
    cudaMemset(tab, 0x5A, bytes);
    const uint64_t threads = uint64_t(grid) * 256;
    const uint64_t target = 1ull << 29;  // sector reads per launch
#define RUN(U, W)                                                                                                              \
  {                                                                                                                            \
    const uint32_t rounds = uint32_t(target / (threads * U)) ? uint32_t(target / (threads * U)) : 1u;                          \
    const float ms = run<U, W>(tab, n_sectors, rounds, sink, grid, iters);                                                     \
    const double sectors = double(threads) * rounds * U;                                                                       \
    const double gbs = sectors * 32.0 / (ms * 1e-3) / 1e9;                                                                     \
    printf("{\"probe\": \"random_sector_gather\", \"gpu\": \"%s\", \"footprint_mb\": %llu, \"load_bytes\": %d, \"in_flight_per_thread\": %d, "   \
           "\"l2_fetch_granularity\": %d, \"threads\": %llu, \"sectors\": %.0f, \"ms\": %.4f, \"gsectors_per_s\": %.3f, \"gb_per_s_32B_sectors\": %.1f, \"frac_of_copy_peak\": %s}\n", \
           prop.name, (unsigned long long)mb, W, U, int(fetch_now), (unsigned long long)threads, sectors, ms, sectors / (ms * 1e-3) / 1e9, gbs,   \
           copy_gbs > 0 ? std::to_string(gbs / copy_gbs).c_str() : "null");                                                    \
    fflush(stdout);                                                                                                            \
  }
    RUN(1, 4) RUN(2, 4) RUN(4, 4) RUN(8, 4) RUN(16, 4)
    RUN(1, 32) RUN(2, 32) RUN(4, 32) RUN(8, 32) RUN(16, 32)
#undef RUN
    cudaFree(tab);
  }
  cudaFree(sink);
  return 0;
}

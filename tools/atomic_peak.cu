// tools/atomic_peak.cu -- measured ceiling for the Bloom build's update pattern on B200: fire-and-forget atomicOr (RED) on
// random 32-bit words of a block of F MB.  MEASUREMENT PROBE, not product code.
//
//   tools/atomic_peak [--mb 8,16,32,48,64,96,128,512] [--iters 3]
//
// Two sources of addresses per footprint: "hash" (computed from a counter, no memory read) and "stream" (4-byte entries
// read once with ld.global.cs from a 2 GB buffer, the access pattern of bloom_apply_kernel: does the stream push the
// block out of L2?).  One JSON line each: G updates/s.  Grid 148 x 8 x 256 as in the product kernel.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <vector>

__device__ __forceinline__ uint64_t mix64(uint64_t x)
{
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

__global__ void __launch_bounds__(256) red_hash_kernel(uint32_t* __restrict__ words, uint64_t n_words, uint64_t n)
{
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t h = mix64(i);
    atomicOr(words + __umul64hi(h, n_words), 1u << (h & 31));
  }
}

__global__ void __launch_bounds__(256) fill_kernel(uint32_t* __restrict__ e, uint64_t n, uint64_t n_words)
{
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) e[i] = uint32_t(__umul64hi(mix64(i), n_words));
}

__global__ void __launch_bounds__(256) red_stream_kernel(uint32_t* __restrict__ words, const uint32_t* __restrict__ e, uint64_t n)
{
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t w = __ldcs(e + i);
    atomicOr(words + w, 1u << (w & 31));
  }
}

int main(int argc, char** argv)
{
  std::vector<uint64_t> mbs = {8, 16, 32, 48, 64, 96, 128, 512};
  int iters = 3;
  for (int i = 1; i < argc; ++i) {
    if (!strcmp(argv[i], "--mb") && i + 1 < argc) {
      mbs.clear();
      for (char* tok = strtok(argv[++i], ","); tok; tok = strtok(nullptr, ",")) mbs.push_back(strtoull(tok, nullptr, 10));
    } else if (!strcmp(argv[i], "--iters") && i + 1 < argc) iters = atoi(argv[++i]);
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 1; }
  const int grid = prop.multiProcessorCount * 8;
  const uint64_t n = 1ull << 29;  // updates per launch
  uint32_t* entries;
  if (cudaMalloc(&entries, n * 4) != cudaSuccess) { fprintf(stderr, "cudaMalloc entries failed\n"); return 1; }
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  for (uint64_t mb : mbs) {
    const uint64_t bytes = mb << 20, n_words = bytes / 4;
    uint32_t* words;
    if (cudaMalloc(&words, bytes) != cudaSuccess) { fprintf(stderr, "cudaMalloc %llu MB failed\n", (unsigned long long)mb); continue; }
    cudaMemset(words, 0, bytes);
    fill_kernel<<<grid, 256>>>(entries, n, n_words);
    for (int mode = 0; mode < 2; ++mode) {
      float best = 1e30f;
      for (int it = 0; it < iters + 1; ++it) {
        cudaEventRecord(a);
        if (mode == 0) red_hash_kernel<<<grid, 256>>>(words, n_words, n);
        else red_stream_kernel<<<grid, 256>>>(words, entries, n);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        if (it && ms < best) best = ms;
      }
      printf("{\"probe\": \"random_word_atomic_or\", \"gpu\": \"%s\", \"footprint_mb\": %llu, \"addresses\": \"%s\", \"updates\": %llu, \"ms\": %.4f, \"gupdates_per_s\": %.2f}\n",
             prop.name, (unsigned long long)mb, mode ? "stream" : "hash", (unsigned long long)n, best, double(n) / (best * 1e-3) / 1e9);
      fflush(stdout);
    }
    cudaFree(words);
  }
  return 0;
}

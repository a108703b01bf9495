// Exploratory round 8: does token-bucket pacing lift plain STG.128 stores off their plateau too?
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)
__device__ __forceinline__ uint64_t gtimer() { uint64_t t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
struct TP {
  uint32_t base, hic, lo0;
  __device__ TP(uint64_t v0, uint32_t seed) { uint32_t hi = (uint32_t)(v0 >> 32); base = (hi ^ seed) * 0x85EBCA77u + seed; hic = hi ^ 0xC2B2AE3Du; lo0 = (uint32_t)v0; }
  __device__ __forceinline__ uint4 word(uint32_t j) const {
    uint32_t lo = lo0 + j, h = lo * 0x9E3779B1u + base; h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    return make_uint4(h, ~h, h * 0x297A2D39u + lo, __funnelshift_l(h, h, 13) ^ hic);
  }
};
// tile_vec words per tile, one tile per token; all threads of the CTA wait for the token
template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_stg_paced(uint4 *p, uint64_t n, uint32_t seed, uint32_t tile_vec, uint32_t period_x16) {
  const uint64_t tiles = n / tile_vec;
  const uint64_t t0 = gtimer();
  uint64_t next = (uint64_t)period_x16 * blockIdx.x / gridDim.x;   // staggered start
  for (uint64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    if (period_x16) {
      uint64_t now;
      while ((now = (gtimer() - t0) * 16) < next) {}
      if (now > next + period_x16) next = now;
      next += period_x16;
    }
    const uint64_t v0 = t * tile_vec;
    TP pat(v0, seed);
#pragma unroll 4
    for (uint32_t j = threadIdx.x; j < tile_vec; j += THREADS) p[v0 + j] = pat.word(j);
  }
}
template <typename F> static double bench(uint64_t bytes, F launch) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  std::vector<float> ts;
  for (int r = 0; r < 5; ++r) { CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (r) ts.push_back(ms); }
  std::sort(ts.begin(), ts.end());
  CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b));
  return bytes / ts[ts.size() / 2] / 1e6;
}
int main(int argc, char **argv) {
  double gib = argc > 1 ? atof(argv[1]) : 8.0;
  uint64_t bytes = (uint64_t)(gib * (1ull << 30)), n = bytes / 16;
  uint4 *p; CK(cudaMalloc(&p, bytes));
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const int rates[] = {0, 6400, 6800, 7000, 7200, 7300, 7400, 7500, 7600, 7800};
  printf("columns: target GB/s:"); for (int r : rates) printf(" %6d", r); printf("\n");
  for (int ctas : {1, 2, 4})
    for (uint32_t kib : {16u, 64u, 256u}) {
      uint32_t tv = kib * 64;
      printf("256 thr, %d CTA/SM, tile %3u KiB :", ctas, kib);
      for (int rate : rates) {
        int grid = sms * ctas;
        uint32_t per = rate ? (uint32_t)((double)tv * 16 * grid / rate * 16.0) : 0;
        printf(" %6.0f", bench(bytes, [&] { k_stg_paced<256><<<grid, 256>>>(p, n, 7, tv, per); }));
      }
      printf("\n"); fflush(stdout);
    }
  for (uint32_t kib : {64u, 256u}) {
    uint32_t tv = kib * 64;
    printf("1024 thr, 1 CTA/SM, tile %3u KiB :", kib);
    for (int rate : rates) {
      uint32_t per = rate ? (uint32_t)((double)tv * 16 * sms / rate * 16.0) : 0;
      printf(" %6.0f", bench(bytes, [&] { k_stg_paced<1024><<<sms, 1024>>>(p, n, 7, tv, per); }));
    }
    printf("\n");
  }
  return 0;
}

// Exploratory round 4: map the pacing landscape of bulk stores.
// knobs: tile size, sub size, stages, wait mode (smem-read vs completion), extra delay.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)
constexpr int kThreads = 256;
__device__ __forceinline__ uint4 pat(uint64_t i, uint32_t seed) {
  const uint32_t lo = (uint32_t)i, hi = (uint32_t)(i >> 32);
  uint32_t h = lo * 0x9E3779B1u + (hi ^ seed) * 0x85EBCA77u + seed;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
  return make_uint4(h, ~h, h * 0x297A2D39u + lo, __funnelshift_l(h, h, 13) ^ hi ^ 0xC2B2AE3Du);
}
// STAGES ring of tile_vec words; WAITC: wait for completion instead of smem-read
template <int STAGES, bool WAITC>
__global__ void __launch_bounds__(kThreads, 1)
k_ring(uint4 *p, uint64_t n, uint32_t seed, uint32_t tile_vec, uint32_t sub_vec, uint32_t delay) {
  extern __shared__ __align__(128) uint4 ring[];
  const uint32_t nsub = tile_vec / sub_vec;
  const uint64_t tiles = n / tile_vec;
  int stage = 0;
  for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const uint64_t v0 = tile * tile_vec;
    uint4 *buf = ring + (size_t)stage * tile_vec;
    if (threadIdx.x < nsub) {
      if (WAITC) asm volatile("cp.async.bulk.wait_group %0;" ::"n"(STAGES - 1) : "memory");
      else asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STAGES - 1) : "memory");
    }
    __syncthreads();
#pragma unroll 4
    for (uint32_t j = threadIdx.x; j < tile_vec; j += kThreads) buf[j] = pat(v0 + j, seed);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < nsub) {
      const uint32_t first = threadIdx.x * sub_vec;
      const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf + first);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + v0 + first), "r"(saddr),
                   "r"(sub_vec * 16u) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    if (delay) {
      long long t0 = clock64();
      while (clock64() - t0 < (long long)delay) {}
    }
    stage = stage + 1 == STAGES ? 0 : stage + 1;
  }
  if (threadIdx.x < nsub) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
template <typename F>
static double bench(uint64_t bytes, F launch) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  std::vector<float> ts;
  for (int r = 0; r < 4; ++r) {
    CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    if (r) ts.push_back(ms);
  }
  std::sort(ts.begin(), ts.end());
  CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b));
  return bytes / ts[ts.size() / 2] / 1e6;
}
template <int S, bool W>
static void run(uint4 *p, uint64_t n, uint64_t bytes, int sms, uint32_t tile_kib, uint32_t sub_kib) {
  const uint32_t tv = tile_kib * 64, sv = sub_kib * 64;
  const size_t smem = (size_t)tv * 16 * S;
  if (smem > 220 * 1024 || sv > tv || tv / sv > 32) return;
  CK(cudaFuncSetAttribute(k_ring<S, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  printf("S=%d wait=%s tile=%3uKiB sub=%2uKiB :", S, W ? "done" : "read", tile_kib, sub_kib);
  for (uint32_t delay : {0u, 100u, 200u, 400u, 800u, 1200u, 1600u, 2400u}) {
    double g = bench(bytes, [&] { k_ring<S, W><<<sms, kThreads, smem>>>(p, n, 7, tv, sv, delay); });
    printf(" %6.0f", g);
  }
  printf("\n");
  fflush(stdout);
}
int main(int argc, char **argv) {
  double gib = argc > 1 ? atof(argv[1]) : 8.0;
  uint64_t bytes = (uint64_t)(gib * (1ull << 30)) & ~((1ull << 21) - 1);
  uint64_t n = bytes / 16;
  uint4 *p; CK(cudaMalloc(&p, bytes));
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  printf("columns: delay cycles 0 100 200 400 800 1200 1600 2400 (GB/s, median of 3)\n");
  for (uint32_t tile : {16u, 32u, 64u})
    for (uint32_t sub : {2u, 4u, 8u, 16u, 32u, 64u}) {
      run<2, false>(p, n, bytes, sms, tile, sub);
      run<3, false>(p, n, bytes, sms, tile, sub);
      run<2, true>(p, n, bytes, sms, tile, sub);
      run<3, true>(p, n, bytes, sms, tile, sub);
      if (tile <= 32) { run<4, false>(p, n, bytes, sms, tile, sub); run<4, true>(p, n, bytes, sms, tile, sub); }
    }
  return 0;
}

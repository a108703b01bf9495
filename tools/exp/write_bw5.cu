// Exploratory round 5: token-bucket pacing of bulk stores on %globaltimer.
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
__device__ __forceinline__ uint64_t gtimer() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__global__ void k_timer_res(uint64_t *out) {
  uint64_t t0 = gtimer(), t = t0, mind = ~0ull;
  int changes = 0;
  while (changes < 64) {
    uint64_t t1 = gtimer();
    if (t1 != t) { uint64_t d = t1 - t; if (d < mind) mind = d; t = t1; ++changes; }
  }
  out[0] = mind; out[1] = t - t0;
}
// period_x16: tile period in 1/16 ns; stagger: spread first issue over one period
template <int STAGES>
__global__ void __launch_bounds__(kThreads, 1)
k_paced(uint4 *p, uint64_t n, uint32_t seed, uint32_t tile_vec, uint32_t sub_vec, uint32_t period_x16, int stagger) {
  extern __shared__ __align__(128) uint4 ring[];
  const uint32_t nsub = tile_vec / sub_vec;
  const uint64_t tiles = n / tile_vec;
  int stage = 0;
  uint64_t next_x16 = gtimer() * 16;
  if (stagger) next_x16 += (uint64_t)period_x16 * blockIdx.x / gridDim.x;
  for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const uint64_t v0 = tile * tile_vec;
    uint4 *buf = ring + (size_t)stage * tile_vec;
    if (threadIdx.x < nsub) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STAGES - 1) : "memory");
    __syncthreads();
#pragma unroll 4
    for (uint32_t j = threadIdx.x; j < tile_vec; j += kThreads) buf[j] = pat(v0 + j, seed);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < nsub) {
      if (period_x16) {
        uint64_t now;
        while ((now = gtimer() * 16) < next_x16) {}
        // behind schedule by more than a period: do not burst to catch up
        if (now > next_x16 + period_x16) next_x16 = now;
        next_x16 += period_x16;
      }
      const uint32_t first = threadIdx.x * sub_vec;
      const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf + first);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + v0 + first), "r"(saddr),
                   "r"(sub_vec * 16u) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    stage = stage + 1 == STAGES ? 0 : stage + 1;
  }
  if (threadIdx.x < nsub) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
template <typename F>
static double bench(uint64_t bytes, F launch, double *best) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  std::vector<float> ts;
  for (int r = 0; r < 6; ++r) {
    CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    if (r) ts.push_back(ms);
  }
  std::sort(ts.begin(), ts.end());
  CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b));
  if (best) *best = bytes / ts[0] / 1e6;
  return bytes / ts[ts.size() / 2] / 1e6;
}
int main(int argc, char **argv) {
  double gib = argc > 1 ? atof(argv[1]) : 8.0;
  uint64_t bytes = (uint64_t)(gib * (1ull << 30)) & ~((1ull << 21) - 1);
  uint64_t n = bytes / 16;
  uint4 *p; CK(cudaMalloc(&p, bytes));
  uint64_t *d; CK(cudaMalloc(&d, 16));
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  k_timer_res<<<1, 1>>>(d);
  uint64_t h[2]; CK(cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost));
  printf("globaltimer min step %llu ns (64 steps in %llu ns)\n", (unsigned long long)h[0], (unsigned long long)h[1]);
  const int S = 3;
  CK(cudaFuncSetAttribute(k_paced<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  const uint32_t combos[][2] = {{64, 4}, {64, 8}, {64, 16}, {64, 64}, {32, 32}, {32, 16}, {32, 8}, {16, 16}};
  const int rates[] = {0, 6600, 6800, 7000, 7100, 7200, 7300, 7400, 7500, 7600, 7800, 8200};
  printf("columns: target aggregate GB/s:");
  for (int r : rates) printf(" %6d", r);
  printf("\n");
  for (auto &c : combos)
    for (int stagger : {0, 1}) {
      const uint32_t tv = c[0] * 64, sv = c[1] * 64;
      const size_t smem = (size_t)tv * 16 * S;
      printf("tile=%2uKiB sub=%2uKiB stagger=%d :", c[0], c[1], stagger);
      for (int rate : rates) {
        uint32_t per = rate ? (uint32_t)((double)tv * 16 * sms / rate * 16.0) : 0;  // ns*16 (GB/s == B/ns)
        double g = bench(bytes, [&] { k_paced<S><<<sms, kThreads, smem>>>(p, n, 7, tv, sv, per, stagger); }, nullptr);
        printf(" %6.0f", g);
      }
      printf("\n");
      fflush(stdout);
    }
  // size sensitivity at the best-looking setting
  for (double g2 : {0.25, 1.0, 4.0}) {
    uint64_t b2 = (uint64_t)(g2 * (1ull << 30)), n2 = b2 / 16;
    for (int rate : {0, 7200, 7300, 7400}) {
      uint32_t per = rate ? (uint32_t)((double)4096 * 16 * sms / rate * 16.0) : 0;
      double best;
      double g = bench(b2, [&] { k_paced<S><<<sms, kThreads, 3 * 65536>>>(p, n2, 7, 4096, 256, per, 1); }, &best);
      printf("size %.2f GiB tile=64 sub=4 stagger=1 rate=%d : med %6.0f best %6.0f\n", g2, rate, g, best);
    }
  }
  return 0;
}

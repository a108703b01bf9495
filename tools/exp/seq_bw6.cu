// Exploratory round 6: does a read pass slow down because it directly follows a write pass?
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)
__device__ __forceinline__ uint4 pat(uint64_t i, uint32_t seed) {
  const uint32_t lo = (uint32_t)i, hi = (uint32_t)(i >> 32);
  uint32_t h = lo * 0x9E3779B1u + (hi ^ seed) * 0x85EBCA77u + seed;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
  return make_uint4(h, ~h, h * 0x297A2D39u + lo, __funnelshift_l(h, h, 13) ^ hi ^ 0xC2B2AE3Du);
}
__global__ void k_write(uint4 *p, uint64_t n, uint32_t seed) {
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = pat(i, seed);
}
// tile order: dir=0 forward, dir=1 reverse
__global__ void k_read(const uint4 *p, uint64_t n, uint32_t seed, unsigned long long *bad, int dir) {
  const uint64_t tiles = n / 4096;
  unsigned long long nb = 0;
  for (uint64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    const uint64_t tt = dir ? tiles - 1 - t : t;
    const uint64_t v0 = tt * 4096;
    for (uint32_t k = 0; k < 4096; k += 1024) {
      uint4 g[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) g[u] = p[v0 + k + u * 256 + threadIdx.x];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint4 w = pat(v0 + k + u * 256 + threadIdx.x, seed);
        nb += (g[u].x != w.x) | (g[u].y != w.y) | (g[u].z != w.z) | (g[u].w != w.w);
      }
    }
  }
  if (nb) atomicAdd(bad, nb);
}
__global__ void k_spin(long long cycles) { long long t0 = clock64(); while (clock64() - t0 < cycles) {} }
int main(int argc, char **argv) {
  double gib = argc > 1 ? atof(argv[1]) : 8.0;
  uint64_t bytes = (uint64_t)(gib * (1ull << 30)) & ~((1ull << 21) - 1), n = bytes / 16;
  uint4 *p; CK(cudaMalloc(&p, bytes));
  unsigned long long *bad; CK(cudaMalloc(&bad, 8)); CK(cudaMemset(bad, 0, 8));
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const char *seqs[] = {"WRRR", "WWWW", "RRRR", "WrRR", "WsRR", "WSRR", "WRWR", "RWRW"};
  std::vector<cudaEvent_t> ev(16);
  for (auto &evt : ev) CK(cudaEventCreate(&evt));
  k_write<<<sms * 2, 256>>>(p, n, 7); CK(cudaDeviceSynchronize());
  for (int rep = 0; rep < 2; ++rep)
    for (const char *s : seqs) {
      int m = (int)strlen(s);
      CK(cudaDeviceSynchronize());
      for (int i = 0; i < m; ++i) {
        CK(cudaEventRecord(ev[i]));
        switch (s[i]) {
          case 'W': k_write<<<sms * 2, 256>>>(p, n, 7); break;
          case 'R': k_read<<<sms * 8, 256>>>(p, n, 7, bad, 0); break;
          case 'r': k_read<<<sms * 8, 256>>>(p, n, 7, bad, 1); break;
          case 's': k_spin<<<1, 1>>>(200000); break;    // ~0.1 ms idle
          case 'S': k_spin<<<1, 1>>>(4000000); break;   // ~2 ms idle
        }
      }
      CK(cudaEventRecord(ev[m])); CK(cudaEventSynchronize(ev[m])); CK(cudaGetLastError());
      printf("%s :", s);
      for (int i = 0; i < m; ++i) {
        float ms; CK(cudaEventElapsedTime(&ms, ev[i], ev[i + 1]));
        if (s[i] == 's' || s[i] == 'S') printf("  [idle %.2fms]", ms);
        else printf("  %c %6.0f", s[i], bytes / ms / 1e6);
      }
      printf("\n");
    }
  unsigned long long h; CK(cudaMemcpy(&h, bad, 8, cudaMemcpyDeviceToHost));
  printf("bad words: %llu\n", h);
  return 0;
}

// Exploratory round 2: bulk-store (TMA engine) design space + memset-only mode for ncu.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint4 pat(uint64_t i, uint32_t seed) {
  const uint32_t lo = (uint32_t)i, hi = (uint32_t)(i >> 32);
  uint32_t h = lo * 0x9E3779B1u + (hi ^ seed) * 0x85EBCA77u + seed;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
  return make_uint4(h, ~h, h * 0x297A2D39u + lo, __funnelshift_l(h, h, 13) ^ hi ^ 0xC2B2AE3Du);
}

template <int STAGES, bool HINT>
__global__ void k_bulk(uint4 *p, uint64_t n, uint32_t seed, uint32_t chunk_vec) {
  extern __shared__ __align__(128) uint4 sm[];
  uint64_t chunks = n / chunk_vec;
  uint64_t pol = 0;
  if (HINT) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  int stage = 0;
  for (uint64_t c = blockIdx.x; c < chunks; c += gridDim.x) {
    uint4 *buf = sm + (size_t)stage * chunk_vec;
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STAGES - 1) : "memory");
    __syncthreads();
    uint64_t v0 = c * chunk_vec;
    for (uint32_t j = threadIdx.x; j < chunk_vec; j += blockDim.x) buf[j] = pat(v0 + j, seed);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf);
      if (HINT)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(p + v0),
                     "r"(saddr), "r"(chunk_vec * 16u), "l"(pol) : "memory");
      else
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + v0), "r"(saddr),
                     "r"(chunk_vec * 16u) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    stage = (stage + 1) % STAGES;
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// several smaller bulk ops per chunk (sub-chunks issued by different warps)
template <int STAGES>
__global__ void k_bulk_split(uint4 *p, uint64_t n, uint32_t seed, uint32_t chunk_vec, uint32_t sub_vec) {
  extern __shared__ __align__(128) uint4 sm[];
  uint64_t chunks = n / chunk_vec;
  int stage = 0;
  const uint32_t nsub = chunk_vec / sub_vec;
  for (uint64_t c = blockIdx.x; c < chunks; c += gridDim.x) {
    uint4 *buf = sm + (size_t)stage * chunk_vec;
    if (threadIdx.x < nsub) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STAGES - 1) : "memory");
    __syncthreads();
    uint64_t v0 = c * chunk_vec;
    for (uint32_t j = threadIdx.x; j < chunk_vec; j += blockDim.x) buf[j] = pat(v0 + j, seed);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < nsub) {
      uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf + threadIdx.x * sub_vec);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + v0 + threadIdx.x * sub_vec),
                   "r"(saddr), "r"(sub_vec * 16u) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    stage = (stage + 1) % STAGES;
  }
  if (threadIdx.x < nsub) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int MODE>
__global__ void k_st(uint4 *p, uint64_t n, uint32_t seed) {
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint4 v = pat(i, seed);
    if (MODE == 0) asm volatile("st.global.wt.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p + i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (MODE == 1) asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p + i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (MODE == 2) asm volatile("st.global.L1::evict_first.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p + i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
}

template <typename F>
static double bench(const char *name, uint64_t bytes, F launch) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  std::vector<float> ts;
  for (int r = 0; r < 5; ++r) {
    CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    if (r) ts.push_back(ms);
  }
  std::sort(ts.begin(), ts.end());
  double med = bytes / ts[ts.size() / 2] / 1e6;
  printf("%-60s best %8.1f  med %8.1f GB/s\n", name, bytes / ts[0] / 1e6, med);
  fflush(stdout);
  CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b));
  return med;
}

int main(int argc, char **argv) {
  double gib = argc > 1 ? atof(argv[1]) : 8.0;
  const char *mode = argc > 2 ? argv[2] : "sweep";
  uint64_t bytes = (uint64_t)(gib * (1ull << 30)) & ~((1ull << 21) - 1);
  uint64_t n = bytes / 16;
  uint4 *p; CK(cudaMalloc(&p, bytes));
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  if (!strcmp(mode, "memset")) {
    for (int i = 0; i < 3; ++i) CK(cudaMemsetAsync(p, 0x5a, bytes));
    CK(cudaDeviceSynchronize());
    printf("memset done\n");
    return 0;
  }
  char name[160];
  const int SM_MAX = 220 * 1024;
  CK(cudaFuncSetAttribute(k_bulk<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_MAX));
  CK(cudaFuncSetAttribute(k_bulk<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_MAX));
  CK(cudaFuncSetAttribute(k_bulk<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_MAX));
  CK(cudaFuncSetAttribute(k_bulk<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_MAX));
  CK(cudaFuncSetAttribute(k_bulk_split<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_MAX));
  CK(cudaFuncSetAttribute(k_bulk_split<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_MAX));
  bench("cudaMemsetAsync", bytes, [&] { CK(cudaMemsetAsync(p, 0x5a, bytes)); });
  for (int k : {1, 2, 4}) {
    snprintf(name, sizeof name, "st.wt ctas/sm=%d", k); bench(name, bytes, [&] { k_st<0><<<sms * k, 256>>>(p, n, 7); });
    snprintf(name, sizeof name, "st.cg ctas/sm=%d", k); bench(name, bytes, [&] { k_st<1><<<sms * k, 256>>>(p, n, 7); });
    snprintf(name, sizeof name, "st.L1::evict_first ctas/sm=%d", k); bench(name, bytes, [&] { k_st<2><<<sms * k, 256>>>(p, n, 7); });
  }
  for (int k : {1, 2})
    for (int th : {128, 256, 512, 1024})
      for (uint32_t kib : {8u, 12u, 16u, 20u, 24u, 32u, 48u, 64u}) {
        uint32_t cv = kib * 64;
        for (int st : {2, 3, 4}) {
          size_t smem = (size_t)cv * 16 * st;
          if (smem * k > (size_t)SM_MAX) continue;
          snprintf(name, sizeof name, "bulk ctas/sm=%d th=%4d chunk=%2uKiB S=%d", k, th, kib, st);
          if (st == 2) bench(name, bytes, [&] { k_bulk<2, false><<<sms * k, th, smem>>>(p, n, 7, cv); });
          if (st == 3) bench(name, bytes, [&] { k_bulk<3, false><<<sms * k, th, smem>>>(p, n, 7, cv); });
          if (st == 4) bench(name, bytes, [&] { k_bulk<4, false><<<sms * k, th, smem>>>(p, n, 7, cv); });
        }
        size_t smem2 = (size_t)cv * 16 * 2;
        if (smem2 * k <= (size_t)SM_MAX && th == 256) {
          snprintf(name, sizeof name, "bulk+evict_first ctas/sm=%d th=%4d chunk=%2uKiB S=2", k, th, kib);
          bench(name, bytes, [&] { k_bulk<2, true><<<sms * k, th, smem2>>>(p, n, 7, cv); });
        }
      }
  for (uint32_t kib : {16u, 32u, 64u})
    for (uint32_t sub : {2u, 4u, 8u}) {
      uint32_t cv = kib * 64, sv = sub * 64;
      for (int st : {2, 3}) {
        size_t smem = (size_t)cv * 16 * st;
        if (smem > (size_t)SM_MAX) continue;
        snprintf(name, sizeof name, "bulk_split ctas/sm=1 th=256 chunk=%2uKiB sub=%uKiB S=%d", kib, sub, st);
        if (st == 2) bench(name, bytes, [&] { k_bulk_split<2><<<sms, 256, smem>>>(p, n, 7, cv, sv); });
        if (st == 3) bench(name, bytes, [&] { k_bulk_split<3><<<sms, 256, smem>>>(p, n, 7, cv, sv); });
      }
    }
  return 0;
}

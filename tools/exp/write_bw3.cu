// Exploratory round 3: why is the in-library bulk kernel slower than k_bulk_split?
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)
constexpr int kThreads = 256;
constexpr uint32_t kTileVec = 4096;
struct SliceDesc { uint4 *base; uint64_t n_vec, tile_begin, tile_end; uint32_t seed, pad; };
struct Batch { SliceDesc s[64]; uint64_t total_tiles; uint32_t n, pad; };
__device__ __forceinline__ uint4 pat(uint64_t i, uint32_t seed) {
  const uint32_t lo = (uint32_t)i, hi = (uint32_t)(i >> 32);
  uint32_t h = lo * 0x9E3779B1u + (hi ^ seed) * 0x85EBCA77u + seed;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
  return make_uint4(h, ~h, h * 0x297A2D39u + lo, __funnelshift_l(h, h, 13) ^ hi ^ 0xC2B2AE3Du);
}
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
// the library kernel, verbatim; MODE bit0: drop minBlocks; bit1: no unroll pragma
template <int STAGES, int MODE>
__global__ void __launch_bounds__(kThreads, 1) k_lib(const __grid_constant__ Batch b, uint32_t sub_vec) {
  extern __shared__ __align__(128) uint4 ring[];
  const uint32_t nsub = kTileVec / sub_vec;
  uint32_t cur = 0;
  int stage = 0;
  for (uint64_t tile = blockIdx.x; tile < b.total_tiles; tile += gridDim.x) {
    while (tile >= b.s[cur].tile_end) ++cur;
    const uint64_t v0 = (tile - b.s[cur].tile_begin) * kTileVec;
    const uint64_t rem64 = b.s[cur].n_vec - v0;
    const uint32_t rem = rem64 >= kTileVec ? kTileVec : (uint32_t)rem64;
    const uint32_t seed = b.s[cur].seed;
    uint4 *buf = ring + (size_t)stage * kTileVec;
    if (threadIdx.x < nsub) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STAGES - 1) : "memory");
    __syncthreads();
    if (rem == kTileVec) {
      if (MODE & 2) {
#pragma unroll 1
        for (uint32_t j = threadIdx.x; j < kTileVec; j += kThreads) buf[j] = pat(v0 + j, seed);
      } else {
#pragma unroll 4
        for (uint32_t j = threadIdx.x; j < kTileVec; j += kThreads) buf[j] = pat(v0 + j, seed);
      }
    } else {
      for (uint32_t j = threadIdx.x; j < rem; j += kThreads) buf[j] = pat(v0 + j, seed);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < nsub) {
      const uint32_t first = threadIdx.x * sub_vec;
      if ((MODE & 4) || first < rem) {
        const uint32_t cnt = (MODE & 4) ? sub_vec : (rem - first < sub_vec ? rem - first : sub_vec);
        const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf + first);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(b.s[cur].base + v0 + first),
                     "r"(saddr), "r"(cnt * 16u) : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    stage = stage + 1 == STAGES ? 0 : stage + 1;
  }
  if (threadIdx.x < nsub) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__global__ void k_stg(uint4 *p, uint64_t n, uint32_t seed) {
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = pat(i, seed);
}
template <typename F>
static double bench(const char *name, uint64_t bytes, F launch, cudaStream_t st) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  std::vector<float> ts;
  for (int r = 0; r < 6; ++r) {
    CK(cudaEventRecord(a, st)); launch(); CK(cudaEventRecord(b, st)); CK(cudaEventSynchronize(b));
    CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    if (r) ts.push_back(ms);
  }
  std::sort(ts.begin(), ts.end());
  double med = bytes / ts[ts.size() / 2] / 1e6;
  printf("%-60s best %8.1f  med %8.1f GB/s\n", name, bytes / ts[0] / 1e6, med);
  fflush(stdout);
  return med;
}
int main(int argc, char **argv) {
  double gib = argc > 1 ? atof(argv[1]) : 8.0;
  uint64_t bytes = (uint64_t)(gib * (1ull << 30)) & ~((1ull << 21) - 1);
  uint64_t n = bytes / 16;
  uint4 *p; CK(cudaMalloc(&p, bytes));
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaStream_t st; CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  Batch b; memset(&b, 0, sizeof b);
  b.s[0].base = p; b.s[0].n_vec = n; b.s[0].tile_begin = 0; b.s[0].tile_end = n / kTileVec; b.s[0].seed = 7;
  b.total_tiles = n / kTileVec; b.n = 1;
  const size_t smem = 3 * 65536;
  if (argc > 2 && !strcmp(argv[2], "prof")) {
    CK(cudaFuncSetAttribute(k_bulk_split<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(k_lib<3, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(k_lib<3, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(k_lib<3, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int r = 0; r < 2; ++r) {
      k_stg<<<sms * 2, 256>>>(p, n, 7);
      k_lib<3, 0><<<sms, 256, smem>>>(b, 256);
      k_lib<3, 4><<<sms, 256, smem>>>(b, 256);
      k_lib<3, 6><<<sms, 256, smem>>>(b, 256);
      k_bulk_split<3><<<sms, 256, smem>>>(p, n, 7, 4096, 256);
    }
    CK(cudaDeviceSynchronize());
    printf("prof done\n");
    return 0;
  }
  for (int big : {0, 1}) {
    int attr = big ? 220 * 1024 : (int)smem;
    CK(cudaFuncSetAttribute(k_bulk_split<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, attr));
    CK(cudaFuncSetAttribute(k_lib<3, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, attr));
    CK(cudaFuncSetAttribute(k_lib<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, attr));
    CK(cudaFuncSetAttribute(k_lib<3, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, attr));
    CK(cudaFuncSetAttribute(k_lib<3, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, attr));
    printf("-- max dyn smem attr %d\n", attr);
    for (int rep = 0; rep < 2; ++rep) {
      bench("k_bulk_split<3> default stream", bytes, [&] { k_bulk_split<3><<<sms, 256, smem>>>(p, n, 7, 4096, 256); }, 0);
      bench("k_bulk_split<3> nonblocking stream", bytes, [&] { k_bulk_split<3><<<sms, 256, smem, st>>>(p, n, 7, 4096, 256); }, st);
      bench("k_lib<3,0> (verbatim)", bytes, [&] { k_lib<3, 0><<<sms, 256, smem, st>>>(b, 256); }, st);
      bench("k_lib<3,2> (no unroll)", bytes, [&] { k_lib<3, 2><<<sms, 256, smem, st>>>(b, 256); }, st);
      bench("k_lib<3,4> (no tail logic)", bytes, [&] { k_lib<3, 4><<<sms, 256, smem, st>>>(b, 256); }, st);
      bench("k_lib<3,6> (no unroll, no tail)", bytes, [&] { k_lib<3, 6><<<sms, 256, smem, st>>>(b, 256); }, st);
    }
  }
  return 0;
}

// Exploratory: which store form fills HBM fastest on B200? (not part of the library)
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint4 pat(uint64_t i, uint32_t seed) {
  const uint32_t lo = (uint32_t)i, hi = (uint32_t)(i >> 32);
  uint32_t h = lo * 0x9E3779B1u + (hi ^ seed) * 0x85EBCA77u + seed;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
  return make_uint4(h, ~h, h * 0x297A2D39u + lo, __funnelshift_l(h, h, 13) ^ hi ^ 0xC2B2AE3Du);
}

// K1: grid-stride, thread-contiguous
template <bool PAT>
__global__ void k_gridstride(uint4 *p, uint64_t n, uint32_t seed) {
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    p[i] = PAT ? pat(i, seed) : make_uint4(seed, seed, seed, seed);
}

// K5: tile-contiguous per CTA, tile_vec runtime
__global__ void k_tile(uint4 *p, uint64_t n, uint32_t seed, uint32_t tile_vec) {
  uint64_t tiles = n / tile_vec;
  for (uint64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    uint64_t v0 = t * tile_vec;
    for (uint32_t k = threadIdx.x; k < tile_vec; k += blockDim.x * 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint32_t j = k + u * blockDim.x;
        if (j < tile_vec) p[v0 + j] = pat(v0 + j, seed);
      }
    }
  }
}

// K3: 256-bit stores
__global__ void k_v8(uint4 *p, uint64_t n, uint32_t seed) {
  uint64_t n2 = n / 2;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
    uint4 a = pat(2 * i, seed), b = pat(2 * i + 1, seed);
    asm volatile("st.global.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p + 2 * i), "r"(a.x), "r"(a.y),
                 "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w) : "memory");
  }
}

// K4: pattern -> smem -> bulk async store (TMA engine, non-tensor form)
template <int STAGES>
__global__ void k_bulk(uint4 *p, uint64_t n, uint32_t seed, uint32_t chunk_vec) {
  extern __shared__ __align__(128) uint4 sm[];
  uint64_t chunks = n / chunk_vec;
  int stage = 0;
  for (uint64_t c = blockIdx.x; c < chunks; c += gridDim.x) {
    uint4 *buf = sm + (size_t)stage * chunk_vec;
    // make sure the bulk store that last read this stage has finished reading smem
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STAGES - 1) : "memory");
    __syncthreads();
    uint64_t v0 = c * chunk_vec;
    for (uint32_t j = threadIdx.x; j < chunk_vec; j += blockDim.x) buf[j] = pat(v0 + j, seed);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + v0), "r"(saddr),
                   "r"(chunk_vec * 16u) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    stage = (stage + 1) % STAGES;
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// read kernel for completeness (sum to defeat DCE)
__global__ void k_read(const uint4 *p, uint64_t n, uint32_t *out) {
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint32_t acc = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint4 v = p[i];
    acc ^= v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0x12345678u) *out = acc;
}

template <typename F>
static void bench(const char *name, uint64_t bytes, F launch) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  std::vector<float> ts;
  for (int r = 0; r < 6; ++r) {
    CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    if (r) ts.push_back(ms);
  }
  std::sort(ts.begin(), ts.end());
  printf("%-44s best %8.1f  med %8.1f GB/s\n", name, bytes / ts[0] / 1e6, bytes / ts[ts.size() / 2] / 1e6);
  fflush(stdout);
}

int main(int argc, char **argv) {
  double gib = argc > 1 ? atof(argv[1]) : 8.0;
  uint64_t bytes = (uint64_t)(gib * (1ull << 30)) & ~((1ull << 21) - 1);
  uint64_t n = bytes / 16;
  uint4 *p; CK(cudaMalloc(&p, bytes));
  uint32_t *out; CK(cudaMalloc(&out, 4));
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  printf("SMs %d bytes %llu\n", sms, (unsigned long long)bytes);
  char name[128];
  bench("cudaMemsetAsync", bytes, [&] { CK(cudaMemsetAsync(p, 0x5a, bytes)); });
  for (int th : {256, 512, 1024})
    for (int k : {1, 2, 4, 8}) {
      if (th * k > 2048) continue;
      snprintf(name, sizeof name, "gridstride pat   th=%d ctas/sm=%d", th, k);
      bench(name, bytes, [&] { k_gridstride<true><<<sms * k, th>>>(p, n, 7); });
      snprintf(name, sizeof name, "gridstride const th=%d ctas/sm=%d", th, k);
      bench(name, bytes, [&] { k_gridstride<false><<<sms * k, th>>>(p, n, 7); });
    }
  for (int k : {1, 2, 4}) {
    snprintf(name, sizeof name, "v8 256-bit th=256 ctas/sm=%d", k);
    bench(name, bytes, [&] { k_v8<<<sms * k, 256>>>(p, n, 7); });
  }
  for (uint32_t tv : {1024u, 4096u, 16384u, 65536u})
    for (int k : {1, 2, 4}) {
      snprintf(name, sizeof name, "tile %4u KiB th=256 ctas/sm=%d", tv * 16 / 1024, k);
      bench(name, bytes, [&] { k_tile<<<sms * k, 256>>>(p, n, 7, tv); });
    }
  CK(cudaFuncSetAttribute(k_bulk<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CK(cudaFuncSetAttribute(k_bulk<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  for (uint32_t cv : {512u, 1024u, 2048u, 4096u})
    for (int k : {1, 2, 4}) {
      size_t sm2 = (size_t)cv * 16 * 2, sm4 = (size_t)cv * 16 * 4;
      if (sm2 * k <= 200 * 1024) {
        snprintf(name, sizeof name, "bulk S=2 chunk %3u KiB th=256 ctas/sm=%d", cv * 16 / 1024, k);
        bench(name, bytes, [&] { k_bulk<2><<<sms * k, 256, sm2>>>(p, n, 7, cv); });
      }
      if (sm4 * k <= 200 * 1024) {
        snprintf(name, sizeof name, "bulk S=4 chunk %3u KiB th=256 ctas/sm=%d", cv * 16 / 1024, k);
        bench(name, bytes, [&] { k_bulk<4><<<sms * k, 256, sm4>>>(p, n, 7, cv); });
      }
    }
  for (int k : {2, 4, 8}) {
    snprintf(name, sizeof name, "read gridstride th=256 ctas/sm=%d", k);
    bench(name, bytes, [&] { k_read<<<sms * k, 256>>>(p, n, out); });
  }
  return 0;
}

// Exploratory round 9: does the DATA written change the sustained (power-capped) store rate?
// Same paced STG.128 kernel, same grid, five patterns of very different bit entropy, each run
// back to back for several seconds so the board settles at its power cap. If all-zero words run
// markedly faster than hashed words under the cap, HBM I/O toggling is a lever; if not, the
// pattern is free to be as random as the fault model wants.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)
__device__ __forceinline__ uint64_t gtimer() { uint64_t t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

template <int KIND> __device__ __forceinline__ uint4 word(uint64_t i, uint32_t seed) {
  const uint32_t lo = (uint32_t)i, hi = (uint32_t)(i >> 32);
  if (KIND == 0) return make_uint4(0, 0, 0, 0);                                   // all zero
  if (KIND == 1) {                                                                // the probe's hash
    uint32_t h = lo * 0x9E3779B1u + (hi ^ seed) * 0x85EBCA77u + seed; h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    return make_uint4(h, ~h, h * 0x297A2D39u + lo, __funnelshift_l(h, h, 13) ^ hi ^ 0xC2B2AE3Du);
  }
  if (KIND == 2) return make_uint4(lo, hi, lo, hi);                               // the address: few bits change
  if (KIND == 3) return (lo & 1) ? make_uint4(~0u, ~0u, ~0u, ~0u) : make_uint4(0, 0, 0, 0);  // 16 B of 0 / 16 B of 1
  if (KIND == 4) return make_uint4(0x55555555u, 0xAAAAAAAAu, 0x55555555u, 0xAAAAAAAAu);  // checkerboard, constant
  if (KIND == 5) {  // per-sweep random base XOR the address: every cell gets a random bit, few bits change per word
    uint32_t s0 = seed * 0x9E3779B1u + 0x7F4A7C15u; s0 ^= s0 >> 15; s0 *= 0x2C1B3C6Du; s0 ^= s0 >> 12;
    const uint32_t s1 = s0 * 0x85EBCA77u + 1, s2 = s1 * 0xC2B2AE3Du + 1, s3 = s2 * 0x27D4EB2Fu + 1;
    return make_uint4(s0 ^ lo, s1 ^ ~lo, s2 ^ hi, s3 ^ __funnelshift_l(lo, lo, 16));
  }
  if (KIND == 7) {  // one random word per sweep; lanes are +-(R ^ small address terms): adjacent 32-bit bus beats
                    // are near-equal or near-complementary (free under data-bus inversion)
    uint32_t r = seed * 0x9E3779B1u + 0x7F4A7C15u; r ^= r >> 15; r *= 0x2C1B3C6Du; r ^= r >> 12;
    const uint32_t x = r ^ lo, z = r ^ hi ^ __funnelshift_l(lo, lo, 16);
    return make_uint4(x, ~x, z, ~z);
  }
  if (KIND == 8) {  // per-sweep random background, the address in one lane only
    uint32_t r = seed * 0x9E3779B1u + 0x7F4A7C15u; r ^= r >> 15; r *= 0x2C1B3C6Du; r ^= r >> 12;
    const uint32_t r1 = r * 0x85EBCA77u + 1, r2 = r1 * 0xC2B2AE3Du + 1, r3 = r2 * 0x27D4EB2Fu + 1;
    return make_uint4(r ^ lo, r1 ^ hi, r2, r3);
  }
  {  // KIND 6: per-64-KiB-tile random base XOR the offset in the tile
    const uint32_t tile = (uint32_t)(i >> 12), j = lo & 0xFFFu;
    uint32_t h = tile * 0x9E3779B1u + seed; h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    return make_uint4(h ^ j, ~h, h * 0x297A2D39u, __funnelshift_l(h, h, 13) ^ (j << 16));
  }
}

template <int KIND>
__global__ void __launch_bounds__(256) k_write(uint4 *p, uint64_t n, uint32_t seed, uint32_t tile_vec, uint32_t period_x16) {
  const uint64_t tiles = n / tile_vec;
  const uint64_t t0 = gtimer();
  uint64_t next = (uint64_t)period_x16 * blockIdx.x / gridDim.x;
  for (uint64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    if (period_x16 && threadIdx.x < 32) {
      uint64_t now;
      while ((now = (gtimer() - t0) * 16) < next) __nanosleep(32);
      if (now > next + period_x16) next = now;
      next += period_x16;
    }
    __syncthreads();
    const uint64_t v0 = t * tile_vec;
#pragma unroll 4
    for (uint32_t j = threadIdx.x; j < tile_vec; j += 256) p[v0 + j] = word<KIND>(v0 + j, seed);
  }
}

// plain loads (the verify pass without the compare): what the read side does under the same cap
__global__ void __launch_bounds__(256) k_read(const uint4 *p, uint64_t n, uint32_t *sink) {
  uint32_t acc = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (uint64_t)gridDim.x * 256) {
    const uint4 v = __ldcs(p + i);
    acc ^= v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0x12345678u) *sink = acc;
}

template <int KIND> static void run(const char *name, uint4 *buf, uint64_t n, int sms, double seconds, uint32_t *sink) {
  const uint32_t tile_vec = 1024;  // 16 KiB per token
  const uint32_t period_x16 = (uint32_t)((double)tile_vec * 16.0 * sms / 7600.0 * 16.0);
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  std::vector<float> wr, rd;
  double total = 0;
  while (total < seconds * 1e3) {
    CK(cudaEventRecord(a));
    k_write<KIND><<<sms, 256>>>(buf, n, 20261106u, tile_vec, period_x16);
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); wr.push_back(ms); total += ms;
    CK(cudaEventRecord(a));
    k_read<<<sms * 8, 256>>>(buf, n, sink);
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    CK(cudaEventElapsedTime(&ms, a, b)); rd.push_back(ms); total += ms;
  }
  CK(cudaGetLastError());
  auto mean = [](const std::vector<float> &v, size_t from, size_t to) { double s = 0; for (size_t i = from; i < to; ++i) s += v[i]; return s / (double)(to - from); };
  const double gb = (double)n * 16 / 1e9;
  const size_t k = wr.size(), head = k < 6 ? k : 6, tail0 = k > 20 ? k - k / 3 : 0;
  printf("%-24s passes %3zu | store first %zu: %6.0f GB/s  last third: %6.0f GB/s | load first: %6.0f  last third: %6.0f GB/s\n",
         name, k, head, gb / mean(wr, 0, head) * 1e3, gb / mean(wr, tail0, k) * 1e3, gb / mean(rd, 0, head) * 1e3,
         gb / mean(rd, tail0, k) * 1e3);
  fflush(stdout);
}

int main(int argc, char **argv) {
  const double seconds = argc > 1 ? atof(argv[1]) : 4.0;
  const uint64_t bytes = (argc > 2 ? (uint64_t)atoll(argv[2]) : 64ull) << 30;
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  uint4 *buf; CK(cudaMalloc(&buf, bytes));
  uint32_t *sink; CK(cudaMalloc(&sink, 4));
  const uint64_t n = bytes / 16;
  printf("%s, %d SMs, %.0f GiB per pass, %.1f s per pattern (store pass + load pass alternating)\n", prop.name,
         prop.multiProcessorCount, bytes / 1073741824.0, seconds);
  run<1>("hash (first version)", buf, n, prop.multiProcessorCount, seconds, sink);
  run<0>("all zero", buf, n, prop.multiProcessorCount, seconds, sink);
  run<2>("address", buf, n, prop.multiProcessorCount, seconds, sink);
  run<4>("0x55 / 0xAA constant", buf, n, prop.multiProcessorCount, seconds, sink);
  run<5>("seed base ^ address", buf, n, prop.multiProcessorCount, seconds, sink);
  run<7>("+-(R ^ address)", buf, n, prop.multiProcessorCount, seconds, sink);
  run<8>("background+addr (probe)", buf, n, prop.multiProcessorCount, seconds, sink);
  run<1>("hash again", buf, n, prop.multiProcessorCount, seconds, sink);
  return 0;
}

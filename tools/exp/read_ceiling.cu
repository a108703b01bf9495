// Exploratory round 10: what bounds the load+compare pass? Runs the product's own kernels
// (csrc/probe_kernels.cuh) over one large slice, each form sustained for `secs` seconds:
//   - the LDG kernel as shipped (U = 2 / 4 / 8, 8 / 4 / 2 CTAs per SM)
//   - the bulk-load ring with its compare                       (what the product would run)
//   - the bulk-load ring with the compare switched off (COMPARE = false: consumers only release
//     the stages) -- the rate at which the TMA engine alone pulls the slice through the SMs,
//     i.e. the ceiling no compare kernel on this path can exceed
// and prints the median rate of the second half of each run with a power / SM-clock sample
// taken while the launches are queued.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iinclude \
//        -Ialiyun-gpushare-device-plugin_b200/csrc tools/exp/read_ceiling.cu -o tools/exp/read_ceiling.bin
//   tools/exp/read_ceiling.bin [GiB=32] [secs=1.5]
#include <algorithm>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "probe_kernels.cuh"

using namespace gs_probe_dev;

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e = (x);                                                              \
    if (e != cudaSuccess) {                                                           \
      printf("CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e));             \
      exit(1);                                                                        \
    }                                                                                 \
  } while (0)

static std::string smi_sample() {
  FILE *f = popen("nvidia-smi -i 0 --query-gpu=power.draw,clocks.sm,clocks_event_reasons.sw_power_cap "
                  "--format=csv,noheader,nounits 2>/dev/null", "r");
  if (!f) return "?";
  char buf[256] = {0};
  if (!fgets(buf, sizeof buf, f)) buf[0] = 0;
  pclose(f);
  std::string s(buf);
  while (!s.empty() && (s.back() == '\n' || s.back() == ' ')) s.pop_back();
  return s;
}

struct Form {
  const char *name;
  const void *fn;
  int threads;
  size_t smem;
  int grid_per_sm;
  uint32_t stages;  // 0: LDG kernel (no third argument)
};

int main(int argc, char **argv) {
  const double gib = argc > 1 ? atof(argv[1]) : 32.0;
  const double secs = argc > 2 ? atof(argv[2]) : 1.5;
  const uint64_t bytes = (uint64_t)(gib * (1ull << 30)) & ~((1ull << 21) - 1);
  uint4 *p;
  CK(cudaMalloc(&p, bytes));
  DevResult *res;
  CK(cudaMalloc(&res, sizeof(DevResult) * GS_PROBE_MAX_SLICES));
  CK(cudaMemset(res, 0, sizeof(DevResult) * GS_PROBE_MAX_SLICES));
  int sms;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));

  Batch b;
  memset(&b, 0, sizeof b);
  b.s[0].base = p;
  b.s[0].n_vec = bytes / 16;
  b.s[0].tile_begin = 0;
  b.s[0].tile_end = (b.s[0].n_vec + kTileVec - 1) / kTileVec;
  b.s[0].seed = 77;
  b.s[0].pattern = GS_PROBE_PATTERN_LOWTOGGLE;
  b.total_tiles = b.s[0].tile_end;
  b.n = 1;
  probe_write_kernel<4, 0><<<sms * 4, kThreads>>>(b);
  CK(cudaDeviceSynchronize());

  const size_t ring = 192u << 10;
#define LDG(U, per) {"ldg U=" #U " x" #per "/SM", (const void *)probe_verify_kernel<U, 0>, kThreads, 0, per, 0}
#define BULK(C, W, H, CMP, per)                                                                    \
  {"bulk " #C "w x" #W "warps hint=" #H " compare=" #CMP " x" #per "/SM",                           \
   (const void *)probe_verify_bulk_kernel<C, W, H, CMP>, 32 + 32 * W, ring / per, per,             \
   (uint32_t)(ring / per / (C * 16))}
  std::vector<Form> forms = {
      LDG(2, 8), LDG(4, 5), LDG(4, 4), LDG(8, 3), LDG(8, 2),
      BULK(1024, 8, false, false, 1), BULK(2048, 8, false, false, 1), BULK(4096, 8, false, false, 1),
      BULK(1024, 8, true, false, 1),  BULK(4096, 8, true, false, 1),  BULK(2048, 8, false, false, 2),
      BULK(1024, 4, false, true, 1),  BULK(1024, 8, false, true, 1),  BULK(1024, 16, false, true, 1),
      BULK(2048, 4, false, true, 1),  BULK(2048, 8, false, true, 1),  BULK(2048, 16, false, true, 1),
      BULK(4096, 4, false, true, 1),  BULK(4096, 8, false, true, 1),  BULK(4096, 16, false, true, 1),
      BULK(2048, 8, true, true, 1),   BULK(2048, 8, false, true, 2),  BULK(1024, 4, false, true, 2),
  };
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int rep = 0; rep < 2; ++rep)
    for (const Form &f : forms) {
      if (f.smem) CK(cudaFuncSetAttribute(f.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f.smem));
      void *args_ldg[] = {(void *)&b, (void *)&res};
      uint32_t stages = f.stages > kBulkMaxStages ? kBulkMaxStages : f.stages;
      void *args_bulk[] = {(void *)&b, (void *)&res, (void *)&stages};
      const size_t smem = f.stages ? (size_t)stages * (f.smem / f.stages) : 0;
      std::vector<float> ms;
      std::string smi;
      cudaEvent_t t0, t1;
      CK(cudaEventCreate(&t0));
      CK(cudaEventCreate(&t1));
      CK(cudaEventRecord(t0));
      float total = 0.f;
      int iter = 0;
      while (total < secs * 1e3f) {
        CK(cudaEventRecord(e0));
        CK(cudaLaunchKernel(f.fn, dim3(sms * f.grid_per_sm), dim3(f.threads), f.stages ? args_bulk : args_ldg,
                            smem, 0));
        CK(cudaEventRecord(e1));
        if (iter == 40) smi = smi_sample();  // while the GPU is busy with this form
        CK(cudaEventSynchronize(e1));
        float m;
        CK(cudaEventElapsedTime(&m, e0, e1));
        ms.push_back(m);
        CK(cudaEventRecord(t1));
        CK(cudaEventSynchronize(t1));
        CK(cudaEventElapsedTime(&total, t0, t1));
        ++iter;
      }
      std::vector<float> tail(ms.begin() + ms.size() / 2, ms.end());
      std::sort(tail.begin(), tail.end());
      const float med = tail[tail.size() / 2];
      printf("rep %d  %-52s first %6.0f  median %6.0f  best %6.0f GB/s | %s\n", rep, f.name,
             bytes / ms[0] / 1e6, bytes / med / 1e6, bytes / tail[0] / 1e6, smi.c_str());
      fflush(stdout);
    }
  DevResult h;
  CK(cudaMemcpy(&h, res, sizeof h, cudaMemcpyDeviceToHost));
  printf("mismatches over all runs: %llu (must be 0)\n", h.mismatches);
  return h.mismatches != 0;
}

// cuda_util.cu -- NVML index -> CUDA ordinal by PCI bus id. The reference hands
// NVIDIA_VISIBLE_DEVICES=<NVML index> to the container runtime and never runs CUDA
// itself; the probe does, and CUDA's enumeration order need not match NVML's.
#include <cuda_runtime.h>

#include "gpushare.h"
#include "internal.h"

extern "C" int gs_cuda_map_ordinals(gs_gpu_info *gpus, uint32_t n) {
  if (!gpus && n) return GS_E_INVAL;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess) {
    gs_internal::set_cuda_error("cudaGetDeviceCount", (int)e, cudaGetErrorString(e));
    return GS_E_CUDA;
  }
  for (uint32_t i = 0; i < n; ++i) {
    gpus[i].cuda_ordinal = -1;
    if (!gpus[i].pci_bus_id[0]) continue;
    int dev = -1;
    e = cudaDeviceGetByPCIBusId(&dev, gpus[i].pci_bus_id);
    if (e == cudaSuccess && dev >= 0 && dev < count) gpus[i].cuda_ordinal = dev;
    else (void)cudaGetLastError();  // hidden by CUDA_VISIBLE_DEVICES: leave -1
  }
  return GS_OK;
}

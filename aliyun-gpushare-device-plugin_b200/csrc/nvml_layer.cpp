// nvml_layer.cpp -- thin layer over NVML, bound at run time with dlopen.
//
// Replaces the go-nvml v0.12.0-1 binding (go.mod:6) at the reference's call sites:
//   nvml.Init / nvml.Shutdown            pkg/gpu/nvidia/gpumanager.go:36,41
//   nvml.DeviceGetCount                  nvidia.go:45,51
//   nvml.DeviceGetHandleByIndex          nvidia.go:57
//   Device.GetUUID                       nvidia.go:60
//   Device.GetMemoryInfo_v2              nvidia.go:64
//   nvml.ErrorString                     nvidia.go:19
// and implements the health watch the fork left commented out (nvidia.go:97-153):
//   EventSetCreate / RegisterEventForDevice(XidCriticalError) / WaitForEvent(5000).
// dlopen keeps libgpushare_core.so loadable on machines without a driver and lets
// tests point it at tests/fake_nvml/libfake_nvml.so (BASELINE configs 1 and 5).
#include <dlfcn.h>
#include <nvml.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "gpushare.h"
#include "internal.h"

namespace {

struct Api {
  void *handle = nullptr;
  nvmlReturn_t (*Init)(void) = nullptr;
  nvmlReturn_t (*Shutdown)(void) = nullptr;
  const char *(*ErrorString)(nvmlReturn_t) = nullptr;
  nvmlReturn_t (*DeviceGetCount)(unsigned int *) = nullptr;
  nvmlReturn_t (*DeviceGetHandleByIndex)(unsigned int, nvmlDevice_t *) = nullptr;
  nvmlReturn_t (*DeviceGetUUID)(nvmlDevice_t, char *, unsigned int) = nullptr;
  nvmlReturn_t (*DeviceGetMemoryInfo_v2)(nvmlDevice_t, nvmlMemory_v2_t *) = nullptr;
  nvmlReturn_t (*DeviceGetMemoryInfo)(nvmlDevice_t, nvmlMemory_t *) = nullptr;
  nvmlReturn_t (*DeviceGetPciInfo)(nvmlDevice_t, nvmlPciInfo_t *) = nullptr;
  nvmlReturn_t (*EventSetCreate)(nvmlEventSet_t *) = nullptr;
  nvmlReturn_t (*DeviceRegisterEvents)(nvmlDevice_t, unsigned long long, nvmlEventSet_t) = nullptr;
  nvmlReturn_t (*EventSetWait)(nvmlEventSet_t, nvmlEventData_t *, unsigned int) = nullptr;
  nvmlReturn_t (*EventSetFree)(nvmlEventSet_t) = nullptr;
};

std::mutex g_mu;
Api g_api;
bool g_open = false;
nvmlEventSet_t g_events = nullptr;
bool g_events_open = false;
std::vector<nvmlDevice_t> g_handles;  // by NVML index, filled by gs_health_open

int fail(const char *what, nvmlReturn_t r) {
  const char *msg = g_api.ErrorString ? g_api.ErrorString(r) : "nvml error";
  gs_internal::set_nvml_error(what, (int)r, msg);
  return GS_E_NVML;
}

template <class F>
bool sym(F &fn, const char *name) {
  fn = reinterpret_cast<F>(dlsym(g_api.handle, name));
  return fn != nullptr;
}

}  // namespace

extern "C" {

int gs_nvml_open(const char *lib_path) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_open) return GS_E_STATE;
  const char *path = (lib_path && *lib_path) ? lib_path : "libnvidia-ml.so.1";
  g_api = Api();
  g_api.handle = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!g_api.handle) {
    gs_internal::set_nvml_error("dlopen", 0, dlerror());
    return GS_E_NOLIB;
  }
  bool ok = sym(g_api.Init, "nvmlInit_v2") && sym(g_api.Shutdown, "nvmlShutdown") &&
            sym(g_api.ErrorString, "nvmlErrorString") &&
            sym(g_api.DeviceGetCount, "nvmlDeviceGetCount_v2") &&
            sym(g_api.DeviceGetHandleByIndex, "nvmlDeviceGetHandleByIndex_v2") &&
            sym(g_api.DeviceGetUUID, "nvmlDeviceGetUUID");
  // memory: v2 preferred (the reference's call), v1 accepted for old drivers
  sym(g_api.DeviceGetMemoryInfo_v2, "nvmlDeviceGetMemoryInfo_v2");
  sym(g_api.DeviceGetMemoryInfo, "nvmlDeviceGetMemoryInfo");
  ok = ok && (g_api.DeviceGetMemoryInfo_v2 || g_api.DeviceGetMemoryInfo);
  // optional: PCI id (ordinal mapping) and events (health watch)
  sym(g_api.DeviceGetPciInfo, "nvmlDeviceGetPciInfo_v3");
  sym(g_api.EventSetCreate, "nvmlEventSetCreate");
  sym(g_api.DeviceRegisterEvents, "nvmlDeviceRegisterEvents");
  if (!sym(g_api.EventSetWait, "nvmlEventSetWait_v2")) sym(g_api.EventSetWait, "nvmlEventSetWait");
  sym(g_api.EventSetFree, "nvmlEventSetFree");
  if (!ok) {
    gs_internal::set_nvml_error("dlsym", 0, "library lacks a required NVML symbol");
    dlclose(g_api.handle);
    g_api = Api();
    return GS_E_NOLIB;
  }
  const nvmlReturn_t r = g_api.Init();
  if (r != NVML_SUCCESS) {
    const int rc = fail("nvmlInit_v2", r);
    dlclose(g_api.handle);
    g_api = Api();
    return rc;
  }
  g_open = true;
  return GS_OK;
}

int gs_nvml_close(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_open) return GS_E_STATE;
  if (g_events_open && g_api.EventSetFree) g_api.EventSetFree(g_events);
  g_events_open = false;
  g_handles.clear();
  const nvmlReturn_t r = g_api.Shutdown();
  dlclose(g_api.handle);
  g_api = Api();
  g_open = false;
  if (r != NVML_SUCCESS) {
    gs_internal::set_nvml_error("nvmlShutdown", (int)r, "shutdown failed");
    return GS_E_NVML;
  }
  return GS_OK;
}

int gs_nvml_device_count(uint32_t *count) {
  if (!count) return GS_E_INVAL;
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_open) return GS_E_STATE;
  unsigned int n = 0;
  const nvmlReturn_t r = g_api.DeviceGetCount(&n);
  if (r != NVML_SUCCESS) return fail("nvmlDeviceGetCount_v2", r);
  *count = n;
  return GS_OK;
}

int gs_nvml_device_info(uint32_t index, gs_gpu_info *out) {
  if (!out) return GS_E_INVAL;
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_open) return GS_E_STATE;
  memset(out, 0, sizeof(*out));
  out->nvml_index = (int32_t)index;
  out->cuda_ordinal = -1;
  nvmlDevice_t d;
  nvmlReturn_t r = g_api.DeviceGetHandleByIndex(index, &d);
  if (r != NVML_SUCCESS) return fail("nvmlDeviceGetHandleByIndex_v2", r);
  r = g_api.DeviceGetUUID(d, out->uuid, GS_UUID_MAX);
  if (r != NVML_SUCCESS) return fail("nvmlDeviceGetUUID", r);
  out->uuid[GS_UUID_MAX - 1] = '\0';
  if (g_api.DeviceGetMemoryInfo_v2) {
    nvmlMemory_v2_t m;
    memset(&m, 0, sizeof(m));
    m.version = nvmlMemory_v2;
    r = g_api.DeviceGetMemoryInfo_v2(d, &m);
    if (r != NVML_SUCCESS) return fail("nvmlDeviceGetMemoryInfo_v2", r);
    out->total_bytes = m.total;
  } else {
    nvmlMemory_t m;
    r = g_api.DeviceGetMemoryInfo(d, &m);
    if (r != NVML_SUCCESS) return fail("nvmlDeviceGetMemoryInfo", r);
    out->total_bytes = m.total;
  }
  if (g_api.DeviceGetPciInfo) {
    nvmlPciInfo_t pci;
    memset(&pci, 0, sizeof(pci));
    if (g_api.DeviceGetPciInfo(d, &pci) == NVML_SUCCESS) {
      snprintf(out->pci_bus_id, sizeof(out->pci_bus_id), "%s", pci.busId);
    }
  }
  return GS_OK;
}

int gs_health_open(uint64_t *unsupported_mask) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_open || g_events_open) return GS_E_STATE;
  if (!g_api.EventSetCreate || !g_api.DeviceRegisterEvents || !g_api.EventSetWait ||
      !g_api.EventSetFree) {
    gs_internal::set_nvml_error("gs_health_open", 0, "library lacks the NVML event API");
    return GS_E_NOLIB;
  }
  unsigned int n = 0;
  nvmlReturn_t r = g_api.DeviceGetCount(&n);
  if (r != NVML_SUCCESS) return fail("nvmlDeviceGetCount_v2", r);
  r = g_api.EventSetCreate(&g_events);
  if (r != NVML_SUCCESS) return fail("nvmlEventSetCreate", r);
  uint64_t mask = 0;
  g_handles.assign(n, nullptr);
  for (unsigned int i = 0; i < n; ++i) {
    r = g_api.DeviceGetHandleByIndex(i, &g_handles[i]);
    if (r != NVML_SUCCESS) {
      g_api.EventSetFree(g_events);
      return fail("nvmlDeviceGetHandleByIndex_v2", r);
    }
    r = g_api.DeviceRegisterEvents(g_handles[i], nvmlEventTypeXidCriticalError, g_events);
    if (r == NVML_ERROR_NOT_SUPPORTED) {  // nvidia.go:107-111: too old => unhealthy
      if (i < 64) mask |= 1ull << i;
      continue;
    }
    if (r != NVML_SUCCESS) {              // nvidia.go:114-116: fatal
      g_api.EventSetFree(g_events);
      return fail("nvmlDeviceRegisterEvents", r);
    }
  }
  if (unsupported_mask) *unsupported_mask = mask;
  g_events_open = true;
  return GS_OK;
}

int gs_health_wait(uint32_t timeout_ms, gs_health_event *out) {
  if (!out) return GS_E_INVAL;
  nvmlEventSet_t set;
  decltype(g_api.EventSetWait) wait_fn;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_open || !g_events_open) return GS_E_STATE;
    set = g_events;
    wait_fn = g_api.EventSetWait;
  }
  nvmlEventData_t data;
  memset(&data, 0, sizeof(data));
  const nvmlReturn_t r = wait_fn(set, &data, timeout_ms);  // blocks outside the lock
  if (r == NVML_ERROR_TIMEOUT) return GS_E_TIMEOUT;
  if (r != NVML_SUCCESS) return fail("nvmlEventSetWait_v2", r);
  if (data.eventType != nvmlEventTypeXidCriticalError) return GS_E_TIMEOUT;  // nvidia.go:126-128
  out->xid = data.eventData;
  // nvidia.go:134-136: application errors leave the GPU healthy
  out->ignored = (data.eventData == 31 || data.eventData == 43 || data.eventData == 45) ? 1 : 0;
  out->gpu_index = GS_ALL_GPUS;  // nvidia.go:138-143: no device identity => all GPUs
  std::lock_guard<std::mutex> lk(g_mu);
  for (size_t i = 0; i < g_handles.size(); ++i)
    if (data.device && g_handles[i] == data.device) out->gpu_index = (int32_t)i;
  return GS_OK;
}

int gs_health_close(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_open || !g_events_open) return GS_E_STATE;
  g_api.EventSetFree(g_events);
  g_events_open = false;
  g_handles.clear();
  return GS_OK;
}

}  // extern "C"

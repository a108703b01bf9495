// errors.cpp -- status strings and per-thread last-error text
#include <stdio.h>
#include <string.h>

#include "gpushare.h"
#include "internal.h"

namespace {
thread_local char g_cuda_err[256] = "";
thread_local char g_nvml_err[256] = "";
}  // namespace

namespace gs_internal {
void set_cuda_error(const char *what, int code, const char *msg) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s (%d)", what, msg ? msg : "?", code);
}
void set_nvml_error(const char *what, int code, const char *msg) {
  snprintf(g_nvml_err, sizeof(g_nvml_err), "%s: %s (%d)", what, msg ? msg : "?", code);
}
}  // namespace gs_internal

extern "C" {

const char *gs_strerror(int status) {
  switch (status) {
    case GS_OK: return "ok";
    case GS_E_INVAL: return "invalid argument";
    case GS_E_NOSPACE: return "output buffer too small";
    case GS_E_NOTFOUND: return "not found";
    case GS_E_WIRE: return "malformed protobuf wire bytes";
    case GS_E_NVML: return "NVML call failed";
    case GS_E_CUDA: return "CUDA call failed";
    case GS_E_NOLIB: return "shared library could not be loaded";
    case GS_E_STATE: return "call made in the wrong state";
    case GS_E_TIMEOUT: return "timed out";
    case GS_E_LIMIT: return "size beyond the documented maximum";
    default: return "unknown status";
  }
}

const char *gs_version(void) { return "gpushare-b200 0.1.0 (sm_100a)"; }
const char *gs_cuda_last_error(void) { return g_cuda_err; }
const char *gs_nvml_last_error(void) { return g_nvml_err; }

}  // extern "C"

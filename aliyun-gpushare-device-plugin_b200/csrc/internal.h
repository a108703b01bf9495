// internal.h -- helpers shared by the translation units of libgpushare_core.so
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <new>

namespace gs_internal {
// last-error slots are thread-local so concurrent callers do not trample them
void set_cuda_error(const char *what, int code, const char *msg);
void set_nvml_error(const char *what, int code, const char *msg);
}  // namespace gs_internal

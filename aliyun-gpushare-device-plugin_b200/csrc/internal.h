// internal.h -- helpers shared by the translation units of libgpushare_core.so
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <new>

#include "gpushare.h"

namespace gs_internal {
// last-error slots are thread-local so concurrent callers do not trample them
void set_cuda_error(const char *what, int code, const char *msg);
void set_nvml_error(const char *what, int code, const char *msg);
// the transport wakes its serving thread through eventfds the plugin signals on every
// health notification and on stop (plugin.cpp)
void plugin_add_wake_fd(struct gs_plugin *p, int fd);
void plugin_remove_wake_fd(struct gs_plugin *p, int fd);
struct gs_table *plugin_table(struct gs_plugin *p);
}  // namespace gs_internal

// internal.h -- helpers shared by the translation units of libgpushare_core.so
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <new>
#include <vector>

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

// One candidate pod of getCandidatePods (podmanager.go:300-323): its parsed assume time and
// its index in the lister's list. The list a decision sorts is these, in lister order.
struct Cand {
  uint64_t time;
  uint32_t pod;
};
// podmanager.go:273-293,306-311 over a lister's list: this node's pods, first occurrence per
// UID, assumed but not assigned -- in list order. gs_allocate does this per call; gs_plugin does
// it once per installed list and drops a pod when it claims it.
void build_candidates(gs_str node_name, const gs_pod_view *pods, size_t n_pods, std::vector<Cand> &out);
// gs_allocate with the candidate list already known (same result as gs_allocate on `pods`).
int allocate_with_candidates(const gs_alloc_ctx *ctx, const uint8_t *req, size_t req_len,
                             const gs_pod_view *pods, size_t n_pods, const Cand *cands, size_t n_cands,
                             uint8_t *resp, size_t resp_cap, char *patch, size_t patch_cap,
                             gs_alloc_result *out);
}  // namespace gs_internal

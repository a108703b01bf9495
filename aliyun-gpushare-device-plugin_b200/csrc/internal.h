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
// The id scan of an AllocateRequest, run while the request is still arriving (the transport feeds
// it what has come in so far; the scan hides behind the socket copy instead of following it). It
// handles exactly the shape kubelet sends -- container_requests of devices_ids with one-byte
// lengths -- and gives up (ok = false) on anything else, in which case the ordinary scan of the
// complete request decides, errors included: the counts it delivers are the ordinary scan's or none.
constexpr uint32_t kScanMaxContainers = 1024;
struct RequestScanCursor {
  size_t pos = 0;    // next unparsed byte of the message
  size_t cend = 0;   // end of the container being walked; 0 = between containers
  uint32_t cur = 0;  // ids counted in it so far
  uint32_t n_containers = 0;
  uint64_t total = 0;
  bool ok = true;
  uint32_t counts[kScanMaxContainers];
};
void request_scan_feed(RequestScanCursor &c, const uint8_t *msg, size_t avail);
// true: the cursor walked exactly `len` bytes; its counts stand for the request
bool request_scan_finish(RequestScanCursor &c, const uint8_t *msg, size_t len);
// gs_plugin_allocate with the id scan already done (same answer, same claim)
int plugin_allocate_prescanned(struct gs_plugin *p, const uint8_t *req, size_t req_len,
                               const RequestScanCursor &scan, uint8_t *resp, size_t resp_cap,
                               gs_alloc_result *out);
// gs_allocate with the candidate list already known (same result as gs_allocate on `pods`);
// `scan` non-null: the id counts are already known too.
int allocate_with_candidates(const gs_alloc_ctx *ctx, const uint8_t *req, size_t req_len,
                             const gs_pod_view *pods, size_t n_pods, const Cand *cands, size_t n_cands,
                             uint8_t *resp, size_t resp_cap, char *patch, size_t patch_cap,
                             gs_alloc_result *out, const RequestScanCursor *scan = nullptr);
}  // namespace gs_internal

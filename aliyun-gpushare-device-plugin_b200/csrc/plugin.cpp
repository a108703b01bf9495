// plugin.cpp -- host-side mirror of the reference's NvidiaDevicePlugin for the hot path.
//
// Replaces (reference paths relative to cjg/aliyun-gpushare-device-plugin):
//   type NvidiaDevicePlugin + sync.RWMutex        pkg/gpu/nvidia/server.go:19-35
//   Allocate (whole RPC under m.Lock())           allocate.go:43-201
//   GetDevicePluginOptions / PreStartContainer    server.go:93-95,199-201
//   ListAndWatch's wait on m.stop / m.health      server.go:183-192
//   Stop() closing m.stop                          server.go:145-155
//
// The pod lister (kubelet /pods or apiserver LIST, podmanager.go:187-245) and the
// patcher (allocate.go:136) are network I/O and stay with the caller; here they are
// an in-memory pod store, so the same store can sit under this plugin and under the
// oracle in every comparison. Pods live in fixed slots whose annotation strings have
// reserved capacity: applying the "assigned" patch never allocates.
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include "gpushare.h"
#include "internal.h"

namespace {

struct StoredPod {
  std::string uid, ns, name, node;
  uint64_t limit = 0;
  bool has_idx = false, has_time = false, has_assigned = false;
  std::string idx, time, assigned;
};

int64_t real_now_ns() {
  struct timespec ts;
  clock_gettime(CLOCK_REALTIME, &ts);
  return (int64_t)ts.tv_sec * 1000000000ll + ts.tv_nsec;
}

void assign(std::string &dst, const gs_str &s) {
  if (s.p) dst.assign(s.p, s.n);
  else dst.clear();
}
gs_str view(const std::string &s) { return gs_str{s.data(), s.size()}; }

}  // namespace

// NvidiaDevicePlugin (server.go:19-35), minus the gRPC transport.
struct gs_plugin {
  gs_table *table = nullptr;                 // devs / devNameMap / devIndxMap
  std::string node_name;                     // nodeName (podmanager.go:26,55)
  bool disable_cgpu_isolation = false;       // server.go:27
  std::mutex mu;                             // sync.RWMutex, taken as in allocate.go:60
  std::vector<StoredPod> pods;               // stand-in for the lister
  std::vector<gs_pod_view> views;            // rebuilt only when pods change shape
  // getCandidatePods over the installed list, in list order: built when the list is installed,
  // one entry dropped each time a pod is claimed -- a decision then costs the sort of the pending
  // pods and the match, not a pass over every pod of the node
  std::vector<gs_internal::Cand> cands;
  int64_t fixed_clock = 0;                   // 0 = time.Now()
  std::mutex health_mu;
  std::condition_variable health_cv;         // m.health channel
  std::atomic<bool> stopped{false};          // m.stop channel
  std::vector<int> wake_fds;                 // eventfds of transports serving this plugin

  void wake() {  // health_mu held
    const uint64_t one = 1;
    for (int fd : wake_fds)
      if (write(fd, &one, sizeof(one)) < 0) {
      }
  }

  void refresh_view(size_t i) {
    const StoredPod &p = pods[i];
    gs_pod_view &v = views[i];
    v.uid = view(p.uid);
    v.ns = view(p.ns);
    v.name = view(p.name);
    v.node_name = view(p.node);
    v.gpu_mem_limit = p.limit;
    v.ann_idx = p.has_idx ? view(p.idx) : gs_str{nullptr, 0};
    v.ann_assume_time = p.has_time ? view(p.time) : gs_str{nullptr, 0};
    v.ann_assigned = p.has_assigned ? view(p.assigned) : gs_str{nullptr, 0};
  }
};

namespace gs_internal {
void plugin_add_wake_fd(gs_plugin *p, int fd) {
  std::lock_guard<std::mutex> lk(p->health_mu);
  p->wake_fds.push_back(fd);
}
void plugin_remove_wake_fd(gs_plugin *p, int fd) {
  std::lock_guard<std::mutex> lk(p->health_mu);
  for (size_t i = 0; i < p->wake_fds.size(); ++i)
    if (p->wake_fds[i] == fd) {
      p->wake_fds.erase(p->wake_fds.begin() + (long)i);
      break;
    }
}
gs_table *plugin_table(gs_plugin *p) { return p->table; }
}  // namespace gs_internal

extern "C" {

int gs_plugin_new(gs_table *table, const char *node_name, size_t node_len,
                  int32_t disable_cgpu_isolation, gs_plugin **out) {
  if (!out) return GS_E_INVAL;
  *out = nullptr;
  if (!table || (!node_name && node_len)) return GS_E_INVAL;
  gs_plugin *p = new (std::nothrow) gs_plugin();
  if (!p) return GS_E_INVAL;
  p->table = table;
  p->node_name.assign(node_name ? node_name : "", node_len);
  p->disable_cgpu_isolation = disable_cgpu_isolation != 0;
  *out = p;
  return GS_OK;
}

void gs_plugin_free(gs_plugin *p) {
  if (!p) return;
  gs_plugin_stop(p);
  delete p;
}

int gs_plugin_set_pods(gs_plugin *p, const gs_pod_view *pods, size_t n) {
  if (!p || (!pods && n)) return GS_E_INVAL;
  std::lock_guard<std::mutex> lk(p->mu);
  // slots and their strings are reused: re-installing a list of the same shape (every
  // lister refresh) allocates nothing
  p->pods.resize(n);
  p->views.resize(n);
  for (size_t i = 0; i < n; ++i) {
    StoredPod &s = p->pods[i];
    assign(s.uid, pods[i].uid);
    assign(s.ns, pods[i].ns);
    assign(s.name, pods[i].name);
    assign(s.node, pods[i].node_name);
    s.limit = pods[i].gpu_mem_limit;
    s.has_idx = pods[i].ann_idx.p != nullptr;
    s.has_time = pods[i].ann_assume_time.p != nullptr;
    s.has_assigned = pods[i].ann_assigned.p != nullptr;
    assign(s.idx, pods[i].ann_idx);
    // room for "true" and a 20-digit timestamp so the patch is applied in place
    if (s.time.capacity() < 32) s.time.reserve(32);
    if (s.assigned.capacity() < 16) s.assigned.reserve(16);
    assign(s.time, pods[i].ann_assume_time);
    assign(s.assigned, pods[i].ann_assigned);
  }
  for (size_t i = 0; i < n; ++i) p->refresh_view(i);
  gs_internal::build_candidates(view(p->node_name), p->views.data(), n, p->cands);
  return GS_OK;
}

int gs_plugin_pod_count(const gs_plugin *p, size_t *n) {
  if (!p || !n) return GS_E_INVAL;
  *n = p->pods.size();
  return GS_OK;
}

int gs_plugin_get_pod(const gs_plugin *p, size_t index, gs_pod_view *out) {
  if (!p || !out) return GS_E_INVAL;
  if (index >= p->views.size()) return GS_E_NOTFOUND;
  *out = p->views[index];
  return GS_OK;
}

int gs_plugin_set_clock(gs_plugin *p, int64_t fixed_unix_nano) {
  if (!p) return GS_E_INVAL;
  p->fixed_clock = fixed_unix_nano;
  return GS_OK;
}

int gs_plugin_allocate(gs_plugin *p, const uint8_t *req, size_t req_len, uint8_t *resp,
                       size_t resp_cap, gs_alloc_result *out) {
  if (!p || !out) return GS_E_INVAL;
  std::lock_guard<std::mutex> lk(p->mu);  // allocate.go:60-61
  gs_alloc_ctx ctx;
  ctx.table = p->table;
  ctx.node_name = view(p->node_name);
  ctx.disable_cgpu_isolation = p->disable_cgpu_isolation ? 1 : 0;
  ctx.now_unix_nano = p->fixed_clock ? p->fixed_clock : real_now_ns();
  char patch[192];
  if ((!req && req_len) || !p->table) return GS_E_INVAL;
  const int rc = gs_internal::allocate_with_candidates(&ctx, req, req_len, p->views.data(), p->views.size(),
                                                       p->cands.data(), p->cands.size(), resp, resp_cap, patch,
                                                       sizeof(patch), out);
  if (rc != GS_OK) return rc;
  if (out->kind == GS_ALLOC_MATCHED) {
    // claimed: ASSIGNED becomes "true", so the pod leaves the candidates (podutils.go:78-119)
    for (size_t i = 0; i < p->cands.size(); ++i)
      if (p->cands[i].pod == (uint32_t)out->pod_index) {
        p->cands.erase(p->cands.begin() + (long)i);
        break;
      }
    // the PATCH of allocate.go:136 applied to the in-memory store:
    // {"metadata":{"annotations":{ASSIGNED:"true", ASSUME_TIME:"<now>"}}}
    StoredPod &s = p->pods[(size_t)out->pod_index];
    s.assigned.assign("true", 4);
    s.has_assigned = true;
    char num[24];
    int n = 0;
    {
      int64_t v = ctx.now_unix_nano;
      uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
      char tmp[24];
      int k = 0;
      do {
        tmp[k++] = (char)('0' + u % 10);
        u /= 10;
      } while (u);
      if (v < 0) num[n++] = '-';
      while (k) num[n++] = tmp[--k];
    }
    s.time.assign(num, (size_t)n);
    s.has_time = true;
    p->refresh_view((size_t)out->pod_index);
  }
  return GS_OK;
}

int gs_plugin_get_device_plugin_options(gs_plugin *p, uint8_t *resp, size_t cap, size_t *len) {
  (void)resp;
  (void)cap;
  if (!p || !len) return GS_E_INVAL;
  *len = 0;  // &pluginapi.DevicePluginOptions{}: both bools false => empty message
  return GS_OK;
}

int gs_plugin_pre_start_container(gs_plugin *p, const uint8_t *req, size_t req_len,
                                  uint8_t *resp, size_t cap, size_t *len) {
  (void)req;
  (void)req_len;
  (void)resp;
  (void)cap;
  if (!p || !len) return GS_E_INVAL;
  *len = 0;  // &pluginapi.PreStartContainerResponse{}
  return GS_OK;
}

int gs_plugin_wait_health_change(gs_plugin *p, uint64_t *seen_generation, uint32_t timeout_ms) {
  if (!p || !seen_generation) return GS_E_INVAL;
  std::unique_lock<std::mutex> lk(p->health_mu);
  const bool woke = p->health_cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), [&] {
    return p->stopped.load() || gs_law_generation(p->table) != *seen_generation;
  });
  if (p->stopped.load()) return GS_E_STATE;  // <-m.stop: ListAndWatch returns
  if (!woke) return GS_E_TIMEOUT;
  *seen_generation = gs_law_generation(p->table);
  return GS_OK;
}

int gs_plugin_notify_health(gs_plugin *p) {
  if (!p) return GS_E_INVAL;
  {
    std::lock_guard<std::mutex> lk(p->health_mu);
    p->wake();
  }
  p->health_cv.notify_all();
  return GS_OK;
}

int gs_plugin_stop(gs_plugin *p) {
  if (!p) return GS_E_INVAL;
  {
    std::lock_guard<std::mutex> lk(p->health_mu);
    p->stopped.store(true);
    p->wake();
  }
  p->health_cv.notify_all();
  return GS_OK;
}

int gs_plugin_stopped(const gs_plugin *p) { return p && p->stopped.load() ? 1 : 0; }

}  // extern "C"

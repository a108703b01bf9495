// gpushare.hpp -- header-only C++ face of the C ABI, in the reference's own vocabulary.
//
// Mirrors pkg/gpu/nvidia of cjg/aliyun-gpushare-device-plugin for the hot path:
//   NvidiaDevicePlugin            server.go:19-35 (struct), 43-78 (constructor)
//     GetDeviceNameByIndex        server.go:80-91
//     GetDevicePluginOptions      server.go:93-95
//     ListAndWatch (one message)  server.go:180-193
//     PreStartContainer           server.go:199-201
//     Allocate                    allocate.go:43-201
//     Stop                        server.go:145-155
// Same names, argument meaning and error behaviour: Allocate never fails for a "bad" pod
// (the answer carries the poison envs of buildErrResponse, allocate.go:25-40); only
// malformed wire bytes or exhausted buffers raise.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "gpushare.h"

namespace gpushare {

struct Error : std::runtime_error {
  int status;
  Error(const char *where, int s)
      : std::runtime_error(std::string(where) + ": " + gs_strerror(s)), status(s) {}
};
inline void check(int rc, const char *where) {
  if (rc != GS_OK) throw Error(where, rc);
}

enum class MemoryUnit { GiB = GS_UNIT_GIB, MiB = GS_UNIT_MIB };  // const.go:8,34-35

struct Pod {  // what the lister hands over for one v1.Pod
  std::string uid, ns, name, node_name;
  uint64_t gpu_mem_limit = 0;
  bool has_idx = false, has_assume_time = false, has_assigned = false;
  std::string idx, assume_time, assigned;
};

class NvidiaDevicePlugin {
 public:
  // getDevices (nvidia.go:50-86) + NewNvidiaDevicePlugin (server.go:43-78)
  NvidiaDevicePlugin(const std::vector<std::string> &uuids, uint64_t raw_mib_gpu0, MemoryUnit unit,
                     const std::string &node_name, bool disable_cgpu_isolation = false) {
    std::vector<const char *> p;
    std::vector<uint32_t> l;
    for (const auto &u : uuids) {
      p.push_back(u.data());
      l.push_back((uint32_t)u.size());
    }
    check(gs_table_build(p.data(), l.data(), (uint32_t)uuids.size(),
                         gs_units_from_raw(raw_mib_gpu0, (int)unit), (int)unit, &table_),
          "gs_table_build");
    const int rc = gs_plugin_new(table_, node_name.data(), node_name.size(),
                                 disable_cgpu_isolation ? 1 : 0, &plugin_);
    if (rc != GS_OK) {
      gs_table_free(table_);
      throw Error("gs_plugin_new", rc);
    }
  }
  ~NvidiaDevicePlugin() {
    if (server_) gs_server_stop(server_);
    gs_plugin_free(plugin_);
    gs_table_free(table_);
  }
  NvidiaDevicePlugin(const NvidiaDevicePlugin &) = delete;
  NvidiaDevicePlugin &operator=(const NvidiaDevicePlugin &) = delete;

  // the in-memory pod source standing in for the lister / patcher
  void SetPods(const std::vector<Pod> &pods) {
    std::vector<gs_pod_view> v(pods.size());
    for (size_t i = 0; i < pods.size(); ++i) {
      const Pod &p = pods[i];
      v[i].uid = {p.uid.data(), p.uid.size()};
      v[i].ns = {p.ns.data(), p.ns.size()};
      v[i].name = {p.name.data(), p.name.size()};
      v[i].node_name = {p.node_name.data(), p.node_name.size()};
      v[i].gpu_mem_limit = p.gpu_mem_limit;
      v[i].ann_idx = p.has_idx ? gs_str{p.idx.data(), p.idx.size()} : gs_str{nullptr, 0};
      v[i].ann_assume_time =
          p.has_assume_time ? gs_str{p.assume_time.data(), p.assume_time.size()} : gs_str{nullptr, 0};
      v[i].ann_assigned =
          p.has_assigned ? gs_str{p.assigned.data(), p.assigned.size()} : gs_str{nullptr, 0};
    }
    check(gs_plugin_set_pods(plugin_, v.data(), v.size()), "gs_plugin_set_pods");
  }
  void SetClock(int64_t unix_nano) { check(gs_plugin_set_clock(plugin_, unix_nano), "gs_plugin_set_clock"); }

  std::pair<std::string, bool> GetDeviceNameByIndex(uint64_t index) const {
    const char *p = nullptr;
    size_t n = 0;
    if (gs_device_name_by_index(table_, index, &p, &n) != GS_OK) return {"", false};
    return {std::string(p, n), true};
  }

  std::vector<uint8_t> GetDevicePluginOptions() const { return {}; }
  std::vector<uint8_t> PreStartContainer(const std::vector<uint8_t> &) const { return {}; }

  // One serialized ListAndWatchResponse for the current health state.
  std::vector<uint8_t> ListAndWatch() const {
    std::vector<uint8_t> out(gs_law_payload_size(table_));
    size_t n = 0;
    check(gs_law_encode(table_, out.data(), out.size(), &n), "gs_law_encode");
    out.resize(n);
    return out;
  }

  // AllocateRequest wire bytes in, AllocateResponse wire bytes out.
  std::vector<uint8_t> Allocate(const std::vector<uint8_t> &request, gs_alloc_result *result = nullptr) {
    std::vector<uint8_t> out(4096);
    gs_alloc_result res;
    int rc = gs_plugin_allocate(plugin_, request.data(), request.size(), out.data(), out.size(), &res);
    if (rc == GS_E_NOSPACE) {
      out.resize(res.resp_len);
      rc = gs_plugin_allocate(plugin_, request.data(), request.size(), out.data(), out.size(), &res);
    }
    check(rc, "gs_plugin_allocate");
    out.resize(res.resp_len);
    if (result) *result = res;
    return out;
  }

  void unhealthy(int32_t gpu_or_all) {  // healthcheck pump, server.go:211-229
    check(gs_law_mark_unhealthy_gpu(table_, gpu_or_all), "gs_law_mark_unhealthy_gpu");
    gs_plugin_notify_health(plugin_);
  }
  // Start (server.go:114-142): serve the five RPCs on `socket_path` with the library's own
  // gRPC transport. Serve (server.go:232-249) = Start + Register, undone when Register fails.
  void Start(const std::string &socket_path = GS_DEVICE_PLUGIN_PATH GS_SERVER_SOCK_NAME,
             gs_server_allocate_fn allocate = nullptr, void *user = nullptr) {
    gs_server_config cfg{};
    cfg.socket_path = socket_path.c_str();
    cfg.allocate = allocate;
    cfg.user = user;
    check(gs_server_start(plugin_, &cfg, &server_), "gs_server_start");
    socket_ = socket_path;
  }
  void Register(const std::string &kubelet_socket = GS_KUBELET_SOCKET,
                const std::string &resource_name = GS_RESOURCE_NAME) {  // server.go:158-177
    const size_t slash = socket_.rfind('/');
    const std::string endpoint = slash == std::string::npos ? socket_ : socket_.substr(slash + 1);
    check(gs_register(kubelet_socket.c_str(), endpoint.c_str(), resource_name.c_str(), 5000), "Register");
  }
  void Serve(const std::string &socket_path = GS_DEVICE_PLUGIN_PATH GS_SERVER_SOCK_NAME,
             const std::string &kubelet_socket = GS_KUBELET_SOCKET) {
    Start(socket_path);
    try {
      Register(kubelet_socket);
    } catch (...) {
      Stop();
      throw;
    }
  }
  // Stop (server.go:145-155): close(m.stop); a running server ends its streams and goes away
  void Stop() {
    if (server_) {
      gs_server_stop(server_);
      server_ = nullptr;
    } else {
      gs_plugin_stop(plugin_);
    }
  }

  gs_table *table() const { return table_; }
  gs_plugin *handle() const { return plugin_; }

 private:
  gs_table *table_ = nullptr;
  gs_plugin *plugin_ = nullptr;
  gs_server *server_ = nullptr;
  std::string socket_;
};

}  // namespace gpushare

// C++ consumer of include/gpushare.hpp: golden case G1 (binpack-1, BASELINE config 1) and the
// error response, decoded by hand. Built and run by tests/test_cpp_face.py.
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "gpushare.hpp"

using Bytes = std::vector<uint8_t>;

static void ld(Bytes &o, int field, const Bytes &payload) {
  o.push_back((uint8_t)((field << 3) | 2));
  size_t n = payload.size();
  while (n >= 0x80) { o.push_back((uint8_t)(n | 0x80)); n >>= 7; }
  o.push_back((uint8_t)n);
  o.insert(o.end(), payload.begin(), payload.end());
}
static Bytes request(const std::vector<std::string> &ids) {
  Bytes inner;
  for (const auto &i : ids) ld(inner, 1, Bytes(i.begin(), i.end()));
  Bytes out;
  ld(out, 1, inner);
  return out;
}
static size_t varint(const Bytes &b, size_t &p) {
  size_t v = 0; int s = 0;
  for (;;) { uint8_t c = b[p++]; v |= (size_t)(c & 0x7F) << s; if (!(c & 0x80)) return v; s += 7; }
}
// first container's env map
static std::map<std::string, std::string> envs(const Bytes &resp) {
  std::map<std::string, std::string> m;
  size_t p = 0;
  if (resp.empty() || resp[p++] != 0x0A) return m;
  size_t end = varint(resp, p); end += p;
  while (p < end) {
    ++p;  // 0x0A: envs entry
    size_t e = varint(resp, p); e += p;
    std::string k, v;
    while (p < e) {
      uint8_t tag = resp[p++];
      size_t n = varint(resp, p);
      (tag == 0x0A ? k : v).assign((const char *)&resp[p], n);
      p += n;
    }
    m[k] = v;
  }
  return m;
}
#define EXPECT(c) do { if (!(c)) { printf("FAIL line %d: %s\n", __LINE__, #c); return 1; } } while (0)

int main() {
  std::vector<std::string> uuids;
  for (int i = 0; i < 8; ++i) {
    char b[64];
    snprintf(b, sizeof b, "GPU-%08x-0000-0000-0000-%012x", i, i);
    uuids.push_back(b);
  }
  gpushare::NvidiaDevicePlugin plugin(uuids, 16384, gpushare::MemoryUnit::GiB, "node-a");
  plugin.SetClock(1700000000123456789ll);
  std::vector<gpushare::Pod> pods(3);
  for (int k = 0; k < 3; ++k) {
    pods[k].uid = "uid-" + std::to_string(k);
    pods[k].ns = "default";
    pods[k].name = "binpack-1-" + std::to_string(k);
    pods[k].node_name = "node-a";
    pods[k].gpu_mem_limit = 2;
    pods[k].has_idx = pods[k].has_assume_time = pods[k].has_assigned = true;
    pods[k].idx = "0";
    pods[k].assume_time = std::to_string(1000 * (k + 1));
    pods[k].assigned = "false";
  }
  plugin.SetPods(pods);
  const Bytes req = request({uuids[5] + "-_-3", uuids[2] + "-_-7"});
  for (int k = 0; k < 3; ++k) {
    gs_alloc_result r;
    auto e = envs(plugin.Allocate(req, &r));
    EXPECT(r.kind == GS_ALLOC_MATCHED && r.pod_index == k && r.gpu_index == 0);
    EXPECT(e.size() == 5 && e["NVIDIA_VISIBLE_DEVICES"] == "0" && e["ALIYUN_COM_GPU_MEM_IDX"] == "0");
    EXPECT(e["ALIYUN_COM_GPU_MEM_POD"] == "2" && e["ALIYUN_COM_GPU_MEM_CONTAINER"] == "2");
    EXPECT(e["ALIYUN_COM_GPU_MEM_DEV"] == "16");
  }
  gs_alloc_result r;
  auto e = envs(plugin.Allocate(req, &r));  // nobody left: poison envs, still a normal answer
  EXPECT(r.kind == GS_ALLOC_ERR_NO_MATCH);
  EXPECT(e["NVIDIA_VISIBLE_DEVICES"] == "no-gpu-has-2GiB-to-run" && e["ALIYUN_COM_GPU_MEM_IDX"] == "-1");
  EXPECT(plugin.GetDeviceNameByIndex(7).first == uuids[7] && !plugin.GetDeviceNameByIndex(8).second);
  EXPECT(plugin.ListAndWatch().size() == 128 * (13 + 44) + 8 * 6);  // ids -_-10..15 are one longer
  plugin.unhealthy(3);
  EXPECT(plugin.ListAndWatch().size() == 128 * 57 + 48 + 2 * 16);
  EXPECT(plugin.GetDevicePluginOptions().empty() && plugin.PreStartContainer({}).empty());
  bool threw = false;
  try { plugin.Allocate(Bytes{0x0A, 0x05, 0x01}); } catch (const gpushare::Error &err) { threw = err.status == GS_E_WIRE; }
  EXPECT(threw);
  {  // the transport from C++: Start on a socket, call it with the library's client, Stop
    std::vector<std::string> one = {uuids[0]};
    gpushare::NvidiaDevicePlugin served(one, 16384, gpushare::MemoryUnit::GiB, "node-a");
    char dir[] = "/tmp/gscppXXXXXX";
    EXPECT(mkdtemp(dir) != nullptr);
    const std::string sock = std::string(dir) + "/aliyungpushare.sock";
    served.Start(sock);
    gs_grpc_client *c = nullptr;
    EXPECT(gs_grpc_client_connect(sock.c_str(), 5000, &c) == GS_OK);
    uint8_t resp[1024];
    size_t n = 0;
    int32_t status = -1;
    char msg[64];
    EXPECT(gs_grpc_client_unary(c, "/v1beta1.DevicePlugin/Allocate", req.data(), req.size(), resp, sizeof(resp),
                                &n, &status, msg, sizeof(msg)) == GS_OK && status == 0);
    auto e2 = envs(Bytes(resp, resp + n));  // no pod, one GPU: the UUID fallback (allocate.go:154-180)
    EXPECT(e2["NVIDIA_VISIBLE_DEVICES"] == uuids[0] && e2["ALIYUN_COM_GPU_MEM_IDX"] == "0");
    bool refused = false;
    try { served.Register(std::string(dir) + "/kubelet.sock"); } catch (const gpushare::Error &err) { refused = err.status == GS_E_TIMEOUT; }
    EXPECT(refused);  // nobody listens there
    gs_grpc_client_close(c);
    served.Stop();
    EXPECT(access(sock.c_str(), F_OK) != 0);
    rmdir(dir);
  }
  printf("CPP FACE OK\n");
  return 0;
}

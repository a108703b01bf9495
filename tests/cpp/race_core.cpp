// ThreadSanitizer run: Allocate from several threads, health marks, ListAndWatch encodes and the
// health-change wait all at once on one table/plugin. Built with -fsanitize=thread by
// tests/test_cpp_face.py; any data race aborts.
#include <stdio.h>

#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "gpushare.h"

int main() {
  std::vector<std::string> uuids;
  for (int i = 0; i < 4; ++i) uuids.push_back("GPU-0000000" + std::to_string(i) + "-0000-0000-0000-00000000000" + std::to_string(i));
  std::vector<const char *> p;
  std::vector<uint32_t> l;
  for (auto &u : uuids) { p.push_back(u.data()); l.push_back((uint32_t)u.size()); }
  gs_table *tab = nullptr;
  if (gs_table_build(p.data(), l.data(), 4, 2000, GS_UNIT_MIB, &tab) != GS_OK) return 1;
  gs_plugin *plug = nullptr;
  gs_plugin_new(tab, "n", 1, 0, &plug);
  std::vector<gs_pod_view> pods(32);
  std::vector<std::string> keep;
  keep.reserve(32 * 4);
  for (int i = 0; i < 32; ++i) {
    keep.push_back("u" + std::to_string(i));
    pods[i].uid = {keep.back().data(), keep.back().size()};
    pods[i].ns = {"d", 1};
    keep.push_back("p" + std::to_string(i));
    pods[i].name = {keep.back().data(), keep.back().size()};
    pods[i].node_name = {"n", 1};
    pods[i].gpu_mem_limit = 2;
    pods[i].ann_idx = {"1", 1};
    keep.push_back(std::to_string(1000 + i));
    pods[i].ann_assume_time = {keep.back().data(), keep.back().size()};
    pods[i].ann_assigned = {"false", 5};
  }
  const uint8_t req[] = {0x0A, 0x06, 0x0A, 0x01, 'a', 0x0A, 0x01, 'b'};
  std::atomic<bool> stop{false};
  std::atomic<int> matched{0};
  std::vector<std::thread> th;
  for (int t = 0; t < 4; ++t)
    th.emplace_back([&] {
      uint8_t resp[512];
      gs_alloc_result r;
      for (int k = 0; k < 400; ++k) {
        if (k % 50 == 0) gs_plugin_set_pods(plug, pods.data(), pods.size());
        if (gs_plugin_allocate(plug, req, sizeof req, resp, sizeof resp, &r) == GS_OK && r.kind == GS_ALLOC_MATCHED) ++matched;
        gs_alloc_ctx ctx{tab, gs_str{"n", 1}, 0, 5};
        char patch[160];
        gs_allocate(&ctx, req, sizeof req, pods.data(), pods.size(), resp, sizeof resp, patch, sizeof patch, &r);
      }
    });
  th.emplace_back([&] {  // health events
    for (int k = 0; k < 300; ++k) {
      gs_law_mark_unhealthy_device(tab, (uint32_t)(k % 4), (uint64_t)(k * 7 % 2000));
      if (k % 60 == 0) gs_law_mark_unhealthy_gpu(tab, k % 4);
      if (k % 100 == 99) gs_law_reset_health(tab);
      gs_plugin_notify_health(plug);
    }
  });
  for (int t = 0; t < 2; ++t)
    th.emplace_back([&] {  // ListAndWatch senders
      std::vector<uint8_t> buf(4 * 2000 * 80);
      uint64_t seen = gs_law_generation(tab);
      while (!stop.load()) {
        size_t len = 0;
        gs_law_encode(tab, buf.data(), buf.size(), &len);
        gs_iovec iov[4];
        uint32_t n = 0;
        gs_law_payload(tab, iov, 4, &n);
        gs_law_payload_size(tab);
        gs_law_device_healthy(tab, 1, 5);
        gs_plugin_wait_health_change(plug, &seen, 1);
      }
    });
  for (size_t i = 0; i < 5; ++i) th[i].join();
  stop.store(true);
  gs_plugin_stop(plug);
  for (size_t i = 5; i < th.size(); ++i) th[i].join();
  gs_plugin_free(plug);
  gs_table_free(tab);
  printf("RACE CORE OK matched=%d\n", matched.load());
  return 0;
}

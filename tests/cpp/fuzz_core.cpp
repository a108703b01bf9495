// Sanitizer run of the host-side core (no CUDA): built with -fsanitize=address,undefined from
// allocate.cpp, table.cpp, gosem.cpp, plugin.cpp, errors.cpp and hammered with random inputs.
// Any out-of-bounds access, overflow or leak aborts the process. Run by tests/test_cpp_face.py.
#include <stdio.h>
#include <string.h>

#include <random>
#include <string>
#include <vector>

#include "gpushare.h"

static std::mt19937_64 rng(20261106);
static uint64_t R(uint64_t n) { return n ? rng() % n : 0; }

static std::string rand_str(const char *alphabet, size_t max) {
  std::string s;
  const size_t n = R(max + 1), m = strlen(alphabet);
  for (size_t i = 0; i < n; ++i) s.push_back(alphabet[R(m)]);
  return s;
}

static void put_varint(std::vector<uint8_t> &o, uint64_t v) {
  while (v >= 0x80) { o.push_back((uint8_t)(v | 0x80)); v >>= 7; }
  o.push_back((uint8_t)v);
}

static std::vector<uint8_t> rand_request() {
  std::vector<uint8_t> out;
  const int nc = (int)R(4);
  for (int c = 0; c < nc; ++c) {
    std::vector<uint8_t> inner;
    const int ni = (int)R(R(10) ? 40 : 3000);
    for (int i = 0; i < ni; ++i) {
      const size_t l = R(10) ? 20 + R(40) : R(300);
      inner.push_back(0x0A);
      put_varint(inner, l);
      for (size_t k = 0; k < l; ++k) inner.push_back(R(8) ? (uint8_t)('a' + R(26)) : (uint8_t)R(256));
    }
    out.push_back(0x0A);
    put_varint(out, inner.size());
    out.insert(out.end(), inner.begin(), inner.end());
  }
  // mutations: truncate, flip, insert
  if (!out.empty() && R(3) == 0) out.resize(R(out.size()));
  for (int k = (int)R(4); k > 0 && !out.empty(); --k) out[R(out.size())] = (uint8_t)R(256);
  return out;
}

int main() {
  long compared = 0;
  // Go-semantics helpers on arbitrary strings
  for (int i = 0; i < 50000; ++i) {
    const std::string s = rand_str("0123456789+-.eEinumkKMGTP _x", 40);
    int64_t v;
    uint64_t u;
    gs_quantity_value(s.data(), s.size(), &v);
    gs_go_atoi(s.data(), s.size(), &v);
    gs_go_parse_uint64(s.data(), s.size(), &u);
  }
  // tables of random shapes, health marks, payload encodes, id lookups
  for (int t = 0; t < 60; ++t) {
    const uint32_t n = (uint32_t)R(9);
    std::vector<std::string> uuids;
    for (uint32_t g = 0; g < n; ++g) uuids.push_back("GPU-" + std::to_string(g) + rand_str("abcdef0123456789-", 80));
    std::vector<const char *> p;
    std::vector<uint32_t> l;
    for (auto &u : uuids) { p.push_back(u.data()); l.push_back((uint32_t)u.size()); }
    const uint64_t units = R(4) ? R(300) : R(20000);
    gs_table *tab = nullptr;
    if (gs_table_build(p.data(), l.data(), n, units, (int)R(2), &tab) != GS_OK) continue;
    for (int k = 0; k < 20 && n && units; ++k) {
      if (R(3) == 0) gs_law_mark_unhealthy_gpu(tab, (int32_t)R(n + 1) - 1);
      else gs_law_mark_unhealthy_device(tab, (uint32_t)R(n), R(units));
      const size_t need = gs_law_payload_size(tab);
      std::vector<uint8_t> buf(need);
      size_t len = 0;
      if (gs_law_encode(tab, buf.data(), buf.size(), &len) != GS_OK || len != need) { printf("FAIL law size\n"); return 1; }
      std::vector<gs_iovec> iov(n);
      uint32_t cnt = 0;
      gs_law_payload(tab, iov.data(), n, &cnt);
      size_t sum = 0;
      for (uint32_t g = 0; g < cnt; ++g) sum += iov[g].len;
      if (sum != need) { printf("FAIL segments\n"); return 1; }
      char id[200];
      size_t il = 0;
      const uint32_t g = (uint32_t)R(n);
      const uint64_t j = R(units);
      gs_fake_id_format(tab, g, j, id, sizeof id, &il);
      uint32_t g2 = 99; uint64_t j2 = 0;
      if (gs_id_lookup(tab, id, il, &g2, &j2) != GS_OK || g2 != g || j2 != j) { printf("FAIL lookup\n"); return 1; }
      const std::string junk = rand_str("GPU-_0123456789abcdef", 60);
      gs_id_lookup(tab, junk.data(), junk.size(), &g2, &j2);
      if (R(6) == 0) gs_law_reset_health(tab);
    }
    // decisions over random pods and random (often malformed) requests
    gs_plugin *plug = nullptr;
    const int32_t plugin_cgpu = (int32_t)R(2);
    gs_plugin_new(tab, "node-a", 6, plugin_cgpu, &plug);
    for (int k = 0; k < 30; ++k) {
      const size_t np = R(R(5) ? 20 : 200);
      std::vector<std::string> keep;
      keep.reserve(np * 8);
      std::vector<gs_pod_view> pods(np);
      auto S = [&](const std::string &s, bool present = true) {
        if (!present) return gs_str{nullptr, 0};
        keep.push_back(s);
        return gs_str{keep.back().data(), keep.back().size()};
      };
      for (size_t i = 0; i < np; ++i) {
        pods[i].uid = S("u" + std::to_string(R(np + 2)));
        pods[i].ns = S("default");
        pods[i].name = S("p" + std::to_string(i));
        pods[i].node_name = S(R(8) ? "node-a" : "node-b");
        pods[i].gpu_mem_limit = R(6);
        pods[i].ann_idx = S(rand_str("0123456789-+x", 4), R(8) != 0);
        pods[i].ann_assume_time = S(rand_str("0123456789", 12), R(10) != 0);
        pods[i].ann_assigned = S(R(6) ? "false" : "true", R(12) != 0);
      }
      gs_plugin_set_pods(plug, pods.data(), np);
      for (int q = 0; q < 8; ++q) {
        const std::vector<uint8_t> req = rand_request();
        uint8_t resp[2048];
        gs_alloc_result res;
        // what the stateless decision says about the plugin's pods as they are right now (after
        // earlier claims): the plugin, which filters its list once and drops claimed pods, must
        // answer the same -- kind, pod, GPU and response bytes
        size_t cur_n = 0;
        gs_plugin_pod_count(plug, &cur_n);
        std::vector<gs_pod_view> cur(cur_n);
        for (size_t i = 0; i < cur_n; ++i) gs_plugin_get_pod(plug, i, &cur[i]);
        gs_plugin_set_clock(plug, 12345);
        gs_alloc_ctx now_ctx{tab, gs_str{"node-a", 6}, plugin_cgpu, 12345};
        uint8_t want_resp[2048];
        char want_patch[192];
        gs_alloc_result want;
        const int want_rc = gs_allocate(&now_ctx, req.data(), req.size(), cur.data(), cur_n, want_resp,
                                        sizeof want_resp, want_patch, sizeof want_patch, &want);
        const int rc = gs_plugin_allocate(plug, req.data(), req.size(), resp, sizeof resp, &res);
        if (rc == GS_OK && res.resp_len > sizeof resp) { printf("FAIL resp_len\n"); return 1; }
        if (rc != want_rc || (rc == GS_OK && (res.kind != want.kind || res.pod_index != want.pod_index ||
                                              res.gpu_index != want.gpu_index || res.resp_len != want.resp_len ||
                                              memcmp(resp, want_resp, res.resp_len) != 0))) {
          printf("FAIL plugin and stateless decision differ (rc %d/%d kind %d/%d pod %d/%d)\n", rc, want_rc,
                 res.kind, want.kind, res.pod_index, want.pod_index);
          return 1;
        }
        ++compared;
        char patch[8];  // deliberately too small: must report, not overrun
        gs_alloc_ctx ctx{tab, gs_str{"node-a", 6}, 0, (int64_t)rng()};
        uint8_t small[16];
        gs_allocate(&ctx, req.data(), req.size(), pods.data(), np, small, sizeof small, patch, sizeof patch, &res);
      }
    }
    gs_plugin_free(plug);
    gs_table_free(tab);
  }
  printf("plugin vs stateless decisions compared: %ld\n", compared);
  printf("FUZZ CORE OK\n");
  return 0;
}

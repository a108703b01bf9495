// allocate.cpp -- the Allocate decision, allocation-free.
//
// Replaces (reference paths relative to cjg/aliyun-gpushare-device-plugin):
//   Allocate                          pkg/gpu/nvidia/allocate.go:43-201
//   buildErrResponse                  allocate.go:25-40
//   getCandidatePods                  podmanager.go:300-323
//   getPendingPodsInNode (filtering)  podmanager.go:273-293
//   makePodOrderdByAge / Less `<=`    podmanager.go:326-347
//   isGPUMemoryAssumedPod             podutils.go:78-119
//   getGPUIDFromPodAnnotation         podutils.go:37-61
//   getAssumeTimeFromPodAnnotation    podutils.go:64-75
//   patchPodAnnotationSpecAssigned    podutils.go:27-35
//
// Where the reference spends its time: gogo unmarshals one Go string
// per device id (<= 40 960 per request at MiB granularity) only to take len(); the
// candidate list is rebuilt, re-parsed and re-sorted per call; envs are formatted
// into fresh maps. Here the request is scanned in place (ids are hopped over by
// their length prefix, never copied), the candidate order is computed on a small
// array of (assume time, pod index) pairs, and the response is written straight
// into the caller's buffer.
#include <string.h>

#include <vector>

#include "gosort.h"
#include "gpushare.h"
#include "internal.h"
#include "table.h"
#include "wire.h"

namespace gs_internal {
bool parse_u64_digits(const char *s, size_t n, uint64_t *out, bool *overflow);
}

namespace {

constexpr uint32_t kMaxContainers = 1024;

struct Scan {
  uint32_t n_containers = 0;
  uint64_t total = 0;
};

// Counting ids is a chain of dependent hops (read a length byte, jump): one hop costs an
// L1 load-to-use latency however little work it is. For a large container the chain is cut
// into kChains pieces that run interleaved so the loads overlap (measured on the B200 box's
// Xeon, tools/exp/scan_bench.cpp: 2.3-2.5 ns per id sequential, 1.3-1.4 ns with four chains
// and a prefetch, at 1 024 as at 40 960 ids, i.e. 21 -> 38 GB/s of request bytes). A piece starts at a
// GUESSED record boundary (the first byte after its nominal start that looks like
// "0x0A <len < 128>" and whose hop lands on another such byte or on the end); the guess
// is proven, not trusted: the previous piece must land EXACTLY on it. Any surprise (an
// unknown field, a long length, a miss) makes the caller fall back to the plain
// sequential walk, so the count is exact for every input.
constexpr size_t kMultiChainMin = 4096;
constexpr int kChains = 4;

inline bool fast_record(const uint8_t *q, const uint8_t *qend) {
  return qend - q >= 2 && q[0] == 0x0A && q[1] < 0x80 && (size_t)(qend - q) >= 2u + q[1];
}

bool count_ids_multichain(const uint8_t *q0, const uint8_t *qend, uint32_t *count) {
  const size_t len = (size_t)(qend - q0);
  const uint8_t *start[kChains + 1];
  start[0] = q0;
  start[kChains] = qend;
  for (int c = 1; c < kChains; ++c) {
    const uint8_t *g = q0 + len / kChains * c;
    const uint8_t *lim = g + 512 < qend ? g + 512 : qend;
    const uint8_t *found = nullptr;
    for (; g < lim; ++g) {
      if (!fast_record(g, qend)) continue;
      const uint8_t *n1 = g + 2 + g[1];
      if (n1 != qend && !fast_record(n1, qend)) continue;
      found = g;
      break;
    }
    if (!found || found <= start[c - 1]) return false;
    start[c] = found;
  }
  const uint8_t *p[kChains];
  uint32_t cnt[kChains];
  for (int c = 0; c < kChains; ++c) {
    p[c] = start[c];
    cnt[c] = 0;
  }
  // lockstep while every piece has work left: kChains independent hops per iteration
  for (;;) {
    bool all = true;
    for (int c = 0; c < kChains; ++c) all &= p[c] < start[c + 1];
    if (!all) break;
#pragma GCC unroll 4
    for (int c = 0; c < kChains; ++c) {
      __builtin_prefetch(p[c] + 256);
      if (!fast_record(p[c], qend)) return false;
      p[c] += 2u + p[c][1];
      ++cnt[c];
    }
  }
  uint32_t total = 0;
  for (int c = 0; c < kChains; ++c) {
    while (p[c] < start[c + 1]) {  // the longer pieces finish on their own
      if (!fast_record(p[c], qend)) return false;
      p[c] += 2u + p[c][1];
      ++cnt[c];
    }
    if (p[c] != start[c + 1]) return false;  // the next piece's start was not a boundary
    total += cnt[c];
  }
  *count = total;
  return true;
}

// AllocateRequest{repeated ContainerAllocateRequest container_requests = 1}
// ContainerAllocateRequest{repeated string devices_ids = 1}
// Counts field-1 entries of each container; the id bytes are never read (H9).
int scan_request(const uint8_t *req, size_t req_len, uint32_t *counts, uint32_t cap, Scan *out) {
  const uint8_t *p = req, *end = req + req_len;
  uint32_t nc = 0;
  uint64_t total = 0;
  while (p < end) {
    uint64_t key;
    p = gs_wire::get_varint(p, end, &key);
    if (!p) return GS_E_WIRE;
    const uint32_t wt = (uint32_t)(key & 7);
    if ((key >> 3) == 1 && wt == gs_wire::kLen) {
      uint64_t clen;
      p = gs_wire::get_varint(p, end, &clen);
      if (!p || clen > (uint64_t)(end - p)) return GS_E_WIRE;
      const uint8_t *q = p, *qend = p + clen;
      uint32_t cnt = 0;
      if (clen >= kMultiChainMin && count_ids_multichain(q, qend, &cnt)) q = qend;
      while (q < qend) {
        // fast path: tag 0x0A (field 1, LEN) followed by a one-byte length
        if (q[0] == 0x0A && qend - q >= 2 && q[1] < 0x80) {
          const size_t adv = 2u + q[1];
          if (adv > (size_t)(qend - q)) return GS_E_WIRE;
          q += adv;
          ++cnt;
          continue;
        }
        uint64_t k2;
        q = gs_wire::get_varint(q, qend, &k2);
        if (!q) return GS_E_WIRE;
        const uint32_t wt2 = (uint32_t)(k2 & 7);
        if ((k2 >> 3) == 1 && wt2 == gs_wire::kLen) ++cnt;
        q = gs_wire::skip_value(q, qend, wt2);
        if (!q) return GS_E_WIRE;
      }
      if (nc < cap && counts) counts[nc] = cnt;
      ++nc;
      total += cnt;
      p = qend;
    } else {
      p = gs_wire::skip_value(p, end, wt);
      if (!p) return GS_E_WIRE;
    }
  }
  out->n_containers = nc;
  out->total = total;
  return GS_OK;
}

bool str_eq(const gs_str &a, const gs_str &b) {
  return a.n == b.n && (a.n == 0 || memcmp(a.p, b.p, a.n) == 0);
}
bool str_is(const gs_str &a, const char *lit, size_t n) {
  return a.p && a.n == n && memcmp(a.p, lit, n) == 0;
}

uint64_t hash_bytes(const char *p, size_t n) {
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < n; ++i) {
    h ^= (unsigned char)p[i];
    h *= 1099511628211ull;
  }
  return h;
}

// getAssumeTimeFromPodAnnotation (podutils.go:64-75): parse failure leaves 0.
uint64_t assume_time(const gs_pod_view &pod) {
  if (!pod.ann_assume_time.p) return 0;
  uint64_t v = 0;
  bool ovf = false;
  if (!gs_internal::parse_u64_digits(pod.ann_assume_time.p, pod.ann_assume_time.n, &v, &ovf) || ovf)
    return 0;
  return v;
}

// isGPUMemoryAssumedPod (podutils.go:78-119)
bool is_assumed(const gs_pod_view &pod) {
  if (pod.gpu_mem_limit == 0) return false;                 // :84
  if (!pod.ann_assume_time.p) return false;                 // :92
  return str_is(pod.ann_assigned, "false", 5);              // :100-117
}

// getGPUIDFromPodAnnotation (podutils.go:37-61)
int64_t gpu_id_from_annotation(const gs_pod_view &pod) {
  if (!pod.ann_idx.p) return -1;
  int64_t v = 0;
  if (gs_go_atoi(pod.ann_idx.p, pod.ann_idx.n, &v) != GS_OK) return -1;
  return v;
}

size_t put_dec_u64(char *p, uint64_t v) {
  char tmp[20];
  size_t n = 0;
  do {
    tmp[n++] = (char)('0' + v % 10);
    v /= 10;
  } while (v);
  for (size_t i = 0; i < n; ++i) p[i] = tmp[n - 1 - i];
  return n;
}
size_t put_dec_i64(char *p, int64_t v) {
  if (v < 0) {
    *p = '-';
    return 1 + put_dec_u64(p + 1, 0 - (uint64_t)v);
  }
  return put_dec_u64(p, (uint64_t)v);
}

struct EnvVal {
  const char *p;
  size_t n;
};

// map<string,string> entry: ContainerAllocateResponse.envs = 1 -> {key = 1, value = 2}
size_t env_entry_size(size_t kn, size_t vn) {
  const size_t inner = gs_wire::field_bytes_size(1, kn) + gs_wire::field_bytes_size(2, vn);
  return gs_wire::field_bytes_size(1, inner);
}
void put_env(gs_wire::Writer &w, const char *k, size_t kn, const char *v, size_t vn) {
  const size_t inner = gs_wire::field_bytes_size(1, kn) + gs_wire::field_bytes_size(2, vn);
  w.varint((1u << 3) | gs_wire::kLen);
  w.varint(inner);
  w.field_bytes(1, k, kn);
  w.field_bytes(2, v, vn);
}

#define LIT(s) s, sizeof(s) - 1

// One ContainerAllocateResponse per container request, in request order
// (allocate.go:27-37,114-129,163-178). Only envs are ever populated.
void encode_response(gs_wire::Writer &w, const uint32_t *counts, uint32_t n_containers,
                     const EnvVal &visible, const EnvVal &idx, const EnvVal &pod_total,
                     const EnvVal &dev_units, bool cgpu_disable) {
  for (uint32_t c = 0; c < n_containers; ++c) {
    char cbuf[20];
    const size_t cn = put_dec_u64(cbuf, counts[c]);
    size_t inner = env_entry_size(sizeof(GS_ENV_NVGPU) - 1, visible.n) +
                   env_entry_size(sizeof(GS_ENV_RESOURCE_INDEX) - 1, idx.n) +
                   env_entry_size(sizeof(GS_ENV_RESOURCE_BY_POD) - 1, pod_total.n) +
                   env_entry_size(sizeof(GS_ENV_RESOURCE_BY_CONTAINER) - 1, cn) +
                   env_entry_size(sizeof(GS_ENV_RESOURCE_BY_DEV) - 1, dev_units.n);
    if (cgpu_disable) inner += env_entry_size(sizeof(GS_ENV_CGPU_DISABLE) - 1, 4);
    w.varint((1u << 3) | gs_wire::kLen);  // AllocateResponse.container_responses = 1
    w.varint(inner);
    put_env(w, LIT(GS_ENV_NVGPU), visible.p, visible.n);
    put_env(w, LIT(GS_ENV_RESOURCE_INDEX), idx.p, idx.n);
    put_env(w, LIT(GS_ENV_RESOURCE_BY_POD), pod_total.p, pod_total.n);
    put_env(w, LIT(GS_ENV_RESOURCE_BY_CONTAINER), cbuf, cn);
    put_env(w, LIT(GS_ENV_RESOURCE_BY_DEV), dev_units.p, dev_units.n);
    if (cgpu_disable) put_env(w, LIT(GS_ENV_CGPU_DISABLE), "true", 4);
  }
}

// buildErrResponse (allocate.go:25-40): no CGPU_DISABLE on this path.
void encode_error(gs_wire::Writer &w, const gs_table *t, const uint32_t *counts,
                  uint32_t n_containers, uint64_t pod_req) {
  char vis[64], tot[20], dev[20];
  size_t vn = 0;
  memcpy(vis, "no-gpu-has-", 11);
  vn = 11;
  vn += put_dec_u64(vis + vn, pod_req);
  const char *un = gs_unit_name(gs_table_unit(t));
  memcpy(vis + vn, un, 3);
  vn += 3;
  memcpy(vis + vn, "-to-run", 7);
  vn += 7;
  const size_t tn = put_dec_u64(tot, pod_req);
  const size_t dn = put_dec_u64(dev, gs_table_units_per_gpu(t));
  encode_response(w, counts, n_containers, EnvVal{vis, vn}, EnvVal{"-1", 2}, EnvVal{tot, tn},
                  EnvVal{dev, dn}, false);
}

// patchPodAnnotationSpecAssigned (podutils.go:27-35); json.Marshal sorts map keys:
// ..._ASSIGNED < ..._ASSUME_TIME.
size_t encode_patch(char *buf, size_t cap, int64_t now_unix_nano) {
  static const char head[] =
      "{\"metadata\":{\"annotations\":{\"" GS_ENV_ASSIGNED_FLAG "\":\"true\",\"" GS_ENV_ASSUME_TIME
      "\":\"";
  static const char tail[] = "\"}}}";
  char num[24];
  const size_t nn = put_dec_i64(num, now_unix_nano);
  const size_t need = sizeof(head) - 1 + nn + sizeof(tail) - 1;
  if (buf && cap >= need) {
    memcpy(buf, head, sizeof(head) - 1);
    memcpy(buf + sizeof(head) - 1, num, nn);
    memcpy(buf + sizeof(head) - 1 + nn, tail, sizeof(tail) - 1);
  }
  return need;
}

using gs_internal::Cand;

// scratch that grows to the largest pod list seen, then is reused: no allocation in
// steady state
thread_local std::vector<uint32_t> tl_slots;
thread_local std::vector<uint64_t> tl_hashes;
thread_local std::vector<Cand> tl_cands;

}  // namespace

extern "C" {

int gs_request_scan(const uint8_t *req, size_t req_len, uint32_t *counts, uint32_t cap,
                    uint32_t *n_containers, uint64_t *total) {
  if (!req && req_len) return GS_E_INVAL;
  Scan s;
  const int rc = scan_request(req, req_len, counts, cap, &s);
  if (rc != GS_OK) return rc;
  if (n_containers) *n_containers = s.n_containers;
  if (total) *total = s.total;
  return counts && s.n_containers > cap ? GS_E_NOSPACE : GS_OK;
}

int gs_allocate_error_response(const gs_alloc_ctx *ctx, const uint8_t *req, size_t req_len,
                               int32_t kind, uint8_t *resp, size_t resp_cap,
                               gs_alloc_result *out) {
  if (!ctx || !ctx->table || !out || (!req && req_len)) return GS_E_INVAL;
  uint32_t counts[kMaxContainers];
  Scan s;
  int rc = scan_request(req, req_len, counts, kMaxContainers, &s);
  if (rc != GS_OK) return rc;
  if (s.n_containers > kMaxContainers) return GS_E_LIMIT;
  gs_wire::Writer w(resp, resp_cap);
  encode_error(w, ctx->table, counts, s.n_containers, s.total);
  memset(out, 0, sizeof(*out));
  out->kind = kind;
  out->pod_index = -1;
  out->gpu_index = -1;
  out->n_containers = s.n_containers;
  out->pod_req_units = s.total;
  out->resp_len = w.len;
  return w.ok() ? GS_OK : GS_E_NOSPACE;
}

int gs_allocate(const gs_alloc_ctx *ctx, const uint8_t *req, size_t req_len,
                const gs_pod_view *pods, size_t n_pods, uint8_t *resp, size_t resp_cap,
                char *patch, size_t patch_cap, gs_alloc_result *out) {
  if (!ctx || !ctx->table || !out || (!req && req_len) || (!pods && n_pods)) return GS_E_INVAL;
  return gs_internal::allocate_with_candidates(ctx, req, req_len, pods, n_pods, nullptr, 0, resp, resp_cap,
                                               patch, patch_cap, out);
}

}  // extern "C"

namespace gs_internal {

void build_candidates(gs_str node_name, const gs_pod_view *pods, size_t n_pods, std::vector<Cand> &cands) {
  // podmanager.go:273-293 -- this node's pods, first occurrence per UID;
  // podmanager.go:306-311 -- keep the assumed-but-unassigned ones
  cands.clear();
  if (n_pods <= 16) {
    for (size_t i = 0; i < n_pods; ++i) {
      if (!str_eq(pods[i].node_name, node_name)) continue;
      bool dup = false;
      for (size_t k = 0; k < i && !dup; ++k)
        dup = str_eq(pods[k].node_name, node_name) && str_eq(pods[k].uid, pods[i].uid);
      if (dup || !is_assumed(pods[i])) continue;
      cands.push_back(Cand{assume_time(pods[i]), (uint32_t)i});
    }
  } else {
    size_t cap = 32;
    while (cap < 2 * n_pods) cap <<= 1;
    tl_slots.assign(cap, UINT32_MAX);
    tl_hashes.resize(n_pods);
    for (size_t i = 0; i < n_pods; ++i) {
      if (!str_eq(pods[i].node_name, node_name)) continue;
      const uint64_t h = hash_bytes(pods[i].uid.p, pods[i].uid.n);
      tl_hashes[i] = h;
      size_t k = (size_t)h & (cap - 1);
      bool dup = false;
      while (tl_slots[k] != UINT32_MAX) {
        const uint32_t o = tl_slots[k];
        if (tl_hashes[o] == h && str_eq(pods[o].uid, pods[i].uid)) {
          dup = true;
          break;
        }
        k = (k + 1) & (cap - 1);
      }
      if (dup) continue;
      tl_slots[k] = (uint32_t)i;
      if (is_assumed(pods[i])) cands.push_back(Cand{assume_time(pods[i]), (uint32_t)i});
    }
  }
}

int allocate_with_candidates(const gs_alloc_ctx *ctx, const uint8_t *req, size_t req_len,
                             const gs_pod_view *pods, size_t n_pods, const Cand *known, size_t n_known,
                             uint8_t *resp, size_t resp_cap, char *patch, size_t patch_cap,
                             gs_alloc_result *out) {
  const gs_table *t = ctx->table;
  memset(out, 0, sizeof(*out));
  out->pod_index = -1;
  out->gpu_index = -1;

  // allocate.go:55-57 -- podReqGPU = sum of len(req.DevicesIDs)
  uint32_t counts[kMaxContainers];
  Scan s;
  int rc = scan_request(req, req_len, counts, kMaxContainers, &s);
  if (rc != GS_OK) return rc;
  if (s.n_containers > kMaxContainers) return GS_E_LIMIT;
  const uint64_t pod_req = s.total;
  out->n_containers = s.n_containers;
  out->pod_req_units = pod_req;

  std::vector<Cand> &cands = tl_cands;
  if (known)
    cands.assign(known, known + n_known);  // the sort below works on a copy: the caller's order stays lister order
  else
    build_candidates(ctx->node_name, pods, n_pods, cands);

  // podmanager.go:326-347 -- sort.Sort with Less(i,j) = time[i] <= time[j]
  {
    Cand *c = cands.data();
    gs_gosort::sort(
        (int)cands.size(), [c](int i, int j) { return c[i].time <= c[j].time; },
        [c](int i, int j) {
          const Cand tmp = c[i];
          c[i] = c[j];
          c[j] = tmp;
        });
  }

  // allocate.go:79-89 -- first candidate whose limits sum equals the request
  int32_t found = -1;
  for (const Cand &c : cands) {
    if (pods[c.pod].gpu_mem_limit == pod_req) {
      found = (int32_t)c.pod;
      break;
    }
  }

  gs_wire::Writer w(resp, resp_cap);
  char tot[20], dev[20];
  const size_t tn = put_dec_u64(tot, pod_req);
  const size_t dn = put_dec_u64(dev, gs_table_units_per_gpu(t));

  if (found >= 0) {
    out->pod_index = found;
    int64_t id = gpu_id_from_annotation(pods[found]);           // allocate.go:92
    if (id >= 0 && (uint64_t)id >= gs_table_gpu_count(t)) id = -1;  // :100-106
    if (id < 0) {                                               // :109-111
      encode_error(w, t, counts, s.n_containers, pod_req);
      out->kind = GS_ALLOC_ERR_BAD_INDEX;
      out->resp_len = w.len;
      return w.ok() ? GS_OK : GS_E_NOSPACE;
    }
    char idb[20];
    const size_t in = put_dec_u64(idb, (uint64_t)id);
    encode_response(w, counts, s.n_containers, EnvVal{idb, in}, EnvVal{idb, in}, EnvVal{tot, tn},
                    EnvVal{dev, dn}, ctx->disable_cgpu_isolation != 0);  // :114-129
    out->kind = GS_ALLOC_MATCHED;
    out->gpu_index = (int32_t)id;
    out->resp_len = w.len;
    out->patch_len = encode_patch(patch, patch_cap, ctx->now_unix_nano);  // :132
    if (!w.ok() || !patch || patch_cap < out->patch_len) return GS_E_NOSPACE;
    return GS_OK;
  }

  if (gs_table_gpu_count(t) == 1) {                             // allocate.go:154-180
    const std::string &uuid = gs_internal::table_uuid(t, 0);
    encode_response(w, counts, s.n_containers, EnvVal{uuid.data(), uuid.size()}, EnvVal{"0", 1},
                    EnvVal{tot, tn}, EnvVal{dev, dn}, ctx->disable_cgpu_isolation != 0);
    out->kind = GS_ALLOC_SINGLE_GPU;
    out->gpu_index = 0;
    out->resp_len = w.len;
    return w.ok() ? GS_OK : GS_E_NOSPACE;
  }

  encode_error(w, t, counts, s.n_containers, pod_req);          // allocate.go:182-187
  out->kind = GS_ALLOC_ERR_NO_MATCH;
  out->resp_len = w.len;
  return w.ok() ? GS_OK : GS_E_NOSPACE;
}

}  // namespace gs_internal

// table.cpp -- device model: fake-device expansion, O(1) id lookup, ListAndWatch payload.
//
// Replaces (reference paths relative to cjg/aliyun-gpushare-device-plugin):
//   generateFakeDeviceID / extractRealDeviceID   pkg/gpu/nvidia/nvidia.go:23-29
//   setGPUMemory / getGPUMemory                  nvidia.go:31-42
//   getDevices                                   nvidia.go:50-86
//   devNameMap / GetDeviceNameByIndex            server.go:22,80-91
//   ListAndWatch's repeated full-list marshal    server.go:180-193
//
// The reference materialises one heap Device{ID, Health} per memory unit (1.47 M at
// MiB granularity on 8 x B200) and re-marshals all of them on every send. Here the
// ids exist only as (uuid, unit) arithmetic plus one pre-encoded wire segment per
// GPU; health is a bitmap with per-GPU counters, and a send is n_gpus borrowed
// buffers. Wire layout per device (deviceplugin/v1beta1):
//   0A <len> { 0A <len(ID)> ID  12 <len(health)> health }   = 13 + len(ID) bytes
//   for "Healthy" (15 + len(ID) for "Unhealthy").
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "gpushare.h"
#include "internal.h"
#include "table.h"
#include "wire.h"

namespace {

constexpr char kSep[] = "-_-";
constexpr size_t kSepLen = 3;
constexpr char kHealthy[] = GS_HEALTHY;
constexpr char kUnhealthy[] = GS_UNHEALTHY;

uint64_t fnv1a(const char *p, size_t n) {
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < n; ++i) {
    h ^= (unsigned char)p[i];
    h *= 1099511628211ull;
  }
  return h;
}

size_t dec_digits(uint64_t v) {
  size_t n = 1;
  while (v >= 10) {
    v /= 10;
    ++n;
  }
  return n;
}

// Writes v in decimal, returns chars written.
size_t put_dec(char *p, uint64_t v) {
  char tmp[20];
  size_t n = 0;
  do {
    tmp[n++] = (char)('0' + v % 10);
    v /= 10;
  } while (v);
  for (size_t i = 0; i < n; ++i) p[i] = tmp[n - 1 - i];
  return n;
}

size_t device_wire_size(size_t id_len, bool healthy) {
  const size_t hl = healthy ? sizeof(kHealthy) - 1 : sizeof(kUnhealthy) - 1;
  const size_t inner = gs_wire::field_bytes_size(1, id_len) + gs_wire::field_bytes_size(2, hl);
  return 1 + gs_wire::varint_size(inner) + inner;
}

// Encodes devices [j0, j1) of one GPU. `bad` == nullptr means uniform health
// `uniform_healthy`; otherwise bit j of bad[] set means unhealthy.
// The id's decimal suffix is kept as ASCII and incremented in place.
uint8_t *encode_range(uint8_t *out, const std::string &uuid, uint64_t j0, uint64_t j1,
                      bool uniform_healthy, const std::atomic<uint64_t> *bad) {
  if (j0 >= j1) return out;
  char digits[24];
  size_t nd = put_dec(digits, j0);
  const size_t base_len = uuid.size() + kSepLen;
  for (uint64_t j = j0; j < j1; ++j) {
    bool healthy = uniform_healthy;
    if (bad) healthy = !((bad[j >> 6].load(std::memory_order_relaxed) >> (j & 63)) & 1u);
    const char *hs = healthy ? kHealthy : kUnhealthy;
    const size_t hl = healthy ? sizeof(kHealthy) - 1 : sizeof(kUnhealthy) - 1;
    const size_t id_len = base_len + nd;
    const size_t inner = gs_wire::field_bytes_size(1, id_len) + gs_wire::field_bytes_size(2, hl);
    *out++ = 0x0A;  // ListAndWatchResponse.devices = 1, LEN
    out = gs_wire::put_varint(out, inner);
    *out++ = 0x0A;  // Device.ID = 1, LEN
    out = gs_wire::put_varint(out, id_len);
    memcpy(out, uuid.data(), uuid.size());
    out += uuid.size();
    memcpy(out, kSep, kSepLen);
    out += kSepLen;
    memcpy(out, digits, nd);
    out += nd;
    *out++ = 0x12;  // Device.health = 2, LEN
    *out++ = (uint8_t)hl;
    memcpy(out, hs, hl);
    out += hl;
    // ++digits (ASCII)
    size_t k = nd;
    while (k > 0 && digits[k - 1] == '9') digits[--k] = '0';
    if (k == 0) {
      memmove(digits + 1, digits, nd);
      digits[0] = '1';
      ++nd;
    } else {
      ++digits[k - 1];
    }
  }
  return out;
}

size_t uniform_segment_size(size_t uuid_len, uint64_t units, bool healthy) {
  // sum over j of device_wire_size(uuid_len + 3 + digits(j)), by digit-count bands
  size_t total = 0;
  uint64_t lo = 0, pow10 = 10;
  size_t nd = 1;
  while (lo < units) {
    const uint64_t hi = units < pow10 ? units : pow10;  // band [lo, hi) has nd digits
    total += (size_t)(hi - lo) * device_wire_size(uuid_len + kSepLen + nd, healthy);
    lo = hi;
    if (pow10 > UINT64_MAX / 10) break;
    pow10 *= 10;
    ++nd;
  }
  return total;
}

}  // namespace

struct gs_table {
  uint32_t n_gpus = 0;
  uint64_t units = 0;
  int unit = GS_UNIT_GIB;
  std::vector<std::string> uuids;
  // uuid -> gpu, open addressing
  std::vector<int32_t> slots;
  uint64_t slot_mask = 0;
  // health
  uint64_t words_per_gpu = 0;
  std::unique_ptr<std::atomic<uint64_t>[]> bad;        // n_gpus * words_per_gpu bits
  std::unique_ptr<std::atomic<uint64_t>[]> bad_count;  // per GPU
  std::atomic<uint64_t> generation{0};
  // wire segments
  std::vector<std::vector<uint8_t>> healthy_seg;  // built once
  std::vector<size_t> healthy_size;
  std::vector<size_t> unhealthy_size;
  mutable std::mutex seg_mu;  // serialises health changes against payload encodes
  std::vector<std::vector<uint8_t>> current_seg;  // lazily built non-healthy forms
  std::vector<uint64_t> current_gen;              // generation the cached form was built at
  std::vector<uint8_t> current_valid;
  std::unique_ptr<std::atomic<uint64_t>[]> gpu_gen;  // bumps when that GPU's health changes
};

namespace gs_internal {

const std::string &table_uuid(const gs_table *t, uint32_t gpu) { return t->uuids[gpu]; }

int table_find_uuid(const gs_table *t, const char *uuid, size_t len) {
  if (!t->slot_mask && t->slots.empty()) return -1;
  uint64_t i = fnv1a(uuid, len) & t->slot_mask;
  for (;;) {
    const int32_t g = t->slots[i];
    if (g < 0) return -1;
    const std::string &u = t->uuids[(size_t)g];
    if (u.size() == len && memcmp(u.data(), uuid, len) == 0) return g;
    i = (i + 1) & t->slot_mask;
  }
}

}  // namespace gs_internal

extern "C" {

int gs_table_build(const char *const *uuids, const uint32_t *uuid_lens, uint32_t n_gpus,
                   uint64_t units_per_gpu, int unit, gs_table **out) {
  if (!out) return GS_E_INVAL;
  *out = nullptr;
  if ((n_gpus && (!uuids || !uuid_lens)) || (unit != GS_UNIT_GIB && unit != GS_UNIT_MIB))
    return GS_E_INVAL;
  if (n_gpus > 4096) return GS_E_LIMIT;
  std::unique_ptr<gs_table> t(new (std::nothrow) gs_table());
  if (!t) return GS_E_INVAL;
  t->n_gpus = n_gpus;
  t->units = units_per_gpu;
  t->unit = unit;
  uint64_t cap = 4;
  while (cap < 2ull * n_gpus) cap <<= 1;
  t->slots.assign(cap, -1);
  t->slot_mask = cap - 1;
  for (uint32_t g = 0; g < n_gpus; ++g) {
    if (!uuids[g] || uuid_lens[g] == 0 || uuid_lens[g] >= GS_UUID_MAX) return GS_E_INVAL;
    // NVML never reports one UUID twice; the reference's map would silently merge them
    if (gs_internal::table_find_uuid(t.get(), uuids[g], uuid_lens[g]) >= 0) return GS_E_INVAL;
    t->uuids.emplace_back(uuids[g], uuid_lens[g]);
    uint64_t i = fnv1a(uuids[g], uuid_lens[g]) & t->slot_mask;
    while (t->slots[i] >= 0) i = (i + 1) & t->slot_mask;
    t->slots[i] = (int32_t)g;
  }
  t->words_per_gpu = (units_per_gpu + 63) / 64;
  const size_t nwords = (size_t)n_gpus * t->words_per_gpu;
  t->bad.reset(new std::atomic<uint64_t>[nwords ? nwords : 1]);
  for (size_t i = 0; i < (nwords ? nwords : 1); ++i) t->bad[i].store(0, std::memory_order_relaxed);
  t->bad_count.reset(new std::atomic<uint64_t>[n_gpus ? n_gpus : 1]);
  t->gpu_gen.reset(new std::atomic<uint64_t>[n_gpus ? n_gpus : 1]);
  t->healthy_seg.resize(n_gpus);
  t->healthy_size.resize(n_gpus);
  t->unhealthy_size.resize(n_gpus);
  t->current_seg.resize(n_gpus);
  t->current_gen.assign(n_gpus, 0);
  t->current_valid.assign(n_gpus, 0);
  std::atomic<bool> broken{false};
  gs_table *tp = t.get();
  auto build_gpu = [tp, units_per_gpu, &broken](uint32_t g) {
    tp->bad_count[g].store(0, std::memory_order_relaxed);
    tp->gpu_gen[g].store(0, std::memory_order_relaxed);
    const size_t sz = uniform_segment_size(tp->uuids[g].size(), units_per_gpu, true);
    tp->healthy_size[g] = sz;
    tp->unhealthy_size[g] = uniform_segment_size(tp->uuids[g].size(), units_per_gpu, false);
    tp->healthy_seg[g].resize(sz);
    uint8_t *end = encode_range(tp->healthy_seg[g].data(), tp->uuids[g], 0, units_per_gpu, true, nullptr);
    if ((size_t)(end - tp->healthy_seg[g].data()) != sz) broken.store(true);  // size model broken
  };
  // The segments are independent: above ~64 k devices (one B200 at MiB is 183 359) encode them
  // on one thread per GPU, up to eight. This is the restart path (kubelet.sock re-created).
  uint32_t workers = 1;
  if (n_gpus > 1 && units_per_gpu * n_gpus >= 65536) {
    workers = std::thread::hardware_concurrency();
    if (workers > 8) workers = 8;
    if (workers > n_gpus) workers = n_gpus;
    if (workers < 1) workers = 1;
  }
  if (workers <= 1) {
    for (uint32_t g = 0; g < n_gpus; ++g) build_gpu(g);
  } else {
    std::vector<std::thread> pool;
    for (uint32_t k = 0; k < workers; ++k)
      pool.emplace_back([k, workers, n_gpus, &build_gpu] {
        for (uint32_t g = k; g < n_gpus; g += workers) build_gpu(g);
      });
    for (auto &th : pool) th.join();
  }
  if (broken.load()) return GS_E_INVAL;
  *out = t.release();
  return GS_OK;
}

int gs_table_build_from_gpus(const gs_gpu_info *gpus, uint32_t n_gpus, int unit,
                             gs_table **out) {
  if (!out) return GS_E_INVAL;
  *out = nullptr;
  if (n_gpus && !gpus) return GS_E_INVAL;
  std::vector<const char *> ptrs(n_gpus);
  std::vector<uint32_t> lens(n_gpus);
  for (uint32_t g = 0; g < n_gpus; ++g) {
    ptrs[g] = gpus[g].uuid;
    lens[g] = (uint32_t)strnlen(gpus[g].uuid, GS_UUID_MAX);
  }
  // nvidia.go:67-69: the unit count is set once, from GPU 0 (H7). H1: MiB-valued raw.
  const uint64_t units = n_gpus ? gs_units_from_raw(gpus[0].total_bytes >> 20, unit) : 0;
  return gs_table_build(ptrs.data(), lens.data(), n_gpus, units, unit, out);
}

void gs_table_free(gs_table *t) { delete t; }
uint32_t gs_table_gpu_count(const gs_table *t) { return t ? t->n_gpus : 0; }
uint64_t gs_table_units_per_gpu(const gs_table *t) { return t ? t->units : 0; }
uint64_t gs_table_device_count(const gs_table *t) { return t ? t->units * t->n_gpus : 0; }
int gs_table_unit(const gs_table *t) { return t ? t->unit : GS_UNIT_GIB; }

int gs_fake_id_format(const gs_table *t, uint32_t gpu, uint64_t unit_index, char *buf,
                      size_t cap, size_t *len) {
  if (!t || !len) return GS_E_INVAL;
  if (gpu >= t->n_gpus) return GS_E_NOTFOUND;
  const std::string &u = t->uuids[gpu];
  const size_t need = u.size() + kSepLen + dec_digits(unit_index);
  *len = need;
  if (!buf || cap < need) return GS_E_NOSPACE;
  memcpy(buf, u.data(), u.size());
  memcpy(buf + u.size(), kSep, kSepLen);
  put_dec(buf + u.size() + kSepLen, unit_index);
  if (cap > need) buf[need] = '\0';
  return GS_OK;
}

size_t gs_extract_real_device_id(const char *id, size_t len) {
  if (!id) return 0;
  for (size_t i = 0; i + kSepLen <= len; ++i)
    if (id[i] == '-' && id[i + 1] == '_' && id[i + 2] == '-') return i;
  return len;
}

int gs_id_lookup(const gs_table *t, const char *id, size_t len, uint32_t *gpu,
                 uint64_t *unit_index) {
  if (!t || !id) return GS_E_INVAL;
  // advertised ids end in "-_-<canonical decimal>"; walk the digits from the tail
  size_t d = len;
  while (d > 0 && (unsigned)(id[d - 1] - '0') <= 9) --d;
  const size_t nd = len - d;
  if (nd == 0 || nd > 20 || d < kSepLen + 1) return GS_E_NOTFOUND;
  if (memcmp(id + d - kSepLen, kSep, kSepLen) != 0) return GS_E_NOTFOUND;
  if (nd > 1 && id[d] == '0') return GS_E_NOTFOUND;  // "%d" never pads
  uint64_t j = 0;
  for (size_t k = d; k < len; ++k) {
    const unsigned c = (unsigned)(id[k] - '0');
    if (j > (UINT64_MAX - c) / 10) return GS_E_NOTFOUND;
    j = j * 10 + c;
  }
  if (j >= t->units) return GS_E_NOTFOUND;
  const int g = gs_internal::table_find_uuid(t, id, d - kSepLen);
  if (g < 0) return GS_E_NOTFOUND;
  if (gpu) *gpu = (uint32_t)g;
  if (unit_index) *unit_index = j;
  return GS_OK;
}

int gs_device_name_by_index(const gs_table *t, uint64_t index, const char **uuid, size_t *len) {
  if (!t) return GS_E_INVAL;
  if (index >= t->n_gpus) return GS_E_NOTFOUND;
  if (uuid) *uuid = t->uuids[(size_t)index].data();
  if (len) *len = t->uuids[(size_t)index].size();
  return GS_OK;
}

int gs_device_index_by_name(const gs_table *t, const char *uuid, size_t len, uint32_t *index) {
  if (!t || !uuid) return GS_E_INVAL;
  const int g = gs_internal::table_find_uuid(t, uuid, len);
  if (g < 0) return GS_E_NOTFOUND;
  if (index) *index = (uint32_t)g;
  return GS_OK;
}

// ---- health ---------------------------------------------------------------------

int gs_law_mark_unhealthy_device(gs_table *t, uint32_t gpu, uint64_t unit_index) {
  if (!t) return GS_E_INVAL;
  if (gpu >= t->n_gpus || unit_index >= t->units) return GS_E_NOTFOUND;
  std::lock_guard<std::mutex> lk(t->seg_mu);
  std::atomic<uint64_t> &w = t->bad[(size_t)gpu * t->words_per_gpu + (unit_index >> 6)];
  const uint64_t bit = 1ull << (unit_index & 63);
  const uint64_t old = w.fetch_or(bit, std::memory_order_acq_rel);
  if (!(old & bit)) {
    t->bad_count[gpu].fetch_add(1, std::memory_order_acq_rel);
    t->gpu_gen[gpu].fetch_add(1, std::memory_order_acq_rel);
    t->generation.fetch_add(1, std::memory_order_acq_rel);
  }
  return GS_OK;
}

int gs_law_mark_unhealthy_id(gs_table *t, const char *id, size_t len) {
  uint32_t g = 0;
  uint64_t j = 0;
  const int rc = gs_id_lookup(t, id, len, &g, &j);
  if (rc != GS_OK) return rc;
  return gs_law_mark_unhealthy_device(t, g, j);
}

int gs_law_mark_unhealthy_gpu(gs_table *t, int32_t gpu) {
  if (!t) return GS_E_INVAL;
  if (gpu != GS_ALL_GPUS && (gpu < 0 || (uint32_t)gpu >= t->n_gpus)) return GS_E_NOTFOUND;
  const uint32_t g0 = gpu == GS_ALL_GPUS ? 0 : (uint32_t)gpu;
  const uint32_t g1 = gpu == GS_ALL_GPUS ? t->n_gpus : (uint32_t)gpu + 1;
  std::lock_guard<std::mutex> lk(t->seg_mu);
  bool changed = false;
  for (uint32_t g = g0; g < g1; ++g) {
    if (t->bad_count[g].load(std::memory_order_acquire) == t->units) continue;
    std::atomic<uint64_t> *w = &t->bad[(size_t)g * t->words_per_gpu];
    uint64_t added = 0;
    for (uint64_t k = 0; k < t->words_per_gpu; ++k) {
      uint64_t mask = ~0ull;
      if (k == t->words_per_gpu - 1 && (t->units & 63)) mask = (1ull << (t->units & 63)) - 1;
      const uint64_t old = w[k].fetch_or(mask, std::memory_order_acq_rel);
      added += (uint64_t)__builtin_popcountll(mask & ~old);
    }
    if (added) {
      t->bad_count[g].fetch_add(added, std::memory_order_acq_rel);
      t->gpu_gen[g].fetch_add(1, std::memory_order_acq_rel);
      changed = true;
    }
  }
  // one generation bump per event: one coalesced re-advertise (H13)
  if (changed) t->generation.fetch_add(1, std::memory_order_acq_rel);
  return GS_OK;
}

int gs_law_reset_health(gs_table *t) {
  if (!t) return GS_E_INVAL;
  std::lock_guard<std::mutex> lk(t->seg_mu);
  for (uint32_t g = 0; g < t->n_gpus; ++g) {
    for (uint64_t k = 0; k < t->words_per_gpu; ++k)
      t->bad[(size_t)g * t->words_per_gpu + k].store(0, std::memory_order_release);
    t->bad_count[g].store(0, std::memory_order_release);
    t->gpu_gen[g].fetch_add(1, std::memory_order_acq_rel);
  }
  t->generation.fetch_add(1, std::memory_order_acq_rel);
  return GS_OK;
}

int gs_law_device_healthy(const gs_table *t, uint32_t gpu, uint64_t unit_index) {
  if (!t || gpu >= t->n_gpus || unit_index >= t->units) return GS_E_NOTFOUND;
  const uint64_t w = t->bad[(size_t)gpu * t->words_per_gpu + (unit_index >> 6)].load(
      std::memory_order_acquire);
  return ((w >> (unit_index & 63)) & 1u) ? 0 : 1;
}

uint64_t gs_law_generation(const gs_table *t) {
  return t ? t->generation.load(std::memory_order_acquire) : 0;
}

// ---- payload ----------------------------------------------------------------------

size_t gs_law_payload_size(const gs_table *t) {
  if (!t) return 0;
  size_t total = 0;
  for (uint32_t g = 0; g < t->n_gpus; ++g) {
    const uint64_t nbad = t->bad_count[g].load(std::memory_order_acquire);
    if (nbad == t->units) {
      total += t->unhealthy_size[g];
    } else if (nbad == 0) {
      total += t->healthy_size[g];
    } else {
      // mixed: the healthy size plus, per unhealthy device, the exact size difference
      size_t sz = t->healthy_size[g];
      const std::atomic<uint64_t> *w = &t->bad[(size_t)g * t->words_per_gpu];
      const size_t uuid_len = t->uuids[g].size();
      for (uint64_t k = 0; k < t->words_per_gpu; ++k) {
        uint64_t bits = w[k].load(std::memory_order_acquire);
        while (bits) {
          const uint64_t j = k * 64 + (uint64_t)__builtin_ctzll(bits);
          bits &= bits - 1;
          const size_t idl = uuid_len + kSepLen + dec_digits(j);
          sz += device_wire_size(idl, false) - device_wire_size(idl, true);
        }
      }
      total += sz;
    }
  }
  return total;
}

static uint8_t *encode_gpu(const gs_table *t, uint32_t g, uint8_t *out) {
  const uint64_t nbad = t->bad_count[g].load(std::memory_order_acquire);
  if (nbad == 0) {
    memcpy(out, t->healthy_seg[g].data(), t->healthy_seg[g].size());
    return out + t->healthy_seg[g].size();
  }
  if (nbad == t->units) return encode_range(out, t->uuids[g], 0, t->units, false, nullptr);
  return encode_range(out, t->uuids[g], 0, t->units, true, &t->bad[(size_t)g * t->words_per_gpu]);
}

int gs_law_encode(const gs_table *t, uint8_t *buf, size_t cap, size_t *len) {
  if (!t || !len) return GS_E_INVAL;
  std::lock_guard<std::mutex> lk(t->seg_mu);  // health cannot change under the encode
  const size_t need = gs_law_payload_size(t);
  *len = need;
  if (!buf || cap < need) return GS_E_NOSPACE;
  uint8_t *p = buf;
  for (uint32_t g = 0; g < t->n_gpus; ++g) p = encode_gpu(t, g, p);
  return (size_t)(p - buf) == need ? GS_OK : GS_E_STATE;
}

int gs_law_payload(gs_table *t, gs_iovec *iov, uint32_t cap, uint32_t *n) {
  if (!t || !n) return GS_E_INVAL;
  *n = t->n_gpus;
  if (!iov || cap < t->n_gpus) return GS_E_NOSPACE;
  std::lock_guard<std::mutex> lk(t->seg_mu);
  std::vector<uint32_t> stale;  // GPUs whose cached segment predates their last health change
  for (uint32_t g = 0; g < t->n_gpus; ++g) {
    if (t->bad_count[g].load(std::memory_order_acquire) == 0) continue;
    const uint64_t gen = t->gpu_gen[g].load(std::memory_order_acquire);
    if (!t->current_valid[g] || t->current_gen[g] != gen) stale.push_back(g);
  }
  auto rebuild = [t](uint32_t g) {
    const uint64_t gen = t->gpu_gen[g].load(std::memory_order_acquire);
    std::vector<uint8_t> &seg = t->current_seg[g];
    seg.resize(t->unhealthy_size[g]);  // upper bound: every device unhealthy
    uint8_t *end = encode_gpu(t, g, seg.data());
    seg.resize((size_t)(end - seg.data()));
    t->current_gen[g] = gen;
    t->current_valid[g] = 1;
  };
  if (stale.size() > 1 && t->units >= 8192) {  // an event without a device takes every GPU down
    std::vector<std::thread> pool;
    const size_t workers = stale.size() < 8 ? stale.size() : 8;
    for (size_t k = 0; k < workers; ++k)
      pool.emplace_back([k, workers, &stale, &rebuild] {
        for (size_t i = k; i < stale.size(); i += workers) rebuild(stale[i]);
      });
    for (auto &th : pool) th.join();
  } else {
    for (uint32_t g : stale) rebuild(g);
  }
  for (uint32_t g = 0; g < t->n_gpus; ++g) {
    if (t->bad_count[g].load(std::memory_order_acquire) == 0) {
      iov[g].base = t->healthy_seg[g].data();
      iov[g].len = t->healthy_seg[g].size();
    } else {
      iov[g].base = t->current_seg[g].data();
      iov[g].len = t->current_seg[g].size();
    }
  }
  return GS_OK;
}

}  // extern "C"

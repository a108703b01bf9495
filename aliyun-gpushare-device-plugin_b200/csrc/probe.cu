// probe.cu -- host side of K1, the HBM slice probe (sm_100a): engines, variant dispatch,
// launches and the extern "C" entry points. The kernels are in probe_kernels.cuh. No
// counterpart in the reference: cjg/aliyun-gpushare-device-plugin runs no GPU code
// (SURVEY.md section 8, row K1).

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <mutex>

#include "gpushare.h"
#include "internal.h"
#include "probe_kernels.cuh"

namespace {

using namespace gs_probe_dev;

// ---- variant dispatch ----------------------------------------------------------
using WriteFn = void (*)(const Batch);
using VerifyFn = void (*)(const Batch, DevResult *);

constexpr int kNumWriteVariants = 4;  // STG forms (codes 0..29)
// Bulk-store codes: 100 * threads_sel + 10 * stages_sel + sub_sel
//   threads_sel 1 -> 256 threads per CTA, 2 -> 512, 3 -> 1024 (measured: no difference in a
//               sustained sweep -- the store pass is power-limited there, not fill-limited)
//   stages_sel  0 -> 3-stage ring (192 KiB smem), 1 -> 2-stage ring (128 KiB),
//               2 -> 3-stage ring with an L2 evict-first hint on the bulk stores
//   sub_sel     0 -> 4 KiB bulk ops, 1 -> 2 KiB, 2 -> 8 KiB, 3 -> 16 KiB, 4 -> 1 KiB
constexpr int kWriteBulkBase = 100;
constexpr int kDefaultWriteVariant = 102;  // 256 threads, 3-stage ring, 8 KiB bulk ops
// aggregate offered store rate the token buckets add up to (GB/s == bytes/ns)
constexpr uint32_t kDefaultWriteRateGBps = 7600;
constexpr uint32_t kBulkSubVec[5] = {256, 128, 512, 1024, 64};
// Paced-STG codes: 400 + sub_sel, sub_sel as for the bulk codes (4 KiB, 2, 8, 16, 1) plus
//   5 -> 32 KiB, 6 -> 64 KiB per token
constexpr int kWritePacedBase = 400;
constexpr uint32_t kPacedSubVec[7] = {256, 128, 512, 1024, 64, 2048, 4096};
bool is_paced_code(int code) { return code >= kWritePacedBase && code < kWritePacedBase + 7; }
bool is_bulk_code(int code) {
  return code >= kWriteBulkBase && code < 4 * kWriteBulkBase && (code % 100) / 10 < 3 &&
         code % 10 < 5;
}
constexpr int kNumVerifyVariants = 5;
constexpr int kNumUnrolls = 3;  // U = 2, 4, 8 -> variant index / 10 selects
// Bulk-load verify codes: 500 + 100 * hint + 10 * chunk_sel + warps_sel
//   hint      0 -> plain bulk loads, 1 -> with an L2 evict-first hint (codes 600..)
//   chunk_sel 0 -> 16 KiB stages, 1 -> 32 KiB, 2 -> 64 KiB (the ring holds as many as fit)
//   warps_sel 0 -> 4 consumer warps, 1 -> 8, 2 -> 16 (plus the producer warp)
constexpr int kVerifyBulkBase = 500;
constexpr uint32_t kVerifyBulkRingBytes = 192u << 10;  // per SM; split between the CTAs of an SM
bool is_bulk_verify_code(int code) {
  return code >= kVerifyBulkBase && code < kVerifyBulkBase + 200 && (code % 100) / 10 < 3 &&
         code % 10 < 3;
}
using VerifyBulkFn = void (*)(const Batch, DevResult *, uint32_t);
template <bool HINT>
VerifyBulkFn verify_bulk_fn_h(int chunk_sel, int warps_sel) {
  static const VerifyBulkFn fns[3][3] = {
      {probe_verify_bulk_kernel<1024, 4, HINT>, probe_verify_bulk_kernel<1024, 8, HINT>,
       probe_verify_bulk_kernel<1024, 16, HINT>},
      {probe_verify_bulk_kernel<2048, 4, HINT>, probe_verify_bulk_kernel<2048, 8, HINT>,
       probe_verify_bulk_kernel<2048, 16, HINT>},
      {probe_verify_bulk_kernel<4096, 4, HINT>, probe_verify_bulk_kernel<4096, 8, HINT>,
       probe_verify_bulk_kernel<4096, 16, HINT>}};
  return fns[chunk_sel][warps_sel];
}
VerifyBulkFn verify_bulk_fn(int code) {
  const int c = (code % 100) / 10, w = code % 10;
  return code >= kVerifyBulkBase + 100 ? verify_bulk_fn_h<true>(c, w) : verify_bulk_fn_h<false>(c, w);
}

template <int U>
WriteFn write_fn_u(int v) {
  switch (v) {
    case 1: return probe_write_kernel<U, 1>;
    case 2: return probe_write_kernel<U, 2>;
    case 3: return probe_write_kernel<U, 3>;
    default: return probe_write_kernel<U, 0>;
  }
}
template <int U>
VerifyFn verify_fn_u(int v) {
  switch (v) {
    case 1: return probe_verify_kernel<U, 1>;
    case 2: return probe_verify_kernel<U, 2>;
    case 3: return probe_verify_kernel<U, 3>;
    case 4: return probe_verify_kernel<U, 4>;
    default: return probe_verify_kernel<U, 0>;
  }
}
// variant code = unroll_sel * 10 + access form; unroll_sel 0 -> U=4, 1 -> U=2, 2 -> U=8
WriteFn write_fn(int code) {
  const int v = code % 10, us = code / 10;
  if (us == 1) return write_fn_u<2>(v);
  if (us == 2) return write_fn_u<8>(v);
  return write_fn_u<4>(v);
}
VerifyFn verify_fn(int code) {
  const int v = code % 10, us = code / 10;
  if (us == 1) return verify_fn_u<2>(v);
  if (us == 2) return verify_fn_u<8>(v);
  return verify_fn_u<4>(v);
}
bool valid_code(int code, int n_forms) {
  return code >= 0 && code / 10 < kNumUnrolls && code % 10 < n_forms;
}

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace

struct gs_probe {
  int dev = 0;
  int sms = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  void *scrub = nullptr;  // kScrubBytes, allocated on the first GS_PROBE_SCRUB
  uint32_t scrub_mode = 1;  // 1 = stream loads (default), 0 = store zeros (GS_PROBE_SCRUB_MODE)
  DevResult *d_res = nullptr;
  DevResult *h_res = nullptr;  // pinned
  uint32_t pending_n = 0;
  uint32_t pending_passes = 0;
  uint32_t pending_launches = 0;
  uint64_t pending_bytes = 0;
  bool pending = false;
  int write_variant = kDefaultWriteVariant;
  bool bulk_attr_set[9] = {false, false, false, false, false, false, false, false, false};
  bool verify_bulk_attr_set[18] = {};
  uint32_t write_rate_gbps = kDefaultWriteRateGBps;  // 0 = unpaced
  bool write_rate_auto = false;  // opt-in: back off for a board that cannot keep up (tune_write_rate)
  int verify_variant = 10;  // U = 2, plain ld.global.v4: 30 registers, 8 CTAs/SM resident;
                            // best sustained rate after a store pass (tools/probe_sustained_grid.py)
  int ctas_per_sm = 0;      // STG/LDG kernels: CTAs per SM in the grid; 0 = occupancy query
  // test hook (gs_probe_test_skip_stores): store passes from the skip_stores_from-th on (1-based)
  // launch nothing; 0 = never
  uint32_t skip_stores_from = 0;
  uint32_t store_passes = 0;
  uint32_t seed_nonce = 0;  // drawn when the engine is opened
  uint32_t seed_count = 0;  // sweeps seeded so far
};

// Something that differs between two engines and between two runs of the process: the
// monotonic clock, the pid and the engine's address, avalanched.
static uint32_t fresh_nonce() {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  static std::atomic<uint32_t> opened{0};  // engines may be opened from several threads at once
  uint64_t x = (uint64_t)ts.tv_nsec ^ ((uint64_t)ts.tv_sec << 30) ^ ((uint64_t)getpid() << 17) ^
               (uint64_t)(uintptr_t)&opened ^ ((uint64_t)(opened.fetch_add(1) + 1) << 48);
  x ^= x >> 33;
  x *= 0xFF51AFD7ED558CCDull;
  x ^= x >> 33;
  x *= 0xC4CEB9FE1A85EC53ull;
  x ^= x >> 33;
  return (uint32_t)x;
}

#define GS_CUDA(call)                                      \
  do {                                                     \
    cudaError_t e_ = (call);                               \
    if (e_ != cudaSuccess) {                               \
      gs_internal::set_cuda_error(#call, (int)e_, cudaGetErrorString(e_)); \
      (void)cudaGetLastError(); /* read it off, or the next launch check reports it again */ \
      return GS_E_CUDA;                                    \
    }                                                      \
  } while (0)

extern "C" {

int gs_probe_open(int cuda_ordinal, gs_probe **out) {
  if (!out || cuda_ordinal < 0) return GS_E_INVAL;
  *out = nullptr;
  int count = 0;
  GS_CUDA(cudaGetDeviceCount(&count));
  if (cuda_ordinal >= count) return GS_E_NOTFOUND;
  DeviceGuard g(cuda_ordinal);
  if (!g.ok) {
    gs_internal::set_cuda_error("cudaSetDevice", -1, "cannot select device");
    return GS_E_CUDA;
  }
  gs_probe *p = new (std::nothrow) gs_probe();
  if (!p) {
    gs_internal::set_cuda_error("gs_probe_open", 0, "out of host memory");
    return GS_E_CUDA;
  }
  p->dev = cuda_ordinal;
  p->seed_nonce = fresh_nonce();
  if (const char *m = getenv("GS_PROBE_TEST_SKIP_STORES")) p->skip_stores_from = (uint32_t)atoi(m);
  if (const char *m = getenv("GS_PROBE_SCRUB_MODE")) p->scrub_mode = (uint32_t)atoi(m);
  cudaError_t e = cudaDeviceGetAttribute(&p->sms, cudaDevAttrMultiProcessorCount, cuda_ordinal);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking);
  for (int i = 0; i < 5 && e == cudaSuccess; ++i) e = cudaEventCreate(&p->ev[i]);
  if (e == cudaSuccess) e = cudaMalloc(&p->d_res, sizeof(DevResult) * GS_PROBE_MAX_SLICES);
  if (e == cudaSuccess) e = cudaMallocHost(&p->h_res, sizeof(DevResult) * GS_PROBE_MAX_SLICES);
  if (e != cudaSuccess) {
    gs_internal::set_cuda_error("gs_probe_open", (int)e, cudaGetErrorString(e));
    (void)cudaGetLastError();
    gs_probe_close(p);
    return GS_E_CUDA;
  }
  *out = p;
  return GS_OK;
}

int gs_probe_close(gs_probe *p) {
  if (!p) return GS_E_INVAL;
  DeviceGuard g(p->dev);
  if (p->stream) cudaStreamSynchronize(p->stream);
  if (p->d_res) cudaFree(p->d_res);
  if (p->scrub) cudaFree(p->scrub);
  if (p->h_res) cudaFreeHost(p->h_res);
  for (auto &e : p->ev)
    if (e) cudaEventDestroy(e);
  if (p->stream) cudaStreamDestroy(p->stream);
  delete p;
  return GS_OK;
}

int gs_probe_sm_count(const gs_probe *p, int *sms) {
  if (!p || !sms) return GS_E_INVAL;
  *sms = p->sms;
  return GS_OK;
}

int gs_probe_test_skip_stores(gs_probe *p, int on) {
  if (!p) return GS_E_INVAL;
  p->skip_stores_from = on ? p->store_passes + 1 : 0;
  return GS_OK;
}

int gs_probe_next_seed(gs_probe *p, uint32_t *seed) {
  if (!p || !seed) return GS_E_INVAL;
  // nonce + odd multiple of the counter: 2^32 distinct values before the first repeat
  *seed = p->seed_nonce + 0x9E3779B1u * ++p->seed_count;
  return GS_OK;
}

int gs_probe_mem_alloc(gs_probe *p, uint64_t bytes, void **dptr) {
  if (!p || !dptr || bytes == 0) return GS_E_INVAL;
  DeviceGuard g(p->dev);
  GS_CUDA(cudaMalloc(dptr, bytes));
  return GS_OK;
}

int gs_probe_mem_free(gs_probe *p, void *dptr) {
  if (!p) return GS_E_INVAL;
  DeviceGuard g(p->dev);
  GS_CUDA(cudaFree(dptr));
  return GS_OK;
}

int gs_probe_mem_info(gs_probe *p, uint64_t *free_bytes, uint64_t *total_bytes) {
  if (!p) return GS_E_INVAL;
  DeviceGuard g(p->dev);
  size_t f = 0, t = 0;
  GS_CUDA(cudaMemGetInfo(&f, &t));
  if (free_bytes) *free_bytes = f;
  if (total_bytes) *total_bytes = t;
  return GS_OK;
}

int gs_probe_set_variant(gs_probe *p, int write_variant, int verify_variant,
                         int ctas_per_sm) {
  if (!p) return GS_E_INVAL;
  if (write_variant >= 0) {
    if (!valid_code(write_variant, kNumWriteVariants) && !is_bulk_code(write_variant) &&
        !is_paced_code(write_variant))
      return GS_E_INVAL;
    p->write_variant = write_variant;
  }
  if (verify_variant >= 0) {
    if (!valid_code(verify_variant, kNumVerifyVariants) && !is_bulk_verify_code(verify_variant))
      return GS_E_INVAL;
    p->verify_variant = verify_variant;
  }
  if (ctas_per_sm >= 0) {
    if (ctas_per_sm > 32) return GS_E_INVAL;
    p->ctas_per_sm = ctas_per_sm;
  }
  return GS_OK;
}

int gs_probe_set_write_rate(gs_probe *p, uint32_t gbps) {
  if (!p) return GS_E_INVAL;
  if (gbps > 100000) return GS_E_INVAL;
  p->write_rate_gbps = gbps;
  p->write_rate_auto = false;  // an explicit target stays put
  return GS_OK;
}

int gs_probe_set_write_rate_auto(gs_probe *p, int on) {
  if (!p) return GS_E_INVAL;
  p->write_rate_auto = on != 0;
  if (on && p->write_rate_gbps == 0) p->write_rate_gbps = kDefaultWriteRateGBps;
  return GS_OK;
}

int gs_probe_get_write_rate(const gs_probe *p, uint32_t *gbps, int *is_auto) {
  if (!p || !gbps) return GS_E_INVAL;
  *gbps = p->write_rate_gbps;
  if (is_auto) *is_auto = p->write_rate_auto ? 1 : 0;
  return GS_OK;
}

// Opt-in (gs_probe_set_write_rate_auto). The write path's ceiling differs by a few percent from
// board to board (7.27-7.60 TB/s across the eight GPUs of one box at the fixed 7 600 GB/s
// target), and offering more than the ceiling costs more than offering less
// (profiles/r01_rate_sweep_sustained.txt: the collapse). With this on, a store pass of at least
// 4 GiB that fell more than 1 % behind its target steps the target down 0.8 %, one that kept up
// lets it creep back 0.05 %, never above 7 600. It is OFF by default because on the healthy
// boards it was tried on it loses to the fixed target (tools/rate_follow.py: 7.39-7.44 TB/s
// against 7.52 fixed): single slow passes, which a fixed target shrugs off, each cost a step
// that takes sixteen good passes to win back. It is kept for a board that sits in the collapse.
static void tune_write_rate(gs_probe *p, uint64_t bytes, float write_ms) {
  if (!p->write_rate_auto || !p->write_rate_gbps || bytes < (4ull << 30) || write_ms <= 0.f) return;
  if (!is_paced_code(p->write_variant) && !is_bulk_code(p->write_variant)) return;  // unpaced forms
  const double achieved = (double)bytes / (double)write_ms / 1e6;
  const double target = (double)p->write_rate_gbps;
  double next = achieved >= 0.990 * target ? target * 1.0005 : target * 0.992;
  if (next < 6400.0) next = 6400.0;
  if (next > (double)kDefaultWriteRateGBps) next = (double)kDefaultWriteRateGBps;
  p->write_rate_gbps = (uint32_t)(next + 0.5);
}

static int grid_for(const gs_probe *p, const void *fn, uint64_t total_tiles, int *grid) {
  int per_sm = p->ctas_per_sm;
  if (per_sm <= 0) {
    GS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kThreads, 0));
    if (per_sm < 1) per_sm = 1;
  }
  uint64_t g = (uint64_t)p->sms * (uint64_t)per_sm;
  if (g > total_tiles) g = total_tiles;
  *grid = (int)g;
  return GS_OK;
}

// Slice list -> tile space. Returns GS_E_INVAL for a slice that is not a 16-byte aligned,
// non-empty multiple of 16 bytes or names an unknown pattern.
static int build_batch(const gs_slice *slices, uint32_t n, Batch *b, uint64_t *bytes_out) {
  memset(b, 0, sizeof(*b));
  uint64_t tiles = 0, bytes = 0;
  for (uint32_t i = 0; i < n; ++i) {
    const gs_slice &s = slices[i];
    if (!s.base || s.bytes == 0 || (s.bytes & 15u) || ((uintptr_t)s.base & 15u)) return GS_E_INVAL;
    if (s.pattern != GS_PROBE_PATTERN_LOWTOGGLE && s.pattern != GS_PROBE_PATTERN_HASH)
      return GS_E_INVAL;
    SliceDesc &d = b->s[i];
    d.base = (uint4 *)s.base;
    d.n_vec = s.bytes / 16;
    d.tile_begin = tiles;
    tiles += (d.n_vec + kTileVec - 1) / kTileVec;
    d.tile_end = tiles;
    d.seed = s.seed;
    d.pattern = s.pattern;
    bytes += s.bytes;
  }
  b->total_tiles = tiles;
  b->n = n;
  if (bytes_out) *bytes_out = bytes;
  return GS_OK;
}

int gs_probe_submit(gs_probe *p, const gs_slice *slices, uint32_t n, uint32_t passes) {
  if (!p || !slices || n == 0 || (passes & ~(GS_PROBE_WRITE | GS_PROBE_VERIFY | GS_PROBE_SCRUB)) ||
      (passes & (GS_PROBE_WRITE | GS_PROBE_VERIFY)) == 0)
    return GS_E_INVAL;
  if (n > GS_PROBE_MAX_SLICES) return GS_E_LIMIT;
  if (p->pending) return GS_E_STATE;
  Batch b;
  uint64_t bytes = 0;
  if (int rc = build_batch(slices, n, &b, &bytes)) return rc;
  const uint64_t tiles = b.total_tiles;

  DeviceGuard g(p->dev);
  if (!g.ok) {
    gs_internal::set_cuda_error("cudaSetDevice", -1, "cannot select device");
    return GS_E_CUDA;
  }
  (void)cudaGetLastError();  // an error some earlier call on this thread left unread is not ours
  uint32_t launches = 0;
  GS_CUDA(cudaMemsetAsync(p->d_res, 0, sizeof(DevResult) * n, p->stream));
  GS_CUDA(cudaEventRecord(p->ev[0], p->stream));
  if (passes & GS_PROBE_WRITE) ++p->store_passes;
  if ((passes & GS_PROBE_WRITE) && p->skip_stores_from && p->store_passes >= p->skip_stores_from) {
    // test hook: a store pass that silently does nothing; the load pass must then report the
    // range (a fresh seed never matches what an earlier sweep left behind)
  } else if ((passes & GS_PROBE_WRITE) && is_paced_code(p->write_variant)) {
    const uint32_t sub_vec = kPacedSubVec[p->write_variant - kWritePacedBase];
    const uint64_t grid = tiles < (uint64_t)p->sms ? tiles : (uint64_t)p->sms;
    uint32_t period_x16 = 0;  // per token of sub_vec words, so that `grid` CTAs offer write_rate_gbps
    if (p->write_rate_gbps)
      period_x16 = (uint32_t)((double)sub_vec * 16.0 * (double)grid / (double)p->write_rate_gbps * 16.0);
    probe_write_paced_kernel<<<(int)grid, kThreads, 0, p->stream>>>(b, sub_vec, period_x16);
    GS_CUDA(cudaGetLastError());
    ++launches;
  } else if ((passes & GS_PROBE_WRITE) && is_bulk_code(p->write_variant)) {
    const int threads_sel = p->write_variant / 100;  // 1, 2, 3
    const int stages_sel = (p->write_variant % 100) / 10;
    const uint32_t sub_vec = kBulkSubVec[p->write_variant % 10];
    const int stages = stages_sel == 1 ? 2 : 3;
    const int threads = 128 << threads_sel;
    const size_t smem = (size_t)stages * kTileVec * sizeof(uint4);
    using BulkFn = void (*)(const Batch, uint32_t, uint32_t);
    static const BulkFn kBulkFns[3][3] = {
        {probe_write_bulk_kernel<3, 256>, probe_write_bulk_kernel<2, 256>,
         probe_write_bulk_kernel<3, 256, true>},
        {probe_write_bulk_kernel<3, 512>, probe_write_bulk_kernel<2, 512>,
         probe_write_bulk_kernel<3, 512, true>},
        {probe_write_bulk_kernel<3, 1024>, probe_write_bulk_kernel<2, 1024>,
         probe_write_bulk_kernel<3, 1024, true>}};
    BulkFn fn = kBulkFns[threads_sel - 1][stages_sel];
    const int slot = (threads_sel - 1) * 3 + stages_sel;
    if (!p->bulk_attr_set[slot]) {
      GS_CUDA(cudaFuncSetAttribute((const void *)fn,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      p->bulk_attr_set[slot] = true;
    }
    // the ring fills an SM's shared memory: exactly one CTA per SM
    const uint64_t grid = tiles < (uint64_t)p->sms ? tiles : (uint64_t)p->sms;
    // per-CTA period so that `grid` CTAs together offer write_rate_gbps
    uint32_t period_x16 = 0;
    if (p->write_rate_gbps)
      period_x16 = (uint32_t)((double)kTileVec * 16.0 * (double)grid / (double)p->write_rate_gbps * 16.0);
    fn<<<(int)grid, threads, smem, p->stream>>>(b, sub_vec, period_x16);
    GS_CUDA(cudaGetLastError());
    ++launches;
  } else if (passes & GS_PROBE_WRITE) {
    WriteFn fn = write_fn(p->write_variant);
    int grid = 0;
    int rc = grid_for(p, (const void *)fn, tiles, &grid);
    if (rc != GS_OK) return rc;
    fn<<<grid, kThreads, 0, p->stream>>>(b);
    GS_CUDA(cudaGetLastError());
    ++launches;
  }
  GS_CUDA(cudaEventRecord(p->ev[1], p->stream));
  if (passes & GS_PROBE_SCRUB) {
    if (!p->scrub) {
      GS_CUDA(cudaMalloc(&p->scrub, kScrubBytes + 256));
      GS_CUDA(cudaMemsetAsync(p->scrub, 0, kScrubBytes + 256, p->stream));
    }
    probe_scrub_kernel<<<p->sms * 8, kThreads, 0, p->stream>>>((uint4 *)p->scrub, kScrubBytes / 16,
                                                               p->scrub_mode);
    GS_CUDA(cudaGetLastError());
    ++launches;
  }
  GS_CUDA(cudaEventRecord(p->ev[4], p->stream));
  if ((passes & GS_PROBE_VERIFY) && is_bulk_verify_code(p->verify_variant)) {
    const int code = p->verify_variant;
    const int chunk_sel = (code % 100) / 10, warps_sel = code % 10;
    const uint32_t chunk_bytes = 16384u << chunk_sel;
    const int per_sm = p->ctas_per_sm == 2 ? 2 : 1;  // the ring is an SM's shared memory, split
    uint32_t stages = kVerifyBulkRingBytes / (uint32_t)per_sm / chunk_bytes;
    if (stages > kBulkMaxStages) stages = kBulkMaxStages;
    const size_t smem = (size_t)stages * chunk_bytes;
    VerifyBulkFn fn = verify_bulk_fn(code);
    // the attribute is a maximum: set it once per kernel to the whole ring
    const int slot = (code >= kVerifyBulkBase + 100 ? 9 : 0) + chunk_sel * 3 + warps_sel;
    if (!p->verify_bulk_attr_set[slot]) {
      GS_CUDA(cudaFuncSetAttribute((const void *)fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)kVerifyBulkRingBytes));
      p->verify_bulk_attr_set[slot] = true;
    }
    uint64_t grid = (uint64_t)p->sms * (uint64_t)per_sm;
    if (grid > tiles) grid = tiles;
    fn<<<(int)grid, 32 + 32 * (4 << warps_sel), smem, p->stream>>>(b, p->d_res, stages);
    GS_CUDA(cudaGetLastError());
    ++launches;
  } else if (passes & GS_PROBE_VERIFY) {
    VerifyFn fn = verify_fn(p->verify_variant);
    int grid = 0;
    int rc = grid_for(p, (const void *)fn, tiles, &grid);
    if (rc != GS_OK) return rc;
    fn<<<grid, kThreads, 0, p->stream>>>(b, p->d_res);
    GS_CUDA(cudaGetLastError());
    ++launches;
  }
  GS_CUDA(cudaEventRecord(p->ev[2], p->stream));
  GS_CUDA(cudaMemcpyAsync(p->h_res, p->d_res, sizeof(DevResult) * n, cudaMemcpyDeviceToHost,
                          p->stream));
  GS_CUDA(cudaEventRecord(p->ev[3], p->stream));
  p->pending = true;
  p->pending_n = n;
  p->pending_passes = passes;
  p->pending_bytes = bytes;
  p->pending_launches = launches;
  return GS_OK;
}

int gs_probe_poll(gs_probe *p) {
  if (!p) return GS_E_INVAL;
  if (!p->pending) return GS_OK;
  DeviceGuard g(p->dev);
  cudaError_t e = cudaEventQuery(p->ev[3]);
  if (e == cudaSuccess) return GS_OK;
  if (e == cudaErrorNotReady) return GS_E_TIMEOUT;
  gs_internal::set_cuda_error("cudaEventQuery", (int)e, cudaGetErrorString(e));
  return GS_E_CUDA;
}

int gs_probe_wait(gs_probe *p, gs_slice_result *results, uint32_t n, gs_probe_timing *timing) {
  if (!p) return GS_E_INVAL;
  if (!p->pending) return GS_E_STATE;
  if (results && n < p->pending_n) return GS_E_NOSPACE;
  DeviceGuard g(p->dev);
  p->pending = false;
  GS_CUDA(cudaEventSynchronize(p->ev[3]));
  if (results) {
    for (uint32_t i = 0; i < p->pending_n; ++i) {
      const DevResult &r = p->h_res[i];
      results[i].mismatches = r.mismatches;
      results[i].first_bad_byte = r.mismatches ? ~r.inv_first_bad : UINT64_MAX;
      results[i].bit_flips = r.bit_flips;
      results[i].reserved = 0;
    }
  }
  if (p->pending_passes & GS_PROBE_WRITE) {
    float w = 0.f;
    GS_CUDA(cudaEventElapsedTime(&w, p->ev[0], p->ev[1]));
    tune_write_rate(p, p->pending_bytes, w);
  }
  if (timing) {
    memset(timing, 0, sizeof(*timing));
    if (p->pending_passes & GS_PROBE_WRITE)
      GS_CUDA(cudaEventElapsedTime(&timing->write_ms, p->ev[0], p->ev[1]));
    if (p->pending_passes & GS_PROBE_VERIFY)
      GS_CUDA(cudaEventElapsedTime(&timing->verify_ms, p->ev[4], p->ev[2]));
    if (p->pending_passes & GS_PROBE_SCRUB)
      GS_CUDA(cudaEventElapsedTime(&timing->scrub_ms, p->ev[1], p->ev[4]));
    timing->bytes = p->pending_bytes;
    timing->launches = p->pending_launches;
  }
  return GS_OK;
}

int gs_probe_run(gs_probe *p, const gs_slice *slices, uint32_t n, uint32_t passes,
                 gs_slice_result *results, gs_probe_timing *timing) {
  int rc = gs_probe_submit(p, slices, n, passes);
  if (rc != GS_OK) return rc;
  return gs_probe_wait(p, results, n, timing);
}

int gs_probe_run_granted(int cuda_ordinal, uint64_t bytes, uint32_t seed,
                         gs_slice_result *result, gs_probe_timing *timing) {
  if (bytes == 0 || (bytes & 15u)) return GS_E_INVAL;
  gs_probe *p = nullptr;
  int rc = gs_probe_open(cuda_ordinal, &p);
  if (rc != GS_OK) return rc;
  void *d = nullptr;
  rc = gs_probe_mem_alloc(p, bytes, &d);
  if (rc == GS_OK) {
    gs_slice s{d, bytes, seed, 0};
    rc = gs_probe_run(p, &s, 1, GS_PROBE_WRITE | GS_PROBE_VERIFY, result, timing);
    gs_probe_mem_free(p, d);
  }
  gs_probe_close(p);
  return rc;
}

int gs_probe_inject(gs_probe *p, void *base, uint64_t byte_offset, uint32_t mask) {
  if (!p || !base || (byte_offset & 3u)) return GS_E_INVAL;
  if (p->pending) return GS_E_STATE;
  DeviceGuard g(p->dev);
  probe_inject_kernel<<<1, 1, 0, p->stream>>>((uint32_t *)((char *)base + byte_offset), mask);
  GS_CUDA(cudaGetLastError());
  GS_CUDA(cudaStreamSynchronize(p->stream));
  return GS_OK;
}

int gs_probe_read(gs_probe *p, const void *dptr, void *host, uint64_t bytes) {
  if (!p || !dptr || !host) return GS_E_INVAL;
  DeviceGuard g(p->dev);
  GS_CUDA(cudaStreamSynchronize(p->stream));
  GS_CUDA(cudaMemcpy(host, dptr, bytes, cudaMemcpyDeviceToHost));
  return GS_OK;
}

}  // extern "C"

// ---- read-back pass for callers that may not allocate -----------------------------
// Who cannot allocate: the plugin's health checker on a GPU whose HBM has been handed
// out to tenants down to the last unit (a cudaMalloc there either fails or takes memory
// a tenant was promised), and any caller on a path where cudaMalloc/cudaFree's implicit
// device-wide synchronisation would stall the tenants' own kernels. For them the
// per-slice verdict table -- the only global memory the load+compare kernel needs besides
// the slices themselves -- lives in the fixed array below instead of a cudaMalloc'ed one.
// Every GPU that loads this library carries the array (2 KiB) until the process exits.
namespace {

constexpr uint32_t kNoAllocMaxSlices = GS_PROBE_NOALLOC_MAX_SLICES;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) DevResult g_noalloc_verdicts[kNoAllocMaxSlices];

// In host code `g_noalloc_verdicts` is only a handle for the runtime, not a device
// address; the address a kernel can use is looked up once per device and kept here.
DevResult *g_noalloc_ptr[kNoAllocMaxDevices];
int g_noalloc_sms[kNoAllocMaxDevices];
std::mutex g_noalloc_mu;

}  // namespace

extern "C" int gs_probe_verify_noalloc(int cuda_ordinal, void *stream, const gs_slice *slices,
                                       uint32_t n, gs_slice_result *results) {
  if (cuda_ordinal < 0 || cuda_ordinal >= kNoAllocMaxDevices || !slices || !results || n == 0)
    return GS_E_INVAL;
  if (n > kNoAllocMaxSlices) {
    gs_internal::set_cuda_error("gs_probe_verify_noalloc", 0,
                                "more slices than the fixed verdict table holds "
                                "(GS_PROBE_NOALLOC_MAX_SLICES)");
    return GS_E_LIMIT;
  }
  Batch b;
  if (int rc = build_batch(slices, n, &b, nullptr)) return rc;
  const uint64_t tiles = b.total_tiles;

  DeviceGuard g(cuda_ordinal);
  if (!g.ok) {
    gs_internal::set_cuda_error("cudaSetDevice", -1, "cannot select device");
    return GS_E_CUDA;
  }
  DevResult *table = nullptr;
  int sms = 0;
  {
    // set-up, first call on this device only, with the device current
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[cuda_ordinal]) {
      DevResult *p = nullptr;
      GS_CUDA(cudaGetSymbolAddress((void **)&p, g_noalloc_verdicts));
      int count = 0;
      GS_CUDA(cudaDeviceGetAttribute(&count, cudaDevAttrMultiProcessorCount, cuda_ordinal));
      g_noalloc_sms[cuda_ordinal] = count;
      g_noalloc_ptr[cuda_ordinal] = p;
    }
    table = g_noalloc_ptr[cuda_ordinal];
    sms = g_noalloc_sms[cuda_ordinal];
  }
  cudaStream_t st = (cudaStream_t)stream;
  (void)cudaGetLastError();  // an error some earlier call on this thread left unread is not ours
  // the table still holds the previous call's verdicts: the rows this call uses start at zero
  GS_CUDA(cudaMemsetAsync(table, 0, sizeof(DevResult) * n, st));
  uint64_t grid = (uint64_t)sms * 8u;
  if (grid > tiles) grid = tiles;
  probe_verify_kernel<4, 0><<<(int)grid, kThreads, 0, st>>>(b, table);
  GS_CUDA(cudaGetLastError());
  DevResult host[kNoAllocMaxSlices];
  GS_CUDA(cudaMemcpyAsync(host, table, sizeof(DevResult) * n, cudaMemcpyDeviceToHost, st));
  GS_CUDA(cudaStreamSynchronize(st));
  for (uint32_t i = 0; i < n; ++i) {
    results[i].mismatches = host[i].mismatches;
    results[i].first_bad_byte = host[i].mismatches ? ~host[i].inv_first_bad : UINT64_MAX;
    results[i].bit_flips = host[i].bit_flips;
    results[i].reserved = 0;
  }
  return GS_OK;
}

// gpushare-device-plugin-v2 -- the node daemon as one native binary over libgpushare_core.so.
//
// Replaces, with the same flags, exit codes and restart behaviour:
//   cmd/nvidia/main.go:15-78            flags, memory unit, kubelet client, manager
//   pkg/gpu/nvidia/gpumanager.go:33-111 NVML, FS/OS watchers, the restart loop
//   pkg/gpu/nvidia/watchers.go, coredump.go
//   pkg/gpu/nvidia/server.go:43-78      NewNvidiaDevicePlugin (gpu-count patch, cGPU label)
//   pkg/gpu/nvidia/allocate.go:60-152   list -> decide -> patch under one lock
// The decision, the device table, the wire format, the gRPC transport and the probe are the
// library's (include/gpushare.h); this file is flags, I/O and lifecycle. The same program
// exists in Python (daemon.py); tests run both against the same fake kubelet and cluster.
#include <execinfo.h>
#include <limits.h>
#include <poll.h>
#include <signal.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/inotify.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "gpushare.h"
#include "k8s.h"

using namespace gsd;

namespace {

// ---- flags: Go's flag package (main.go:15-26) ------------------------------------------------
struct Flags {
  bool mps = false, health_check = false, query_kubelet = false, logtostderr = false;
  std::string memory_unit = "GiB", kubelet_address = "0.0.0.0", client_cert, client_key, token;
  long kubelet_port = 10250, timeout = 10, v = 0;
  // additions (not in the reference)
  std::string pod_source_file, nvml_lib, socket_dir = "/var/lib/kubelet/device-plugins", transport = "native";
  bool probe = false, no_register = false;
  long health_probe_interval = 0, probe_gather_ms = 2;
};

int g_v = 0;
void logv(int level, const char *fmt, ...) {
  if (g_v < level) return;
  char line[1024];
  va_list ap;
  va_start(ap, fmt);
  int n = vsnprintf(line, sizeof(line) - 1, fmt, ap);
  va_end(ap);
  if (n < 0) return;
  if (n > (int)sizeof(line) - 2) n = (int)sizeof(line) - 2;
  line[n] = '\n';
  fwrite(line, 1, (size_t)n + 1, stderr);  // one write per line: the probe threads log too
  fflush(stderr);
}

// flag.PrintDefaults of the ten reference flags (main.go:15-26) and the additions
void print_usage() {
  fputs("Usage of gpushare-device-plugin-v2:\n"
        "  -client-cert string\n    \tKubelet TLS client certificate\n"
        "  -client-key string\n    \tKubelet TLS client key\n"
        "  -health-check\n    \tEnable or disable Health check\n"
        "  -kubelet-address string\n    \tKubelet IP Address (default \"0.0.0.0\")\n"
        "  -kubelet-port uint\n    \tKubelet listened Port (default 10250)\n"
        "  -memory-unit string\n    \tSet memoryUnit of the GPU Memroy, support 'GiB' and 'MiB' (default \"GiB\")\n"
        "  -mps\n    \tEnable or Disable MPS\n"
        "  -query-kubelet\n    \tQuery pending pods from kubelet instead of kube-apiserver\n"
        "  -timeout int\n    \tKubelet client http timeout duration (default 10)\n"
        "  -token string\n    \tKubelet client bearer token\n"
        "  -logtostderr, -v int\n    \tglog flags the DaemonSet passes; -v sets the verbosity\n"
        "additions:\n"
        "  -health-probe-interval int\n    \twith -probe and -health-check: every N seconds sweep each GPU's free memory (0 = off)\n"
        "  -no-register\n    \tserve without registering with kubelet\n"
        "  -nvml-lib string\n    \tpath of libnvidia-ml\n"
        "  -pod-source-file string\n    \tJSON file with the node's pods (stands in for the lister)\n"
        "  -probe\n    \tsweep each granted slice on the GPU after Allocate\n"
        "  -probe-gather-ms int\n    \twith -probe: wait this long for more grants before a GPU's batch is swept (default 2)\n"
        "  -socket-dir string\n    \tdevice-plugins directory (default \"/var/lib/kubelet/device-plugins\")\n",
        stderr);
}

[[noreturn]] void usage_error(const std::string &msg) {
  fprintf(stderr, "%s\n", msg.c_str());
  print_usage();
  exit(2);  // Go's flag.ExitOnError
}

bool parse_bool(const std::string &s, bool &out) {  // strconv.ParseBool
  if (s == "1" || s == "t" || s == "T" || s == "true" || s == "TRUE" || s == "True") return out = true, true;
  if (s == "0" || s == "f" || s == "F" || s == "false" || s == "FALSE" || s == "False") return out = false, true;
  return false;
}

Flags parse_flags(int argc, char **argv) {
  Flags f;
  std::map<std::string, bool *> bools = {{"mps", &f.mps}, {"health-check", &f.health_check},
                                         {"query-kubelet", &f.query_kubelet}, {"logtostderr", &f.logtostderr},
                                         {"probe", &f.probe}, {"no-register", &f.no_register}};
  std::map<std::string, std::string *> strs = {{"memory-unit", &f.memory_unit}, {"kubelet-address", &f.kubelet_address},
                                               {"client-cert", &f.client_cert}, {"client-key", &f.client_key},
                                               {"token", &f.token}, {"pod-source-file", &f.pod_source_file},
                                               {"nvml-lib", &f.nvml_lib}, {"socket-dir", &f.socket_dir},
                                               {"transport", &f.transport}};
  std::map<std::string, long *> ints = {{"kubelet-port", &f.kubelet_port}, {"timeout", &f.timeout}, {"v", &f.v},
                                        {"health-probe-interval", &f.health_probe_interval},
                                        {"probe-gather-ms", &f.probe_gather_ms}};
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    if (a.size() < 2 || a[0] != '-') break;  // first non-flag argument ends parsing
    if (a == "--") break;
    a = a.substr(a[1] == '-' ? 2 : 1);
    if (a == "h" || a == "help") {  // flag.ErrHelp: usage, exit status 0
      print_usage();
      exit(0);
    }
    std::string val;
    bool has_val = false;
    const size_t eq = a.find('=');
    if (eq != std::string::npos) {
      val = a.substr(eq + 1);
      a.resize(eq);
      has_val = true;
    }
    if (bools.count(a)) {
      bool b = true;
      if (has_val && !parse_bool(val, b)) usage_error("invalid boolean value \"" + val + "\" for -" + a);
      *bools[a] = b;
      continue;
    }
    if (!strs.count(a) && !ints.count(a)) usage_error("flag provided but not defined: -" + a);
    if (!has_val) {
      if (i + 1 >= argc) usage_error("flag needs an argument: -" + a);
      val = argv[++i];
    }
    if (strs.count(a)) {
      *strs[a] = val;
    } else {
      char *end = nullptr;
      const long n = strtol(val.c_str(), &end, 10);
      if (val.empty() || *end) usage_error("invalid value \"" + val + "\" for flag -" + a);
      *ints[a] = n;
    }
  }
  if (f.transport != "native") usage_error("this binary serves over the library's own transport only (-transport=native)");
  return f;
}

// ---- the slice probe beside Allocate (no counterpart in the reference) --------------------------
// One worker thread PER GPU owns every CUDA call for that GPU and never waits for a sweep: a grant
// is queued when its answer has left, the worker gathers what has queued up into one batch (at
// most GS_PROBE_MAX_SLICES slices per launch pair, an L2 scrub between the passes when the batch is
// smaller than the L2 can hide), submits it and goes on; finished batches are collected by polling.
// Health-time sweeps of free memory fire on every GPU at the same instant (the ticks are counted
// from one shared epoch), each on its own thread -- allocation, sweep and release included -- so a
// node's sweep takes far less than the sum (8 B200s: 0.73 s against 8 x 0.36 s; the driver
// serialises part of each allocation, the sweeps overlap). Every sweep draws its seeds from the engine
// (gs_probe_next_seed): what an earlier sweep left in a range never matches.
struct ProbeJob {
  int gpu;
  uint64_t bytes;
  int32_t pod_index;
};

struct ProbeWorker {
  static constexpr uint64_t kScrubBelow = 512ull << 20;     // batch bytes below which the L2 is scrubbed
  static constexpr uint64_t kGrantMargin = 1ull << 30;      // free HBM left alone when sweeping a grant
  static constexpr uint64_t kHealthMargin = 2ull << 30;     // ... when sweeping free memory
  static constexpr uint64_t kSentinelBytes = 64ull << 20;   // kept per GPU for the no-allocation read-back
  static constexpr int kGrowWindowSeconds = 600;            // how long a fresh grant is held free for its tenant
  using Clock = std::chrono::steady_clock;

  struct Gpu {
    int index = -1;  // NVML index
    gs_probe *probe = nullptr;
    int cuda_ordinal = -1;
    uint64_t total_bytes = 0;
    // grants of the last kGrowWindow: a tenant that was just admitted may not have allocated its
    // memory yet, and a health sweep must leave that much alone
    std::deque<std::pair<Clock::time_point, uint64_t>> recent_grants;
    std::deque<ProbeJob> pending;        // grants waiting for the engine
    Clock::time_point first_pending;     // when the oldest of them arrived
    std::vector<void *> arenas;          // ranges of the batch in flight (returned when it ends)
    std::vector<int32_t> tags;
    bool inflight = false;
    bool inflight_is_health = false;
    Clock::time_point health_tick;       // the tick the health sweep in flight belongs to
    void *sentinel = nullptr;            // written once at start, read back without allocating
    uint32_t sentinel_seed = 0;
    uint64_t dropped = 0;
    std::mutex mu;                       // guards queue
    std::condition_variable cv;
    std::deque<ProbeJob> queue;          // from the Allocate path
    std::thread th;
  };
  std::map<int, std::unique_ptr<Gpu>> gpus;  // NVML index -> state; the map itself is fixed once the threads run
  std::atomic<bool> stop{false};
  gs_table *table = nullptr;
  gs_plugin *plugin = nullptr;
  long health_interval = 0;
  long gather_ms = 2;  // after the first grant of a batch arrives, wait this long for company
  Clock::time_point epoch;

  void add_gpu(int nvml_index, int cuda_ordinal, uint64_t total_bytes, gs_probe *p) {
    std::unique_ptr<Gpu> g(new Gpu());
    g->index = nvml_index;
    g->probe = p;
    g->cuda_ordinal = cuda_ordinal;
    g->total_bytes = total_bytes;
    gpus[nvml_index] = std::move(g);
  }
  bool has(int nvml_index) const { return gpus.count(nvml_index) != 0; }
  void mark_bad(int gpu) {
    gs_law_mark_unhealthy_gpu(table, gpu);
    gs_plugin_notify_health(plugin);
  }
  void release(Gpu &g) {
    // the ranges belong to the tenants: give them back the moment the sweep is over
    for (void *d : g.arenas) gs_probe_mem_free(g.probe, d);
    g.arenas.clear();
    g.tags.clear();
    g.inflight = false;
  }
  void harvest(Gpu &g, bool block) {
    if (!g.inflight) return;
    if (!block && gs_probe_poll(g.probe) != GS_OK) return;
    gs_slice_result r[GS_PROBE_MAX_SLICES];
    gs_probe_timing tm;
    const char *label = g.inflight_is_health ? "health probe" : "probe";
    if (gs_probe_wait(g.probe, r, GS_PROBE_MAX_SLICES, &tm) == GS_OK) {
      const double ms = tm.write_ms + tm.verify_ms + tm.scrub_ms;
      unsigned long long bad = 0;
      for (size_t i = 0; i < g.arenas.size(); ++i) bad += r[i].mismatches;
      logv(1, "%s: gpu %d swept %llu bytes in %zu slice(s), %u launches, at %.0f GB/s, %llu mismatching words", label,
           g.index, (unsigned long long)tm.bytes, g.arenas.size(), tm.launches,
           ms > 0 ? 2.0 * (double)tm.bytes / ms / 1e6 : 0.0, bad);
      if (bad) mark_bad(g.index);
    } else {
      logv(1, "%s: gpu %d: %s", label, g.index, gs_cuda_last_error());
    }
    const bool was_health = g.inflight_is_health;
    release(g);
    if (was_health)
      logv(1, "health probe: gpu %d done %.0f ms after the tick (allocation, sweep and release)", g.index,
           std::chrono::duration<double, std::milli>(Clock::now() - g.health_tick).count());
  }
  // the grants queued for one GPU -> one batch -> one launch pair
  void start_grants(Gpu &g) {
    if (g.inflight || g.pending.empty()) return;
    // a pod's containers and the pods of one scheduling burst arrive within milliseconds of each
    // other: a short gather window turns them into one launch pair (nobody waits for the sweep)
    if (g.pending.size() < GS_PROBE_MAX_SLICES &&
        Clock::now() - g.first_pending < std::chrono::milliseconds(gather_ms))
      return;
    gs_slice sl[GS_PROBE_MAX_SLICES];
    memset(sl, 0, sizeof(sl));
    uint32_t n = 0;
    uint64_t total = 0;
    while (!g.pending.empty() && n < GS_PROBE_MAX_SLICES) {
      const ProbeJob j = g.pending.front();
      g.pending.pop_front();
      uint64_t free_b = 0, total_b = 0;
      if (gs_probe_mem_info(g.probe, &free_b, &total_b) != GS_OK) {
        logv(1, "probe skipped: %s", gs_cuda_last_error());
        continue;
      }
      uint64_t bytes = j.bytes;  // exactly the granted range, capped by what is free right now
      const uint64_t room = free_b > kGrantMargin ? free_b - kGrantMargin : 0;
      if (bytes > room) bytes = room;
      bytes &= ~0xFFFFFull;
      if (!bytes) {
        logv(1, "probe skipped: gpu %d has no free memory left for the %llu bytes granted to pod %d", g.index,
             (unsigned long long)j.bytes, j.pod_index);
        continue;
      }
      void *d = nullptr;
      if (gs_probe_mem_alloc(g.probe, bytes, &d) != GS_OK) {
        logv(1, "probe skipped: %s", gs_cuda_last_error());
        continue;
      }
      g.arenas.push_back(d);
      g.tags.push_back(j.pod_index);
      sl[n].base = d;
      sl[n].bytes = bytes;
      gs_probe_next_seed(g.probe, &sl[n].seed);
      sl[n].tag = (uint32_t)j.pod_index;
      sl[n].pattern = GS_PROBE_PATTERN_LOWTOGGLE;
      total += bytes;
      logv(1, "probe: gpu %d, %llu bytes granted to pod %d queued for a sweep", g.index, (unsigned long long)bytes,
           j.pod_index);
      ++n;
    }
    if (!n) return;
    uint32_t passes = GS_PROBE_WRITE | GS_PROBE_VERIFY;
    if (total < kScrubBelow) passes |= GS_PROBE_SCRUB;
    if (gs_probe_submit(g.probe, sl, n, passes) != GS_OK) {
      logv(1, "probe skipped: %s", gs_cuda_last_error());
      release(g);  // nothing was launched: the ranges go back at once
      return;
    }
    g.inflight = true;
    g.inflight_is_health = false;
    logv(1, "probe: gpu %d, %u slice(s), %llu bytes in one launch pair", g.index, n, (unsigned long long)total);
  }
  // Health-time probe. What may be written is memory nobody was promised: free HBM minus a
  // margin and minus what this plugin granted in the last ten minutes (a tenant that has not yet
  // grown into its grant must still find it). The sweep covers HBM that is free now -- not the
  // pages a tenant will later get.
  void start_health(Gpu &g, Clock::time_point tick) {
    if (g.inflight || !g.pending.empty()) return;
    uint64_t free_b = 0, total_b = 0;
    if (gs_probe_mem_info(g.probe, &free_b, &total_b) != GS_OK) return;
    const auto now = Clock::now();
    while (!g.recent_grants.empty() && now - g.recent_grants.front().first > std::chrono::seconds(kGrowWindowSeconds))
      g.recent_grants.pop_front();
    uint64_t promised = 0;
    for (const auto &rg : g.recent_grants) promised += rg.second;
    uint64_t bytes = free_b > promised ? free_b - promised : 0;
    bytes = bytes > kHealthMargin ? (bytes - kHealthMargin) & ~0xFFFFFull : 0;
    if (bytes < (1ull << 30)) {
      sentinel_check(g);  // nothing left to sweep: read the resident range back, allocating nothing
      return;
    }
    void *d = nullptr;
    if (gs_probe_mem_alloc(g.probe, bytes, &d) != GS_OK) {
      logv(1, "health probe skipped: %s", gs_cuda_last_error());
      sentinel_check(g);
      return;
    }
    gs_slice s;
    memset(&s, 0, sizeof(s));
    s.base = d;
    s.bytes = bytes;
    gs_probe_next_seed(g.probe, &s.seed);
    s.pattern = GS_PROBE_PATTERN_HASH;  // the strong pattern: nobody is waiting for this sweep
    g.arenas.push_back(d);
    g.tags.push_back(-1);
    if (gs_probe_submit(g.probe, &s, 1, GS_PROBE_WRITE | GS_PROBE_VERIFY) != GS_OK) {
      logv(1, "health probe skipped: %s", gs_cuda_last_error());
      release(g);
      return;
    }
    g.inflight = true;
    g.inflight_is_health = true;
    g.health_tick = tick;
  }
  // A GPU granted out to the last unit leaves the health checker nothing to allocate. A small range
  // taken and written once while the GPU was still empty is read back instead through the entry
  // point that allocates nothing (gs_probe_verify_noalloc): the whole path -- launch, loads from
  // HBM, compare, verdict -- still answers, and the range itself must still hold what was written.
  void sentinel_setup(Gpu &g) {
    void *d = nullptr;
    if (gs_probe_mem_alloc(g.probe, kSentinelBytes, &d) != GS_OK) return;
    gs_slice s;
    memset(&s, 0, sizeof(s));
    s.base = d;
    s.bytes = kSentinelBytes;
    gs_probe_next_seed(g.probe, &s.seed);
    s.pattern = GS_PROBE_PATTERN_HASH;
    gs_slice_result r;
    if (gs_probe_run(g.probe, &s, 1, GS_PROBE_WRITE | GS_PROBE_VERIFY, &r, nullptr) != GS_OK || r.mismatches) {
      logv(1, "health probe: gpu %d: sentinel range could not be written", g.index);
      gs_probe_mem_free(g.probe, d);
      return;
    }
    g.sentinel = d;
    g.sentinel_seed = s.seed;
  }
  void sentinel_check(Gpu &g) {
    if (!g.sentinel) return;
    gs_slice s;
    memset(&s, 0, sizeof(s));
    s.base = g.sentinel;
    s.bytes = kSentinelBytes;
    s.seed = g.sentinel_seed;
    s.pattern = GS_PROBE_PATTERN_HASH;
    gs_slice_result r;
    const int rc = gs_probe_verify_noalloc(g.cuda_ordinal, nullptr, &s, 1, &r);
    if (rc != GS_OK) {
      logv(1, "health probe: gpu %d: %s", g.index, gs_cuda_last_error());
      return;
    }
    logv(1, "health probe: gpu %d sentinel of %llu bytes read back without allocating, %llu mismatching words",
         g.index, (unsigned long long)kSentinelBytes, (unsigned long long)r.mismatches);
    if (r.mismatches) mark_bad(g.index);
  }
  void run(Gpu &g) {
    sentinel_setup(g);
    long ticks_done = 0;
    while (!stop.load()) {
      const bool busy = g.inflight || !g.pending.empty();
      // the next health tick of every GPU falls on the same instant: epoch + k * interval
      auto wake = Clock::now() + std::chrono::milliseconds(busy ? 2 : 200);
      Clock::time_point next_tick{};
      if (health_interval > 0) {
        next_tick = epoch + std::chrono::seconds(health_interval * (ticks_done + 1));
        if (next_tick < wake) wake = next_tick;
      }
      {
        std::unique_lock<std::mutex> lk(g.mu);
        // while a sweep is in flight come back soon to collect it; otherwise sleep until a grant arrives
        g.cv.wait_until(lk, wake, [&] { return stop.load() || !g.queue.empty(); });
        while (!g.queue.empty()) {
          const ProbeJob j = g.queue.front();
          g.queue.pop_front();
          g.recent_grants.emplace_back(Clock::now(), j.bytes);
          if (g.pending.empty()) g.first_pending = Clock::now();
          if (g.pending.size() >= GS_PROBE_MAX_SLICES) {  // HBM cannot be swept as fast as it is granted:
            g.pending.pop_front();                        // the newest grants win, the rest are counted
            ++g.dropped;
          }
          g.pending.push_back(j);
        }
      }
      harvest(g, false);
      start_grants(g);
      if (health_interval > 0 && Clock::now() >= next_tick) {
        // ticks missed while a long sweep ran are skipped, not queued up
        while (epoch + std::chrono::seconds(health_interval * (ticks_done + 1)) <= Clock::now()) ++ticks_done;
        start_health(g, epoch + std::chrono::seconds(health_interval * ticks_done));
      }
    }
    harvest(g, true);
    if (g.sentinel) gs_probe_mem_free(g.probe, g.sentinel);
    gs_probe_close(g.probe);
    g.probe = nullptr;
  }
  void start() {
    epoch = Clock::now();
    for (auto &kv : gpus) {
      Gpu *g = kv.second.get();
      g->th = std::thread([this, g] { run(*g); });
    }
  }
  void shutdown() {  // idempotent
    stop.store(true);
    for (auto &kv : gpus) {
      {
        std::lock_guard<std::mutex> lk(kv.second->mu);
      }
      kv.second->cv.notify_all();
    }
    for (auto &kv : gpus) {
      if (kv.second->th.joinable()) {
        kv.second->th.join();  // the thread closed its engine on the way out
      } else if (kv.second->probe) {
        gs_probe_close(kv.second->probe);  // never started: nothing else owns it
        kv.second->probe = nullptr;
      }
    }
  }
  ~ProbeWorker() { shutdown(); }
  void push(const ProbeJob &j) {
    auto it = gpus.find(j.gpu);
    if (it == gpus.end()) return;
    {
      std::lock_guard<std::mutex> lk(it->second->mu);
      it->second->queue.push_back(j);
    }
    it->second->cv.notify_one();
  }
};

// ---- one NewNvidiaDevicePlugin + Serve (gpumanager.go:70-79): what a restart rebuilds ------------
struct PluginInstance {
  const Flags &f;
  std::string node;
  int unit;
  std::vector<gs_gpu_info> infos;
  gs_table *table = nullptr;
  gs_plugin *plugin = nullptr;
  gs_server *server = nullptr;
  std::string socket_path, kubelet_socket;
  bool cluster = false;
  bool cgpu_disable = false;
  Transport transport;
  PodLister lister;
  std::mutex lock;  // m.Lock(), allocate.go:60
  std::vector<Pod> pods;
  std::vector<gs_pod_view> views;
  std::unique_ptr<ProbeWorker> probes;
  uint64_t unit_bytes;
  std::thread health_thread;
  std::atomic<bool> stop{false};
  bool health_open = false;

  PluginInstance(const Flags &flags, const std::string &node_name, int mem_unit)
      : f(flags), node(node_name), unit(mem_unit), unit_bytes(mem_unit == GS_UNIT_MIB ? 1ull << 20 : 1ull << 30) {}

  // getDevices + NewNvidiaDevicePlugin (nvidia.go:50-86, server.go:43-78); exit code 1 on failure
  bool init(std::string &err) {
    uint32_t n = 0;
    if (gs_nvml_device_count(&n) != GS_OK) {
      err = std::string("nvml: ") + gs_nvml_last_error();
      return false;
    }
    infos.resize(n);
    for (uint32_t i = 0; i < n; ++i)
      if (gs_nvml_device_info(i, &infos[i]) != GS_OK) {
        err = std::string("nvml: ") + gs_nvml_last_error();
        return false;
      }
    // nvidia.go:12-15,67-69: gpuMemory is a process global, set once from the first GPU and kept
    // across plugin restarts (SIGHUP, kubelet restart) -- a restart re-reads UUIDs, never the unit count.
    static uint64_t g_gpu_memory = 0;
    if (g_gpu_memory == 0) {
      if (gs_table_build_from_gpus(infos.data(), n, unit, &table) != GS_OK) {
        err = "cannot build the device table";
        return false;
      }
      g_gpu_memory = gs_table_units_per_gpu(table);
    } else {
      std::vector<const char *> uu(n);
      std::vector<uint32_t> ul(n);
      for (uint32_t i = 0; i < n; ++i) {
        uu[i] = infos[i].uuid;
        ul[i] = (uint32_t)strlen(infos[i].uuid);
      }
      if (gs_table_build(uu.data(), ul.data(), n, g_gpu_memory, unit, &table) != GS_OK) {
        err = "cannot build the device table";
        return false;
      }
    }
    socket_path = f.socket_dir + "/" GS_SERVER_SOCK_NAME;
    kubelet_socket = f.socket_dir + "/kubelet.sock";
    if (f.pod_source_file.empty()) {
      cluster = true;
      std::string token = f.token;
      if (f.client_cert.empty() && f.client_key.empty() && token.empty()) {  // main.go:28-36
        if (!read_file(std::string(kSaDir) + "/token", token)) {
          fprintf(stderr, "in cluster mode, find token failed, error: open %s/token: no such file or directory\n",
                  kSaDir);
          exit(1);
        }
      }
      if (!ApiConfig::load(transport.api, err)) return false;
      char url[320];
      snprintf(url, sizeof(url), "https://%s:%ld", f.kubelet_address.c_str(), f.kubelet_port);
      transport.kubelet_url = url;
      transport.kubelet_token = token;
      transport.kubelet_tls.insecure = true;  // main.go:41
      transport.kubelet_tls.cert_file = f.client_cert;
      transport.kubelet_tls.key_file = f.client_key;
      transport.timeout_s = (double)f.timeout;
      if (const char *u = getenv("GS_KUBELET_URL")) transport.kubelet_url = u;  // tests: a plain-http kubelet
      if (!patch_gpu_count(transport, node, (int)n, err)) return false;              // server.go:57-60
      if (!disable_cgpu_isolation_or_not(transport, node, cgpu_disable, err)) return false;  // :61-64
      lister.t = &transport;
      lister.node_name = node;
      lister.query_kubelet = f.query_kubelet;
    }
    if (gs_plugin_new(table, node.data(), node.size(), cgpu_disable ? 1 : 0, &plugin) != GS_OK) {
      err = "cannot create the plugin";
      return false;
    }
    if (f.probe) {
      std::vector<gs_gpu_info> mapped = infos;
      if (gs_cuda_map_ordinals(mapped.data(), (uint32_t)mapped.size()) == GS_OK) {
        probes.reset(new ProbeWorker());
        probes->table = table;
        probes->plugin = plugin;
        probes->health_interval = f.health_check ? f.health_probe_interval : 0;
        probes->gather_ms = f.probe_gather_ms < 0 ? 0 : f.probe_gather_ms;
        for (const gs_gpu_info &g : mapped) {
          gs_probe *p = nullptr;
          if (g.cuda_ordinal >= 0 && gs_probe_open(g.cuda_ordinal, &p) == GS_OK)
            probes->add_gpu(g.nvml_index, g.cuda_ordinal, g.total_bytes, p);
        }
      } else {
        logv(1, "probe disabled: %s", gs_cuda_last_error());
      }
    }
    return true;
  }

  bool load_pod_file(std::string &err) {
    std::string text;
    JVal doc;
    if (!read_file(f.pod_source_file, text) || !json_parse(text, doc)) {
      err = "cannot read " + f.pod_source_file;
      return false;
    }
    const JVal *items = doc.type == JVal::Arr ? &doc : doc.get("items");
    pods.clear();
    if (items && items->type == JVal::Arr)
      for (const JVal &p : items->arr) {
        pods.emplace_back();
        if (!pod_from_json(p, pods.back(), err)) return false;
      }
    return true;
  }

  // Allocate (allocate.go:43-201) with this daemon's pod source; runs on the serving thread
  int allocate(const uint8_t *req, size_t req_len, uint8_t *resp, size_t cap, size_t *resp_len) {
    std::lock_guard<std::mutex> lk(lock);
    gs_alloc_result res;
    memset(&res, 0, sizeof(res));
    int rc;
    if (!cluster) {
      std::string err;
      if (load_pod_file(err)) {  // re-read per Allocate, like a LIST
        pod_views(pods, views);
        gs_plugin_set_pods(plugin, views.data(), views.size());
      }
      rc = gs_plugin_allocate(plugin, req, req_len, resp, cap, &res);
    } else {
      gs_alloc_ctx ctx;
      memset(&ctx, 0, sizeof(ctx));
      ctx.table = table;
      ctx.node_name = gs_str{node.data(), node.size()};
      ctx.disable_cgpu_isolation = cgpu_disable ? 1 : 0;
      struct timespec ts;
      clock_gettime(CLOCK_REALTIME, &ts);
      ctx.now_unix_nano = (int64_t)ts.tv_sec * 1000000000ll + ts.tv_nsec;
      std::string err;
      if (!lister.list(pods, err)) {  // allocate.go:64-67
        logv(1, "invalid allocation requst: Failed to find candidate pods due to %s", err.c_str());
        rc = gs_allocate_error_response(&ctx, req, req_len, GS_ALLOC_ERR_PODS, resp, cap, &res);
      } else {
        pod_views(pods, views);
        char patch[256];
        rc = gs_allocate(&ctx, req, req_len, views.data(), views.size(), resp, cap, patch, sizeof(patch), &res);
        if (rc == GS_OK && res.kind == GS_ALLOC_MATCHED) {
          const Pod &p = pods[(size_t)res.pod_index];
          if (!patch_pod_assigned(transport, p.ns, p.name, std::string(patch, res.patch_len), err)) {
            logv(1, "Failed due to %s", err.c_str());  // allocate.go:138-151
            rc = gs_allocate_error_response(&ctx, req, req_len, GS_ALLOC_ERR_PATCH, resp, cap, &res);
          }
        }
      }
    }
    *resp_len = res.resp_len;
    if (rc == GS_OK && res.kind == GS_ALLOC_MATCHED && probes && probes->has(res.gpu_index))
      probes->push(ProbeJob{res.gpu_index, res.pod_req_units * unit_bytes, res.pod_index});
    return rc;
  }
  static int allocate_cb(void *user, const uint8_t *req, size_t req_len, uint8_t *resp, size_t cap,
                         size_t *resp_len) {
    return static_cast<PluginInstance *>(user)->allocate(req, req_len, resp, cap, resp_len);
  }

  // Serve (server.go:232-249) + healthcheck (server.go:211-229)
  bool serve(std::string &err) {
    gs_server_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.socket_path = socket_path.c_str();
    cfg.allocate = &PluginInstance::allocate_cb;
    cfg.user = this;
    const int rc = gs_server_start(plugin, &cfg, &server);
    if (rc != GS_OK) {
      err = std::string("could not start device plugin: ") + gs_strerror(rc);
      return false;
    }
    logv(1, "Starting to serve on %s", socket_path.c_str());
    if (!f.no_register) {
      const int rr = gs_register(kubelet_socket.c_str(), GS_SERVER_SOCK_NAME, GS_RESOURCE_NAME, 5000);
      if (rr != GS_OK) {
        err = std::string("Could not register device plugin: ") + gs_grpc_last_error();
        gs_server_stop(server);
        server = nullptr;
        return false;
      }
      logv(1, "Registered device plugin with Kubelet");
    }
    if (probes) probes->start();
    if (f.health_check) {
      uint64_t bad_mask = 0;
      if (gs_health_open(&bad_mask) == GS_OK) {
        health_open = true;
        for (size_t g = 0; g < infos.size(); ++g)
          if (bad_mask >> g & 1) {  // nvidia.go:107-111
            gs_law_mark_unhealthy_gpu(table, (int32_t)g);
            gs_plugin_notify_health(plugin);
          }
        health_thread = std::thread([this] {
          bool complained = false;
          while (!stop.load()) {
            gs_health_event ev;
            const int rc = gs_health_wait(1000, &ev);
            if (rc == GS_OK && !ev.ignored) {
              gs_law_mark_unhealthy_gpu(table, ev.gpu_index);
              gs_plugin_notify_health(plugin);
            } else if (rc != GS_OK && rc != GS_E_TIMEOUT) {
              // the event wait itself fails (GPU lost, driver reload): that is not a quiet second.
              // Say so once, treat it as an event without a device (every GPU unhealthy, as
              // nvidia.go:138-143 does for one), and do not spin on the error.
              if (!complained) {
                logv(0, "health watch: %s; marking all GPUs unhealthy", gs_nvml_last_error());
                complained = true;
                gs_law_mark_unhealthy_gpu(table, GS_ALL_GPUS);
                gs_plugin_notify_health(plugin);
              }
              for (int i = 0; i < 10 && !stop.load(); ++i) usleep(100 * 1000);
            }
          }
        });
      }
    }
    return true;
  }

  void shutdown() {  // Stop (server.go:145-155)
    stop.store(true);
    if (server) {
      gs_server_stop(server);
      server = nullptr;
    }
    if (health_thread.joinable()) health_thread.join();
    if (health_open) gs_health_close();
    if (probes) {
      probes->shutdown();
      probes.reset();
    }
    if (plugin) gs_plugin_free(plugin);
    if (table) gs_table_free(table);
    plugin = nullptr;
    table = nullptr;
  }
};

// coredump.go:10-30, for this process: the stack of the supervising thread
void coredump(const std::string &path) {
  FILE *fp = fopen(path.c_str(), "w");
  if (!fp) {
    fprintf(stderr, "failed to write %s: %s\n", path.c_str(), strerror(errno));
    return;
  }
  void *frames[64];
  const int n = backtrace(frames, 64);
  fflush(fp);
  backtrace_symbols_fd(frames, n, fileno(fp));
  fclose(fp);
}

int g_sig_pipe[2] = {-1, -1};
void on_signal(int signum) {
  const unsigned char b = (unsigned char)signum;
  if (write(g_sig_pipe[1], &b, 1) < 0) {
  }
}

[[noreturn]] void wait_forever() {  // select {} (gpumanager.go:39,46)
  for (;;) pause();
}

}  // namespace

int main(int argc, char **argv) {
  const Flags f = parse_flags(argc, argv);
  g_v = (int)f.v;
  int unit = GS_UNIT_GIB, warned = 0;
  gs_translate_memory_unit(f.memory_unit.data(), f.memory_unit.size(), &unit, &warned);  // main.go:67-78
  if (warned) logv(0, "Unsupported memory unit: %s, use memoryUnit Gi as default", f.memory_unit.c_str());
  const char *node_env = getenv("NODE_NAME");
  if (!node_env || !*node_env) {
    fprintf(stderr, "Please set env NODE_NAME\n");  // podmanager.go:55-58
    return 1;
  }
  const std::string node = node_env;
  logv(1, "Start gpushare device plugin");
  // ---- sharedGPUManager.Run (gpumanager.go:33-111) ----------------------------------------
  logv(1, "Loading NVML");
  if (gs_nvml_open(f.nvml_lib.empty() ? nullptr : f.nvml_lib.c_str()) != GS_OK) {
    logv(1, "Failed to initialize NVML: %s.", gs_nvml_last_error());
    logv(1, "If this is a GPU node, did you set the docker default runtime to `nvidia`?");
    wait_forever();
  }
  logv(1, "Fetching devices.");
  uint32_t count = 0;
  if (gs_nvml_device_count(&count) != GS_OK || count == 0) {
    logv(1, "No devices found. Waiting indefinitely.");
    wait_forever();
  }
  logv(1, "Starting FS watcher.");
  const int ino = inotify_init1(IN_NONBLOCK | IN_CLOEXEC);  // newFSWatcher (watchers.go:10-25)
  if (ino < 0 || inotify_add_watch(ino, f.socket_dir.c_str(), IN_CREATE | IN_DELETE | IN_MOVED_TO) < 0) {
    logv(0, "Failed to created FS watcher.");
    return 1;
  }
  logv(1, "Starting OS watcher.");
  if (pipe(g_sig_pipe) != 0) return 1;
  struct sigaction sa;
  memset(&sa, 0, sizeof(sa));
  sa.sa_handler = on_signal;
  for (int s : {SIGHUP, SIGINT, SIGTERM, SIGQUIT}) sigaction(s, &sa, nullptr);
  signal(SIGPIPE, SIG_IGN);

  std::unique_ptr<PluginInstance> inst;
  bool restart = true;
  for (;;) {
    if (restart) {  // gpumanager.go:66-79
      if (inst) inst->shutdown();
      inst.reset(new PluginInstance(f, node, unit));
      std::string err;
      if (!inst->init(err)) {
        logv(0, "Failed to get device plugin due to %s", err.c_str());
        return 1;  // :71-73
      }
      if (!inst->serve(err)) {
        logv(0, "Failed to start device plugin due to %s", err.c_str());
        return 2;  // :74-76
      }
      restart = false;
    }
    struct pollfd pf[2] = {{ino, POLLIN, 0}, {g_sig_pipe[0], POLLIN, 0}};
    const int r = poll(pf, 2, 1000);
    if (r <= 0) continue;
    if (pf[0].revents & POLLIN) {  // :82-87
      alignas(struct inotify_event) char buf[8192];
      const ssize_t n = read(ino, buf, sizeof(buf));
      for (ssize_t off = 0; off + (ssize_t)sizeof(struct inotify_event) <= n;) {
        const struct inotify_event *ev = (const struct inotify_event *)(buf + off);
        if (ev->len && (ev->mask & IN_CREATE) && strcmp(ev->name, "kubelet.sock") == 0) {
          logv(1, "inotify: %s created, restarting.", GS_KUBELET_SOCKET);
          restart = true;
        }
        off += (ssize_t)sizeof(struct inotify_event) + ev->len;
      }
    }
    if (pf[1].revents & POLLIN) {
      unsigned char sigs[64];
      const ssize_t n = read(g_sig_pipe[0], sigs, sizeof(sigs));
      for (ssize_t i = 0; i < n; ++i) {
        if (sigs[i] == SIGHUP) {  // :94-96
          logv(1, "Received SIGHUP, restarting.");
          restart = true;
        } else if (sigs[i] == SIGQUIT) {  // :97-101
          logv(0, "generate core dump");
          char name[96];
          time_t now = time(nullptr);
          struct tm tmv;
          localtime_r(&now, &tmv);
          strftime(name, sizeof(name), "/etc/kubernetes/go_%Y%m%d%H%M%S.txt", &tmv);
          coredump(name);
        } else {  // :102-106
          logv(1, "Received signal \"%d\", shutting down.", (int)sigs[i]);
          inst->shutdown();
          close(ino);
          logv(1, "Shutdown of NVML returned: %d", gs_nvml_close());
          return 0;
        }
      }
    }
  }
}

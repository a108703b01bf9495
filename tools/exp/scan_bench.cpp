// Exploratory: how fast can the id-count hop chain go on one core? (host side)
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <chrono>
#include <random>
#include <string>
#include <vector>
static inline bool fast_record(const uint8_t *q, const uint8_t *qend) {
  return qend - q >= 2 && q[0] == 0x0A && q[1] < 0x80 && (size_t)(qend - q) >= 2u + q[1];
}
static uint32_t seq(const uint8_t *q, const uint8_t *qend) {
  uint32_t c = 0;
  while (q < qend) { if (!fast_record(q, qend)) return ~0u; q += 2u + q[1]; ++c; }
  return c;
}
template <int K, int PF>
static uint32_t multi(const uint8_t *q0, const uint8_t *qend) {
  const size_t len = qend - q0;
  const uint8_t *start[K + 1]; start[0] = q0; start[K] = qend;
  for (int c = 1; c < K; ++c) {
    const uint8_t *g = q0 + len / K * c, *lim = g + 512 < qend ? g + 512 : qend, *f = nullptr;
    for (; g < lim; ++g) { if (!fast_record(g, qend)) continue; const uint8_t *n1 = g + 2 + g[1];
      if (n1 != qend && !fast_record(n1, qend)) continue; f = g; break; }
    if (!f || f <= start[c - 1]) return ~0u; start[c] = f;
  }
  const uint8_t *p[K]; uint32_t cnt[K];
  for (int c = 0; c < K; ++c) { p[c] = start[c]; cnt[c] = 0; }
  // lockstep phase: all chains alive; no per-chain liveness branch
  for (;;) {
    bool all = true;
    for (int c = 0; c < K; ++c) all &= (p[c] < start[c + 1]);
    if (!all) break;
#pragma GCC unroll 8
    for (int c = 0; c < K; ++c) {
      if (PF) __builtin_prefetch(p[c] + PF);
      if (!fast_record(p[c], qend)) return ~0u;
      p[c] += 2u + p[c][1]; ++cnt[c];
    }
  }
  uint32_t total = 0;
  for (int c = 0; c < K; ++c) {
    while (p[c] < start[c + 1]) { if (!fast_record(p[c], qend)) return ~0u; p[c] += 2u + p[c][1]; ++cnt[c]; }
    if (p[c] != start[c + 1]) return ~0u; total += cnt[c];
  }
  return total;
}
template <class F> static void run(const char *name, const std::vector<std::vector<uint8_t>> &bufs, F f) {
  uint64_t sum = 0; size_t bytes = 0;
  auto t0 = std::chrono::steady_clock::now();
  int reps = 200;
  for (int r = 0; r < reps; ++r) for (auto &b : bufs) { sum += f(b.data(), b.data() + b.size()); bytes += b.size(); }
  double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  printf("%-22s %7.2f GB/s  %6.2f ns/id  (ids %llu)\n", name, bytes / dt / 1e9, dt * 1e9 / sum, (unsigned long long)sum / reps);
}
int main() {
  std::mt19937_64 rng(1);
  for (int n : {1024, 16384, 40960}) {
    std::vector<std::vector<uint8_t>> bufs;
    for (int k = 0; k < 8; ++k) {   // 8 distinct requests so the working set is like a wave
      std::vector<uint8_t> b;
      for (int i = 0; i < n; ++i) {
        char id[64]; int l = snprintf(id, sizeof id, "GPU-%08x-0000-0000-0000-%012x-_-%d", k, k, (int)(rng() % 183359));
        b.push_back(0x0A); b.push_back((uint8_t)l); b.insert(b.end(), id, id + l);
      }
      bufs.push_back(b);
    }
    printf("-- %d ids per request, %zu bytes each\n", n, bufs[0].size());
    run("sequential", bufs, seq);
    run("multi K=2", bufs, multi<2, 0>);
    run("multi K=4", bufs, multi<4, 0>);
    run("multi K=8", bufs, multi<8, 0>);
    run("multi K=4 pf256", bufs, multi<4, 256>);
    run("multi K=8 pf512", bufs, multi<8, 512>);
  }
}

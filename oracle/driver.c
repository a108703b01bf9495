/*
 * oracle/driver.c -- C timing loops for the Allocate sweep (tools/allocate_sweep.py).
 * TEST INFRASTRUCTURE ONLY. Calling the C ABI once per decision from Python costs ~1.5 us per
 * call, which is most of a small decision; these loops call the product through function
 * pointers handed over by ctypes (this file does not link the product) so neither side of the
 * comparison pays for the interpreter.
 */
#include <stddef.h>
#include <stdint.h>
#include <time.h>

typedef int (*gs_set_pods_fn)(void *plugin, const void *views, size_t n);
typedef int (*gs_allocate_fn)(void *plugin, const uint8_t *req, size_t len, uint8_t *resp, size_t cap,
                              void *result);

static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* Rounds of {install the pod list, `per_round` decisions on `wire`} for about budget_s seconds.
 * Returns decisions per second; *failed counts calls that did not return 0 / kind 0. */
double gs_drive_rounds(gs_set_pods_fn set_pods, gs_allocate_fn allocate, void *plugin, const void *views,
                       size_t n_pods, const uint8_t *wire, size_t len, int per_round, double budget_s,
                       uint64_t *failed) {
  uint8_t resp[4096];
  int32_t result[16]; /* gs_alloc_result: kind is the first int32 */
  uint64_t done = 0, bad = 0;
  const double t0 = now_s();
  double t = t0;
  while (t - t0 < budget_s) {
    if (set_pods(plugin, views, n_pods) != 0) ++bad;
    for (int k = 0; k < per_round; ++k) {
      if (allocate(plugin, wire, len, resp, sizeof resp, result) != 0 || result[0] != 0) ++bad;
      ++done;
    }
    t = now_s();
  }
  if (failed) *failed = bad;
  return (double)done / (t - t0);
}

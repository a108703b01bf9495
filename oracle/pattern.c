/*
 * oracle/pattern.c -- CPU restatement of the slice-probe pattern and verdict.
 *
 * TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
 * cpu_baseline leg of bench.py. The product (libgpushare_core.so) never links or
 * calls anything in oracle/.
 *
 * The probe kernel has NO counterpart in the
 * reference (the reference runs no GPU code), so there is no reference file:line
 * to follow here and no reference golden vector: this file is the written
 * specification of pattern(i, seed) -- both forms -- that the CUDA kernels must
 * reproduce bit for bit, pinned by the known-answer vectors in
 * tests/golden/pattern_kat.json.
 *
 *   i    = index of the 16-byte word inside the slice (slice byte offset / 16)
 *   seed = per-slice 32-bit seed
 */
#include <stddef.h>
#include <stdint.h>

/* background(seed): four 32-bit words drawn once per sweep. */
static void background(uint32_t seed, uint32_t r[4]) {
  uint32_t x = seed * 0x9E3779B1u + 0x7F4A7C15u;
  x ^= x >> 15;
  x *= 0x2C1B3C6Du;
  x ^= x >> 12;
  r[0] = x;
  r[1] = r[0] * 0x85EBCA77u + 1u;
  r[2] = r[1] * 0xC2B2AE3Du + 1u;
  r[3] = r[2] * 0x27D4EB2Fu + 1u;
}

static inline uint32_t rotl32(uint32_t x, unsigned r) { return (x << r) | (x >> (32u - r)); }

/* Four little-endian 32-bit words of the 16-byte pattern word i, for either pattern
 * (include/gpushare.h GS_PROBE_PATTERN_*; why two: csrc/probe_kernels.cuh).
 *   0 LOWTOGGLE  { r0 ^ lo(i), r1 ^ hi(i), r2, r3 }: a per-sweep random background with the
 *                word's address in it
 *   1 HASH       { h, ~h, h * 0x297A2D39 + lo(i), rotl(h, 13) ^ hi(i) ^ 0xC2B2AE3D } with h an
 *                avalanche hash of (i, seed)
 * Any other pattern code leaves out[] untouched and returns -1. */
int gs_oracle_pattern_word_p(uint64_t i, uint32_t seed, uint32_t pattern, uint32_t out[4]) {
  const uint32_t lo = (uint32_t)i;
  const uint32_t hi = (uint32_t)(i >> 32);
  if (pattern == 0) {
    uint32_t r[4];
    background(seed, r);
    out[0] = r[0] ^ lo;
    out[1] = r[1] ^ hi;
    out[2] = r[2];
    out[3] = r[3];
    return 0;
  }
  if (pattern == 1) {
    uint32_t h = lo * 0x9E3779B1u + (hi ^ seed) * 0x85EBCA77u + seed;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 12;
    out[0] = h;
    out[1] = ~h;
    out[2] = h * 0x297A2D39u + lo;
    out[3] = rotl32(h, 13) ^ hi ^ 0xC2B2AE3Du;
    return 0;
  }
  return -1;
}

void gs_oracle_pattern_word(uint64_t i, uint32_t seed, uint32_t out[4]) {
  gs_oracle_pattern_word_p(i, seed, 0, out);
}

/* Fill n_words 16-byte words starting at word index first. */
void gs_oracle_pattern_fill_p(uint32_t *buf, uint64_t first, uint64_t n_words, uint32_t seed,
                              uint32_t pattern) {
  for (uint64_t k = 0; k < n_words; ++k)
    gs_oracle_pattern_word_p(first + k, seed, pattern, buf + 4 * k);
}

void gs_oracle_pattern_fill(uint32_t *buf, uint64_t first, uint64_t n_words, uint32_t seed) {
  gs_oracle_pattern_fill_p(buf, first, n_words, seed, 0);
}

/* Verdict of a read-back: counts 16-byte words that differ, total flipped bits and
 * the lowest differing byte offset (UINT64_MAX when clean) -- the same three
 * numbers gs_slice_result carries. `first` is the word index buf[0] stands for. */
void gs_oracle_pattern_check_p(const uint32_t *buf, uint64_t first, uint64_t n_words,
                               uint32_t seed, uint32_t pattern, uint64_t *mismatches,
                               uint64_t *first_bad_byte, uint64_t *bit_flips) {
  uint64_t mm = 0, fb = UINT64_MAX, flips = 0;
  for (uint64_t k = 0; k < n_words; ++k) {
    uint32_t want[4];
    gs_oracle_pattern_word_p(first + k, seed, pattern, want);
    int bad = 0;
    for (int w = 0; w < 4; ++w) {
      uint32_t d = buf[4 * k + w] ^ want[w];
      if (!d) continue;
      flips += (uint64_t)__builtin_popcount(d);
      if (!bad) {
        bad = 1;
        if (fb == UINT64_MAX)
          fb = (first + k) * 16u + (uint64_t)w * 4u + (uint64_t)(__builtin_ctz(d) >> 3);
      }
    }
    mm += (uint64_t)bad;
  }
  *mismatches = mm;
  *first_bad_byte = fb;
  *bit_flips = flips;
}

void gs_oracle_pattern_check(const uint32_t *buf, uint64_t first, uint64_t n_words,
                             uint32_t seed, uint64_t *mismatches, uint64_t *first_bad_byte,
                             uint64_t *bit_flips) {
  gs_oracle_pattern_check_p(buf, first, n_words, seed, 0, mismatches, first_bad_byte, bit_flips);
}

/* CPU write+verify over host memory: the cpu_baseline for probe GB/s. Returns the
 * number of mismatching words (0 on sane host memory). */
uint64_t gs_oracle_probe_host(uint32_t *buf, uint64_t n_words, uint32_t seed) {
  uint64_t mm, fb, flips;
  gs_oracle_pattern_fill(buf, 0, n_words, seed);
  gs_oracle_pattern_check(buf, 0, n_words, seed, &mm, &fb, &flips);
  return mm;
}

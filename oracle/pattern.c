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
 * specification of pattern(i, seed) that the CUDA kernel must reproduce bit for
 * bit, pinned by the known-answer vectors in tests/golden/pattern_kat.json.
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

/* Four little-endian 32-bit words of the 16-byte pattern word i:
 * { r0 ^ lo(i), r1 ^ hi(i), r2, r3 } -- a per-sweep random background with the word's
 * address in it (why this shape: csrc/probe.cu). */
void gs_oracle_pattern_word(uint64_t i, uint32_t seed, uint32_t out[4]) {
  uint32_t r[4];
  background(seed, r);
  out[0] = r[0] ^ (uint32_t)i;
  out[1] = r[1] ^ (uint32_t)(i >> 32);
  out[2] = r[2];
  out[3] = r[3];
}

/* Fill n_words 16-byte words starting at word index first. */
void gs_oracle_pattern_fill(uint32_t *buf, uint64_t first, uint64_t n_words, uint32_t seed) {
  for (uint64_t k = 0; k < n_words; ++k) gs_oracle_pattern_word(first + k, seed, buf + 4 * k);
}

/* Verdict of a read-back: counts 16-byte words that differ, total flipped bits and
 * the lowest differing byte offset (UINT64_MAX when clean) -- the same three
 * numbers gs_slice_result carries. `first` is the word index buf[0] stands for. */
void gs_oracle_pattern_check(const uint32_t *buf, uint64_t first, uint64_t n_words,
                             uint32_t seed, uint64_t *mismatches, uint64_t *first_bad_byte,
                             uint64_t *bit_flips) {
  uint64_t mm = 0, fb = UINT64_MAX, flips = 0;
  for (uint64_t k = 0; k < n_words; ++k) {
    uint32_t want[4];
    gs_oracle_pattern_word(first + k, seed, want);
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

/* CPU write+verify over host memory: the cpu_baseline for probe GB/s. Returns the
 * number of mismatching words (0 on sane host memory). */
uint64_t gs_oracle_probe_host(uint32_t *buf, uint64_t n_words, uint32_t seed) {
  uint64_t mm, fb, flips;
  gs_oracle_pattern_fill(buf, 0, n_words, seed);
  gs_oracle_pattern_check(buf, 0, n_words, seed, &mm, &fb, &flips);
  return mm;
}

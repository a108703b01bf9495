// gosem.cpp -- the Go / Kubernetes library semantics the reference's hot path leans
// on, restated for the C ABI (sources are dependencies, not in the reference tree):
//   strconv.Atoi / strconv.ParseUint   Go 1.21 (go.mod:3), used at podutils.go:44,66
//   resource.Quantity parse + Value()  k8s.io/apimachinery v0.28.4 (go.mod:12),
//                                      used at podutils.go:126-127
//   translatememoryUnits               cmd/nvidia/main.go:67-78
//   setGPUMemory                       pkg/gpu/nvidia/nvidia.go:31-38
#include <string.h>

#include "gpushare.h"
#include "internal.h"

namespace gs_internal {

// digits only, base 10, no sign/underscore; *overflow set when > UINT64_MAX
bool parse_u64_digits(const char *s, size_t n, uint64_t *out, bool *overflow) {
  *overflow = false;
  if (n == 0) return false;
  uint64_t v = 0;
  for (size_t i = 0; i < n; ++i) {
    const unsigned d = (unsigned)(unsigned char)s[i] - '0';
    if (d > 9) return false;
    if (v > (UINT64_MAX - d) / 10) {
      // keep validating the syntax: Go reports syntax errors before range errors
      for (size_t k = i; k < n; ++k)
        if ((unsigned)(unsigned char)s[k] - '0' > 9) return false;
      *overflow = true;
      *out = UINT64_MAX;
      return true;
    }
    v = v * 10 + d;
  }
  *out = v;
  return true;
}

}  // namespace gs_internal

using gs_internal::parse_u64_digits;

extern "C" {

int gs_go_parse_uint64(const char *s, size_t len, uint64_t *value) {
  if (!value || (!s && len)) return GS_E_INVAL;
  bool ovf = false;
  uint64_t v = 0;
  if (!parse_u64_digits(s, len, &v, &ovf)) return GS_E_INVAL;  // strconv.ErrSyntax
  if (ovf) return GS_E_LIMIT;                                  // strconv.ErrRange
  *value = v;
  return GS_OK;
}

int gs_go_atoi(const char *s, size_t len, int64_t *value) {
  if (!value || (!s && len)) return GS_E_INVAL;
  if (len == 0) return GS_E_INVAL;
  bool neg = false;
  if (s[0] == '+' || s[0] == '-') {
    neg = s[0] == '-';
    ++s;
    --len;
  }
  bool ovf = false;
  uint64_t v = 0;
  if (!parse_u64_digits(s, len, &v, &ovf)) return GS_E_INVAL;
  if (ovf) return GS_E_LIMIT;
  if (!neg && v > (uint64_t)INT64_MAX) return GS_E_LIMIT;
  if (neg && v > (uint64_t)INT64_MAX + 1) return GS_E_LIMIT;
  *value = neg ? (int64_t)(0 - v) : (int64_t)v;
  return GS_OK;
}

// resource.Quantity: <sign><digits>[.<digits>]<suffix>, suffix = decimal SI
// (n u m "" k M G T P E), binary SI (Ki Mi Gi Ti Pi Ei) or a decimal exponent
// (e|E)[sign]digits. Value() is the integer rounded away from zero. Magnitudes that
// do not fit int64 are reported as GS_E_LIMIT instead of reproducing Go's overflow.
int gs_quantity_value(const char *s, size_t len, int64_t *value) {
  if (!value || !s || len == 0) return GS_E_INVAL;
  size_t i = 0;
  bool positive = true;
  if (s[0] == '+' || s[0] == '-') {
    positive = s[0] == '+';
    i = 1;
  }
  // mantissa as an unsigned 128-bit integer with a decimal shift
  unsigned __int128 mant = 0;
  int exp10 = 0;
  const unsigned __int128 kMantMax = ((unsigned __int128)1 << 120);
  size_t j = i;
  for (; j < len && (unsigned)(s[j] - '0') <= 9; ++j) {
    if (mant >= kMantMax / 10) return GS_E_LIMIT;
    mant = mant * 10 + (unsigned)(s[j] - '0');
  }
  if (j < len && s[j] == '.') {
    ++j;
    for (; j < len && (unsigned)(s[j] - '0') <= 9; ++j) {
      if (mant >= kMantMax / 10) return GS_E_LIMIT;
      mant = mant * 10 + (unsigned)(s[j] - '0');
        --exp10;
    }
  }
  // no digits at all is accepted: apimachinery's parser turns an empty numerator into "0"
  const char *suf = s + j;
  const size_t sl = len - j;
  int exp2 = 0;
  if (sl == 0) {
  } else if (sl == 1 && suf[0] != 'e') {
    switch (suf[0]) {
      case 'n': exp10 -= 9; break;
      case 'u': exp10 -= 6; break;
      case 'm': exp10 -= 3; break;
      case 'k': exp10 += 3; break;
      case 'M': exp10 += 6; break;
      case 'G': exp10 += 9; break;
      case 'T': exp10 += 12; break;
      case 'P': exp10 += 15; break;
      case 'E': exp10 += 18; break;
      default: return GS_E_INVAL;
    }
  } else if (sl == 2 && suf[1] == 'i') {
    switch (suf[0]) {
      case 'K': exp2 = 10; break;
      case 'M': exp2 = 20; break;
      case 'G': exp2 = 30; break;
      case 'T': exp2 = 40; break;
      case 'P': exp2 = 50; break;
      case 'E': exp2 = 60; break;
      default: return GS_E_INVAL;
    }
  } else if ((suf[0] == 'e' || suf[0] == 'E') && sl > 1) {
    size_t k = 1;
    int sign = 1;
    if (suf[k] == '+' || suf[k] == '-') {
      sign = suf[k] == '-' ? -1 : 1;
      ++k;
    }
    if (k >= sl) return GS_E_INVAL;
    int e = 0;
    for (; k < sl; ++k) {
      if ((unsigned)(suf[k] - '0') > 9) return GS_E_INVAL;
      if (e > 100000) return GS_E_LIMIT;
      e = e * 10 + (suf[k] - '0');
    }
    exp10 += sign * e;
  } else {
    return GS_E_INVAL;
  }
  unsigned __int128 v = mant;
  bool inexact = false;
  if (v != 0) {
    const unsigned __int128 kLimit = (unsigned __int128)1 << 64;
    for (; exp10 > 0; --exp10) {
      if (v >= kLimit) return GS_E_LIMIT;
      v *= 10;
    }
    if (exp2) {
      if (v >= ((unsigned __int128)1 << 64)) return GS_E_LIMIT;
      v <<= exp2;  // < 2^124
    }
    for (; exp10 < 0 && v != 0; ++exp10) {
      if (v % 10) inexact = true;
      v /= 10;
    }
    if (inexact) v += 1;  // round away from zero (a positive remainder was dropped)
  }
  if (positive) {
    if (v > (unsigned __int128)INT64_MAX) return GS_E_LIMIT;
    *value = (int64_t)v;
  } else {
    if (v > (unsigned __int128)INT64_MAX + 1) return GS_E_LIMIT;
    *value = (int64_t)(0 - (uint64_t)v);
  }
  return GS_OK;
}

int gs_translate_memory_unit(const char *s, size_t len, int *unit, int *warned) {
  if (!unit || (!s && len)) return GS_E_INVAL;
  int w = 0, u = GS_UNIT_GIB;
  if (len == 3 && memcmp(s, "MiB", 3) == 0) {
    u = GS_UNIT_MIB;
  } else if (len == 3 && memcmp(s, "GiB", 3) == 0) {
    u = GS_UNIT_GIB;
  } else {
    w = 1;  // "Unsupported memory unit ... use memoryUnit Gi as default"
  }
  *unit = u;
  if (warned) *warned = w;
  return GS_OK;
}

const char *gs_unit_name(int unit) { return unit == GS_UNIT_MIB ? "MiB" : "GiB"; }

uint64_t gs_units_from_raw(uint64_t raw_mib, int unit) {
  return unit == GS_UNIT_GIB ? raw_mib / 1024 : raw_mib;
}

}  // extern "C"

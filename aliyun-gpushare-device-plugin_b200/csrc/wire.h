// wire.h -- protobuf wire-format primitives (hand-rolled; the image has no protobuf
// C++ library). Only what deviceplugin/v1beta1 needs: varints and length-delimited
// fields. Message shapes: k8s.io/kubelet v0.28.3 pkg/apis/deviceplugin/v1beta1/api.proto
// (dependency of the reference, go.mod:14; used at server.go:126,165-170,181).
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <string.h>

namespace gs_wire {

constexpr uint32_t kVarint = 0, kFixed64 = 1, kLen = 2, kFixed32 = 5;

inline size_t varint_size(uint64_t v) {
  size_t n = 1;
  while (v >= 0x80) {
    v >>= 7;
    ++n;
  }
  return n;
}

inline uint8_t *put_varint(uint8_t *p, uint64_t v) {
  while (v >= 0x80) {
    *p++ = (uint8_t)(v | 0x80);
    v >>= 7;
  }
  *p++ = (uint8_t)v;
  return p;
}

// Reads one varint; returns nullptr on truncation or overlong encoding.
inline const uint8_t *get_varint(const uint8_t *p, const uint8_t *end, uint64_t *out) {
  uint64_t v = 0;
  for (unsigned shift = 0; shift < 64; shift += 7) {
    if (p >= end) return nullptr;
    const uint8_t b = *p++;
    v |= (uint64_t)(b & 0x7F) << shift;
    if (!(b & 0x80)) {
      *out = v;
      return p;
    }
  }
  return nullptr;
}

// Skips the value of a field with the given wire type; nullptr on malformed input.
inline const uint8_t *skip_value(const uint8_t *p, const uint8_t *end, uint32_t wt) {
  uint64_t v;
  switch (wt) {
    case kVarint: return get_varint(p, end, &v);
    case kFixed64: return (size_t)(end - p) >= 8 ? p + 8 : nullptr;
    case kFixed32: return (size_t)(end - p) >= 4 ? p + 4 : nullptr;
    case kLen:
      p = get_varint(p, end, &v);
      if (!p || v > (uint64_t)(end - p)) return nullptr;
      return p + v;
    default: return nullptr;  // groups are not used by this API
  }
}

// Bounded writer: keeps counting past the capacity so callers learn the required size.
struct Writer {
  uint8_t *buf;
  size_t cap;
  size_t len = 0;
  Writer(uint8_t *b, size_t c) : buf(b), cap(b ? c : 0) {}
  bool ok() const { return len <= cap; }
  void byte(uint8_t b) {
    if (len < cap) buf[len] = b;
    ++len;
  }
  void varint(uint64_t v) {
    while (v >= 0x80) {
      byte((uint8_t)(v | 0x80));
      v >>= 7;
    }
    byte((uint8_t)v);
  }
  void bytes(const void *p, size_t n) {
    if (len + n <= cap) memcpy(buf + len, p, n);
    len += n;
  }
  // length-delimited field holding raw bytes
  void field_bytes(uint32_t field, const void *p, size_t n) {
    varint(((uint64_t)field << 3) | kLen);
    varint(n);
    bytes(p, n);
  }
};

inline size_t field_bytes_size(uint32_t field, size_t n) {
  return varint_size(((uint64_t)field << 3) | kLen) + varint_size(n) + n;
}

}  // namespace gs_wire

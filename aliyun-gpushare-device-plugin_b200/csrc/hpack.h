// hpack.h -- HPACK header compression (RFC 7541) for the native gRPC transport.
//
// The reference gets HTTP/2 + HPACK from google.golang.org/grpc (go.mod:48, used by
// pkg/gpu/nvidia/server.go:114-177,232-249); neither image has grpc C++ or a Go toolchain,
// so the transport under the C ABI is written here from the RFCs.
//
// Decoder: complete for the representations of RFC 7541 section 6 (indexed, literal with /
// without / never indexing, dynamic table size update), with the dynamic table. Huffman
// strings (section 5.2, Appendix B) are decoded for all 256 byte values with the canonical
// form of the code, rebuilt from the 257 code lengths.
//
// Encoder: what a gRPC endpoint needs -- static-table references and literals without
// indexing, never Huffman -- so the peer's decoder keeps no state for us.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include <deque>
#include <string>
#include <utility>
#include <vector>

namespace gs_h2 {

// ---- static table (RFC 7541 Appendix A) --------------------------------------------------
struct StaticEntry {
  const char *name;
  const char *value;
};
inline const StaticEntry *static_table() {
  static const StaticEntry t[62] = {
      {"", ""},
      {":authority", ""},
      {":method", "GET"},
      {":method", "POST"},
      {":path", "/"},
      {":path", "/index.html"},
      {":scheme", "http"},
      {":scheme", "https"},
      {":status", "200"},
      {":status", "204"},
      {":status", "206"},
      {":status", "304"},
      {":status", "400"},
      {":status", "404"},
      {":status", "500"},
      {"accept-charset", ""},
      {"accept-encoding", "gzip, deflate"},
      {"accept-language", ""},
      {"accept-ranges", ""},
      {"accept", ""},
      {"access-control-allow-origin", ""},
      {"age", ""},
      {"allow", ""},
      {"authorization", ""},
      {"cache-control", ""},
      {"content-disposition", ""},
      {"content-encoding", ""},
      {"content-language", ""},
      {"content-length", ""},
      {"content-location", ""},
      {"content-range", ""},
      {"content-type", ""},
      {"cookie", ""},
      {"date", ""},
      {"etag", ""},
      {"expect", ""},
      {"expires", ""},
      {"from", ""},
      {"host", ""},
      {"if-match", ""},
      {"if-modified-since", ""},
      {"if-none-match", ""},
      {"if-range", ""},
      {"if-unmodified-since", ""},
      {"last-modified", ""},
      {"link", ""},
      {"location", ""},
      {"max-forwards", ""},
      {"proxy-authenticate", ""},
      {"proxy-authorization", ""},
      {"range", ""},
      {"referer", ""},
      {"refresh", ""},
      {"retry-after", ""},
      {"server", ""},
      {"set-cookie", ""},
      {"strict-transport-security", ""},
      {"transfer-encoding", ""},
      {"user-agent", ""},
      {"vary", ""},
      {"via", ""},
      {"www-authenticate", ""},
  };
  return t;
}
constexpr uint32_t kStaticCount = 61;

// ---- Huffman (RFC 7541 Appendix B) ----------------------------------------------------------
// The code of Appendix B is canonical: within one code length the codes are consecutive in symbol
// order, and the first code of a length continues from the last code of the length before. The
// 257 code LENGTHS therefore determine it (they add up to a complete code: sum of 2^-len == 1,
// checked by tests/cpp/hpack_kat.cpp together with the RFC's own examples and a round trip of all
// 256 byte values against nghttp2's decoder). All symbols decode, not only printable ASCII: a
// peer may Huffman-code any header byte.
constexpr int kHuffMaxLen = 30;
constexpr int kHuffEos = 256;
struct Huffman {
  uint32_t first_code[kHuffMaxLen + 1];  // first code of each length
  uint16_t first_sym[kHuffMaxLen + 1];   // index of that code's symbol in syms
  uint16_t count[kHuffMaxLen + 1];
  uint16_t syms[257];                    // symbols ordered by (length, value)
  uint32_t code_of[257];                 // for the encoder side of tests
  uint8_t len_of[257];
  Huffman() {
    static const uint8_t lengths[257] = {
      13, 23, 28, 28, 28, 28, 28, 28, 28, 24, 30, 28, 28, 30, 28, 28,  // 0..15
      28, 28, 28, 28, 28, 28, 30, 28, 28, 28, 28, 28, 28, 28, 28, 28,  // 16..31
       6, 10, 10, 12, 13,  6,  8, 11, 10, 10,  8, 11,  8,  6,  6,  6,  // 32..47
       5,  5,  5,  6,  6,  6,  6,  6,  6,  6,  7,  8, 15,  6, 12, 10,  // 48..63
      13,  6,  7,  7,  7,  7,  7,  7,  7,  7,  7,  7,  7,  7,  7,  7,  // 64..79
       7,  7,  7,  7,  7,  7,  7,  7,  8,  7,  8, 13, 19, 13, 14,  6,  // 80..95
      15,  5,  6,  5,  6,  5,  6,  6,  6,  5,  7,  7,  6,  6,  6,  5,  // 96..111
       6,  7,  6,  5,  5,  6,  7,  7,  7,  7,  7, 15, 11, 14, 13, 28,  // 112..127
      20, 22, 20, 20, 22, 22, 22, 23, 22, 23, 23, 23, 23, 23, 24, 23,  // 128..143
      24, 24, 22, 23, 24, 23, 23, 23, 23, 21, 22, 23, 22, 23, 23, 24,  // 144..159
      22, 21, 20, 22, 22, 23, 23, 21, 23, 22, 22, 24, 21, 22, 23, 23,  // 160..175
      21, 21, 22, 21, 23, 22, 23, 23, 20, 22, 22, 22, 23, 22, 22, 23,  // 176..191
      26, 26, 20, 19, 22, 23, 22, 25, 26, 26, 26, 27, 27, 26, 24, 25,  // 192..207
      19, 21, 26, 27, 27, 26, 27, 24, 21, 21, 26, 26, 28, 27, 27, 27,  // 208..223
      20, 24, 20, 21, 22, 21, 21, 23, 22, 22, 25, 25, 24, 24, 26, 23,  // 224..239
      26, 27, 26, 26, 27, 27, 27, 27, 27, 28, 27, 27, 27, 27, 27, 26,  // 240..255
      30,  // 256 = EOS
    };
    uint32_t code = 0;
    uint16_t k = 0;
    first_code[0] = 0, first_sym[0] = 0, count[0] = 0;
    for (int len = 1; len <= kHuffMaxLen; ++len) {
      code <<= 1;
      first_code[len] = code;
      first_sym[len] = k;
      uint16_t n = 0;
      for (int sym = 0; sym < 257; ++sym)
        if (lengths[sym] == len) {
          syms[k++] = (uint16_t)sym;
          code_of[sym] = code + n;
          len_of[sym] = (uint8_t)len;
          ++n;
        }
      count[len] = n;
      code += n;
    }
  }
};
inline const Huffman &huffman() {
  static const Huffman h;
  return h;
}

// false: padding that is not the leading bits of EOS or is 8 bits or longer, or an EOS inside the
// string (RFC 7541 5.2: both are decoding errors)
inline bool huffman_decode(const uint8_t *p, size_t n, std::string &out) {
  const Huffman &h = huffman();
  uint32_t code = 0;
  int len = 0;
  for (size_t i = 0; i < n; ++i) {
    const uint8_t b = p[i];
    for (int bit = 7; bit >= 0; --bit) {
      code = (code << 1) | ((b >> bit) & 1u);
      ++len;
      if (len >= 5) {
        const uint32_t off = code - h.first_code[len];
        if (code >= h.first_code[len] && off < h.count[len]) {
          const uint16_t sym = h.syms[h.first_sym[len] + off];
          if (sym == kHuffEos) return false;
          out.push_back((char)(uint8_t)sym);
          code = 0;
          len = 0;
        } else if (len == kHuffMaxLen) {
          return false;  // unreachable for a complete code; kept as a bound
        }
      }
    }
  }
  return len <= 7 && code == ((1u << len) - 1u);
}

// Huffman-encode (the library's own encoder never does; tests and tools use it)
inline void huffman_encode(const uint8_t *p, size_t n, std::vector<uint8_t> &out) {
  const Huffman &h = huffman();
  uint64_t acc = 0;
  int bits = 0;
  for (size_t i = 0; i < n; ++i) {
    acc = (acc << h.len_of[p[i]]) | h.code_of[p[i]];
    bits += h.len_of[p[i]];
    while (bits >= 8) {
      out.push_back((uint8_t)(acc >> (bits - 8)));
      bits -= 8;
    }
  }
  if (bits) out.push_back((uint8_t)((acc << (8 - bits)) | ((1u << (8 - bits)) - 1u)));
}

// ---- integers (RFC 7541 5.1) ---------------------------------------------------------------
inline bool decode_int(const uint8_t *&p, const uint8_t *end, int prefix_bits, uint64_t &v) {
  if (p >= end) return false;
  const uint32_t mask = (1u << prefix_bits) - 1u;
  v = *p++ & mask;
  if (v < mask) return true;
  int shift = 0;
  for (;;) {
    if (p >= end || shift > 28) return false;
    const uint8_t b = *p++;
    v += (uint64_t)(b & 0x7f) << shift;
    shift += 7;
    if (!(b & 0x80)) return true;
  }
}

inline void encode_int(std::vector<uint8_t> &out, uint8_t first_byte_flags, int prefix_bits,
                       uint64_t v) {
  const uint32_t mask = (1u << prefix_bits) - 1u;
  if (v < mask) {
    out.push_back((uint8_t)(first_byte_flags | v));
    return;
  }
  out.push_back((uint8_t)(first_byte_flags | mask));
  v -= mask;
  while (v >= 128) {
    out.push_back((uint8_t)(v % 128 + 128));
    v /= 128;
  }
  out.push_back((uint8_t)v);
}

// ---- decoder with its dynamic table (RFC 7541 sections 2.3, 4, 6) ---------------------------
struct Header {
  std::string name, value;
};

class Decoder {
 public:
  // SETTINGS_HEADER_TABLE_SIZE we advertise (we never change the default)
  explicit Decoder(uint32_t max_table = 4096, size_t max_list_bytes = 64 << 10)
      : limit_(max_table), max_size_(max_table), max_list_(max_list_bytes) {}

  // Decodes one complete header block, appending to out. false = COMPRESSION_ERROR.
  bool decode(const uint8_t *p, size_t n, std::vector<Header> &out) {
    const uint8_t *end = p + n;
    size_t list_bytes = 0;
    bool fields_seen = false;
    while (p < end) {
      const uint8_t b = *p;
      if (b & 0x80) {  // 6.1 indexed
        uint64_t idx;
        if (!decode_int(p, end, 7, idx) || idx == 0) return false;
        Header h;
        if (!lookup(idx, h, true)) return false;
        list_bytes += h.name.size() + h.value.size() + 32;
        out.push_back(std::move(h));
        fields_seen = true;
      } else if ((b & 0xe0) == 0x20) {  // 6.3 dynamic table size update
        if (fields_seen) return false;  // only at the start of a block (4.2)
        uint64_t sz;
        if (!decode_int(p, end, 5, sz) || sz > limit_) return false;
        max_size_ = (uint32_t)sz;
        evict();
      } else {
        const bool index_it = (b & 0xc0) == 0x40;  // 6.2.1; else 6.2.2 / 6.2.3
        uint64_t idx;
        if (!decode_int(p, end, index_it ? 6 : 4, idx)) return false;
        Header h;
        if (idx) {
          if (!lookup(idx, h, false)) return false;
        } else if (!decode_string(p, end, h.name)) {
          return false;
        }
        if (!decode_string(p, end, h.value)) return false;
        list_bytes += h.name.size() + h.value.size() + 32;
        if (index_it) insert(h);
        out.push_back(std::move(h));
        fields_seen = true;
      }
      if (list_bytes > max_list_) return false;
    }
    return true;
  }

  size_t dynamic_entries() const { return dyn_.size(); }
  uint32_t dynamic_size() const { return size_; }

 private:
  bool lookup(uint64_t idx, Header &h, bool with_value) const {
    if (idx <= kStaticCount) {
      const StaticEntry &e = static_table()[idx];
      h.name = e.name;
      if (with_value) h.value = e.value;
      return true;
    }
    const uint64_t d = idx - kStaticCount - 1;
    if (d >= dyn_.size()) return false;
    h.name = dyn_[d].name;
    if (with_value) h.value = dyn_[d].value;
    return true;
  }
  bool decode_string(const uint8_t *&p, const uint8_t *end, std::string &s) const {
    if (p >= end) return false;
    const bool huff = (*p & 0x80) != 0;
    uint64_t len;
    if (!decode_int(p, end, 7, len) || len > (uint64_t)(end - p) || len > max_list_) return false;
    s.clear();
    if (huff) {
      if (!huffman_decode(p, (size_t)len, s)) return false;
    } else {
      s.assign((const char *)p, (size_t)len);
    }
    p += len;
    return true;
  }
  void insert(const Header &h) {
    const uint32_t sz = (uint32_t)(h.name.size() + h.value.size() + 32);
    if (sz > max_size_) {  // 4.4: an entry larger than the table empties it
      dyn_.clear();
      size_ = 0;
      return;
    }
    dyn_.push_front(h);
    size_ += sz;
    evict();
  }
  void evict() {
    while (size_ > max_size_ && !dyn_.empty()) {
      size_ -= (uint32_t)(dyn_.back().name.size() + dyn_.back().value.size() + 32);
      dyn_.pop_back();
    }
  }
  std::deque<Header> dyn_;
  uint32_t size_ = 0;
  uint32_t limit_, max_size_;
  size_t max_list_;
};

// ---- encoder: stateless ------------------------------------------------------------------
inline void encode_string(std::vector<uint8_t> &out, const char *s, size_t n) {
  encode_int(out, 0x00, 7, n);  // H = 0
  out.insert(out.end(), (const uint8_t *)s, (const uint8_t *)s + n);
}
inline void encode_indexed(std::vector<uint8_t> &out, uint32_t static_index) {
  encode_int(out, 0x80, 7, static_index);
}
// literal without indexing (6.2.2), name from the static table
inline void encode_literal_static_name(std::vector<uint8_t> &out, uint32_t name_index,
                                       const char *v, size_t vn) {
  encode_int(out, 0x00, 4, name_index);
  encode_string(out, v, vn);
}
inline void encode_literal(std::vector<uint8_t> &out, const char *name, size_t nn, const char *v,
                           size_t vn) {
  out.push_back(0x00);
  encode_string(out, name, nn);
  encode_string(out, v, vn);
}

}  // namespace gs_h2

// hpack_kat.cpp -- known-answer tests for csrc/hpack.h from RFC 7541 Appendix C (the
// published examples of the algorithm the reference's grpc-go dependency implements).
// Built and run by tests/test_native_transport.py; prints one line per check, exits 1 on
// the first failure.
#include <stdio.h>
#include <stdlib.h>

#include "hpack.h"

using namespace gs_h2;

static std::vector<uint8_t> hex(const char *s) {
  std::vector<uint8_t> v;
  unsigned hi = 0;
  int have = 0;
  for (; *s; ++s) {
    int d;
    if (*s >= '0' && *s <= '9') d = *s - '0';
    else if (*s >= 'a' && *s <= 'f') d = *s - 'a' + 10;
    else continue;
    if (have) v.push_back((uint8_t)(hi << 4 | d)), have = 0;
    else hi = (unsigned)d, have = 1;
  }
  return v;
}

static int failures = 0;
static void expect(Decoder &d, const char *what, const char *block, const char *want,
                   size_t want_entries, uint32_t want_size) {
  std::vector<uint8_t> b = hex(block);
  std::vector<Header> hs;
  std::string got;
  if (!d.decode(b.data(), b.size(), hs)) got = "<compression error>";
  for (auto &h : hs) got += h.name + ": " + h.value + "\n";
  const bool ok = got == want && d.dynamic_entries() == want_entries && d.dynamic_size() == want_size;
  printf("%s %s (table %zu entries, %u bytes)\n", ok ? "ok  " : "FAIL", what, d.dynamic_entries(),
         d.dynamic_size());
  if (!ok) {
    printf("--- got\n%s--- want\n%s", got.c_str(), want);
    ++failures;
  }
}

int main() {
  {  // C.1 integer examples
    std::vector<uint8_t> o;
    encode_int(o, 0, 5, 10);
    encode_int(o, 0, 5, 1337);
    encode_int(o, 0, 8 - 0, 42);
    const uint8_t want[] = {0x0a, 0x1f, 0x9a, 0x0a, 0x2a};
    bool ok = o.size() == 5 && !memcmp(o.data(), want, 5);
    const uint8_t *p = o.data();
    uint64_t a, b;
    ok = ok && decode_int(p, o.data() + o.size(), 5, a) && a == 10;
    ok = ok && decode_int(p, o.data() + o.size(), 5, b) && b == 1337;
    printf("%s C.1 integers\n", ok ? "ok  " : "FAIL");
    failures += !ok;
  }
  {  // C.2.1-C.2.4 single representations
    Decoder d;
    expect(d, "C.2.1 literal with indexing",
           "400a 6375 7374 6f6d 2d6b 6579 0d63 7573 746f 6d2d 6865 6164 6572",
           "custom-key: custom-header\n", 1, 55);
    Decoder e;
    expect(e, "C.2.2 literal without indexing", "040c 2f73 616d 706c 652f 7061 7468",
           ":path: /sample/path\n", 0, 0);
    expect(e, "C.2.3 never indexed", "1008 7061 7373 776f 7264 0673 6563 7265 74",
           "password: secret\n", 0, 0);
    expect(e, "C.2.4 indexed", "82", ":method: GET\n", 0, 0);
  }
  {  // C.3 requests without Huffman, one connection
    Decoder d;
    expect(d, "C.3.1", "8286 8441 0f77 7777 2e65 7861 6d70 6c65 2e63 6f6d",
           ":method: GET\n:scheme: http\n:path: /\n:authority: www.example.com\n", 1, 57);
    expect(d, "C.3.2", "8286 84be 5808 6e6f 2d63 6163 6865",
           ":method: GET\n:scheme: http\n:path: /\n:authority: www.example.com\ncache-control: no-cache\n",
           2, 110);
    expect(d, "C.3.3", "8287 85bf 400a 6375 7374 6f6d 2d6b 6579 0c63 7573 746f 6d2d 7661 6c75 65",
           ":method: GET\n:scheme: https\n:path: /index.html\n:authority: www.example.com\n"
           "custom-key: custom-value\n",
           3, 164);
  }
  {  // C.4 the same requests with Huffman strings
    Decoder d;
    expect(d, "C.4.1", "8286 8441 8cf1 e3c2 e5f2 3a6b a0ab 90f4 ff",
           ":method: GET\n:scheme: http\n:path: /\n:authority: www.example.com\n", 1, 57);
    expect(d, "C.4.2", "8286 84be 5886 a8eb 1064 9cbf",
           ":method: GET\n:scheme: http\n:path: /\n:authority: www.example.com\ncache-control: no-cache\n",
           2, 110);
    expect(d, "C.4.3", "8287 85bf 4088 25a8 49e9 5ba9 7d7f 8925 a849 e95b b8e8 b4bf",
           ":method: GET\n:scheme: https\n:path: /index.html\n:authority: www.example.com\n"
           "custom-key: custom-value\n",
           3, 164);
  }
  {  // C.6 responses with Huffman and a 256-byte table: evictions
    Decoder d(256);
    expect(d, "C.6.1",
           "4882 6402 5885 aec3 771a 4b61 96d0 7abe 9410 54d4 44a8 2005 9504 0b81 66e0 82a6 2d1b ff6e "
           "919d 29ad 1718 63c7 8f0b 97c8 e9ae 82ae 43d3",
           ":status: 302\ncache-control: private\ndate: Mon, 21 Oct 2013 20:13:21 GMT\n"
           "location: https://www.example.com\n",
           4, 222);
    expect(d, "C.6.2", "4883 640e ffc1 c0bf",
           ":status: 307\ncache-control: private\ndate: Mon, 21 Oct 2013 20:13:21 GMT\n"
           "location: https://www.example.com\n",
           4, 222);
  }
  {  // Appendix B as a whole: the lengths form a complete prefix code, EOS is thirty 1-bits, the
     // codes the RFC prints for a few symbols come out of the canonical reconstruction
    const Huffman &h = huffman();
    unsigned long long kraft = 0;  // in units of 2^-30
    for (int s = 0; s < 257; ++s) kraft += 1ull << (30 - h.len_of[s]);
    bool ok = kraft == (1ull << 30) && h.len_of[256] == 30 && h.code_of[256] == 0x3fffffffu;
    ok = ok && h.len_of['0'] == 5 && h.code_of['0'] == 0x0 && h.len_of['a'] == 5 && h.code_of['a'] == 0x3;
    ok = ok && h.len_of[' '] == 6 && h.code_of[' '] == 0x14 && h.len_of[0] == 13 && h.code_of[0] == 0x1ff8;
    ok = ok && h.len_of[1] == 23 && h.code_of[1] == 0x7fffd8 && h.len_of[255] == 26 && h.code_of[255] == 0x3ffffee;
    ok = ok && h.len_of[10] == 30 && h.code_of[10] == 0x3ffffffc && h.len_of[22] == 30 && h.code_of[22] == 0x3ffffffe;
    printf("%s Appendix B: complete code, EOS, spot codes\n", ok ? "ok  " : "FAIL");
    failures += !ok;
  }
  {  // every byte value, Huffman-coded in one literal: a header block a peer may send
    std::vector<uint8_t> all(256), enc, block;
    for (int i = 0; i < 256; ++i) all[i] = (uint8_t)((i * 7 + 3) & 0xff);  // a permutation of 0..255
    huffman_encode(all.data(), all.size(), enc);
    block.push_back(0x00);  // literal without indexing, new name
    encode_int(block, 0x00, 7, 5);
    block.insert(block.end(), {'x', '-', 'b', 'i', 'n'});
    encode_int(block, 0x80, 7, enc.size());  // H = 1
    block.insert(block.end(), enc.begin(), enc.end());
    Decoder d;
    std::vector<Header> hs;
    bool ok = d.decode(block.data(), block.size(), hs) && hs.size() == 1 && hs[0].name == "x-bin" &&
              hs[0].value.size() == 256 && !memcmp(hs[0].value.data(), all.data(), 256);
    printf("%s all 256 byte values Huffman-coded (%zu bytes)\n", ok ? "ok  " : "FAIL", enc.size());
    failures += !ok;
    printf("block ");
    for (uint8_t b : block) printf("%02x", b);
    printf("\n");
    // an EOS inside a string and 8 bits of padding are decoding errors (RFC 7541 5.2)
    const uint8_t eos_inside[] = {0x00, 0x01, 'x', 0x85, 0xff, 0xff, 0xff, 0xff, 0xff};   // 30 ones + more
    const uint8_t long_pad[] = {0x00, 0x01, 'x', 0x82, 0x07, 0xff};                        // '0' + 3 pad, then 8 more
    ok = !d.decode(eos_inside, sizeof eos_inside, hs) && !d.decode(long_pad, sizeof long_pad, hs);
    printf("%s EOS inside a string / over-long padding rejected\n", ok ? "ok  " : "FAIL");
    failures += !ok;
  }
  {  // malformed input never reads out of bounds and is reported
    Decoder d;
    std::vector<Header> hs;
    const uint8_t bad1[] = {0x80};                    // index 0
    const uint8_t bad2[] = {0xff, 0xff, 0xff};        // integer runs off the block
    const uint8_t bad3[] = {0x40, 0x05, 'a'};         // string longer than the block
    const uint8_t bad4[] = {0x00, 0x81, 0x00, 0x00};  // Huffman padding that is not EOS bits
    const uint8_t bad5[] = {0xc0};                    // index past both tables
    const uint8_t bad6[] = {0x3f, 0xe1, 0x7f};        // table size update above the limit
    bool ok = !d.decode(bad1, 1, hs) && !d.decode(bad2, 3, hs) && !d.decode(bad3, 3, hs) &&
              !d.decode(bad4, 4, hs) && !d.decode(bad5, 1, hs) && !d.decode(bad6, 3, hs);
    printf("%s malformed blocks rejected\n", ok ? "ok  " : "FAIL");
    failures += !ok;
  }
  {  // encoder output decodes to what was meant
    std::vector<uint8_t> o;
    encode_indexed(o, 8);
    encode_literal_static_name(o, 31, "application/grpc", 16);
    encode_literal(o, "grpc-status", 11, "0", 1);
    Decoder d;
    std::vector<Header> hs;
    bool ok = d.decode(o.data(), o.size(), hs) && hs.size() == 3 && hs[0].name == ":status" &&
              hs[0].value == "200" && hs[1].name == "content-type" && hs[1].value == "application/grpc" &&
              hs[2].name == "grpc-status" && hs[2].value == "0" && d.dynamic_entries() == 0;
    printf("%s encoder round trip\n", ok ? "ok  " : "FAIL");
    failures += !ok;
  }
  return failures ? 1 : 0;
}

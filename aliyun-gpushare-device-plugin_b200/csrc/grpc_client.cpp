// grpc_client.cpp -- a small gRPC client over a unix socket, behind the C ABI.
//
// Replaces, in the reference's pkg/gpu/nvidia/server.go:
//   dial      252-265  grpc.Dial(unix socket, insecure, blocking, timeout)
//   Register  158-177  Registration/Register{version, endpoint, resource_name} to kubelet
// and gives tests and benchmarks a kubelet-side caller (unary Allocate, the ListAndWatch
// stream) that costs no more than the server it measures.
//
// One caller at a time per client (no internal locking). A unary call and one or more open
// server streams may share the connection: frames for streams other than the one being
// waited on are filed under their stream.
#include <poll.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include <memory>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "gpushare.h"
#include "h2conn.h"
#include "hpack.h"
#include "internal.h"

using namespace gs_h2;

namespace {

struct RBuf {  // received message bytes of one stream; never zero-filled
  uint8_t *p = nullptr;
  size_t n = 0, cap = 0, rpos = 0;
  ~RBuf() { free(p); }
  bool reserve(size_t want) {
    if (want <= cap) return true;
    size_t c = cap ? cap : 4096;
    while (c < want) c *= 2;
    uint8_t *q = (uint8_t *)realloc(p, c);
    if (!q) return false;
    p = q;
    cap = c;
    return true;
  }
  void compact() {
    if (rpos == n) rpos = n = 0;
  }
  void swap(RBuf &o) {
    std::swap(p, o.p);
    std::swap(n, o.n);
    std::swap(cap, o.cap);
    std::swap(rpos, o.rpos);
  }
};

struct CStream {
  uint32_t id = 0;
  RBuf data;
  bool closed = false;       // END_STREAM or RST_STREAM seen
  int grpc_status = -1;      // from the trailers
  int http_status = 0;
  std::string grpc_message;
  int64_t send_window = kDefaultWindow;
  uint32_t recv_unacked = 0;
};

int hexval(char ch) {
  if (ch >= '0' && ch <= '9') return ch - '0';
  if (ch >= 'a' && ch <= 'f') return ch - 'a' + 10;
  if (ch >= 'A' && ch <= 'F') return ch - 'A' + 10;
  return -1;
}
std::string percent_decode(const std::string &s) {
  std::string o;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] == '%' && i + 2 < s.size() && hexval(s[i + 1]) >= 0 && hexval(s[i + 2]) >= 0) {
      o.push_back((char)(hexval(s[i + 1]) * 16 + hexval(s[i + 2])));
      i += 2;
    } else {
      o.push_back(s[i]);
    }
  }
  return o;
}

int64_t now_ms() {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (int64_t)ts.tv_sec * 1000 + ts.tv_nsec / 1000000;
}

}  // namespace

struct gs_grpc_client {
  Wire w;
  Decoder hpack;
  bool settings_seen = false;
  bool broken = false;
  bool goaway = false;
  uint32_t goaway_last = 0x7fffffffu;
  uint32_t next_id = 1;
  int64_t send_window = kDefaultWindow;
  uint32_t peer_initial_window = kDefaultWindow;
  uint32_t peer_max_frame = kDefaultMaxFrame;
  uint64_t recv_unacked = 0;
  std::unordered_map<uint32_t, std::unique_ptr<CStream>> streams;
  std::vector<uint8_t> header_block;
  uint32_t cont_stream = 0;
  uint8_t cont_flags = 0;
  std::vector<Header> headers;
  std::string last_error;
  RBuf spare;  // the largest message buffer of a finished stream, handed to the next one:
               // a 90 MB list does not fault its pages in again on every stream

  void drop_stream(uint32_t id) {
    auto it = streams.find(id);
    if (it == streams.end()) return;
    RBuf &d = it->second->data;
    if (d.cap > spare.cap) {
      d.n = d.rpos = 0;
      spare.swap(d);
    }
    streams.erase(it);
  }

  bool fail(const char *what) {
    if (!broken) last_error = what;
    broken = true;
    return false;
  }

  // waits until the socket is readable or buffered bytes exist; false = timeout
  bool wait_readable(int64_t deadline_ms) {
    if (w.buffered()) return true;
    for (;;) {
      const int64_t left = deadline_ms - now_ms();
      if (left <= 0) return false;
      struct pollfd p = {w.fd, POLLIN, 0};
      const int r = poll(&p, 1, (int)(left > 1000000 ? 1000000 : left));
      if (r > 0) return true;
      if (r < 0 && errno != EINTR) return false;
    }
  }

  bool headers_complete(uint32_t id, uint8_t flags) {
    headers.clear();
    if (!hpack.decode(header_block.data(), header_block.size(), headers))
      return fail("HPACK: compression error in the peer's header block");
    auto it = streams.find(id);
    if (it == streams.end()) return true;
    CStream &s = *it->second;
    for (const Header &h : headers) {
      if (h.name == ":status") s.http_status = atoi(h.value.c_str());
      else if (h.name == "grpc-status") s.grpc_status = atoi(h.value.c_str());
      else if (h.name == "grpc-message") s.grpc_message = percent_decode(h.value);
    }
    if (flags & kFlagEndStream) s.closed = true;
    return true;
  }

  // reads and files exactly one frame
  bool pump_one() {
    FrameHeader h;
    if (!w.read_frame_header(h)) return fail("connection closed or timed out");
    if (cont_stream && h.type != kContinuation) return fail("protocol: interleaved header block");
    switch (h.type) {
      case kData: {
        size_t len = h.length, pad = 0;
        recv_unacked += h.length;
        if (h.flags & kFlagPadded) {
          if (len < 1 || !w.fill(1)) return fail("short DATA frame");
          pad = w.peek()[0];
          w.consume(1);
          --len;
          if (pad > len) return fail("protocol: padding longer than the frame");
          len -= pad;
        }
        auto it = streams.find(h.stream);
        if (it == streams.end() || it->second->closed) {
          if (!w.skip(len + pad)) return fail("connection closed inside a DATA frame");
        } else {
          CStream &s = *it->second;
          RBuf &d = s.data;
          d.compact();
          size_t want = d.n + len;
          if (d.n - d.rpos >= 5) {  // the message length is known: size the buffer once
            const size_t whole = d.rpos + 5 + be32(d.p + d.rpos + 1);
            if (whole > want && whole <= ((size_t)1 << 31)) want = whole;
          } else if (d.n == d.rpos && len >= 5 && w.buffered() >= 5) {
            const size_t whole = d.n + 5 + be32(w.peek() + 1);
            if (whole > want && whole <= ((size_t)1 << 31)) want = whole;
          }
          if (!d.reserve(want)) return fail("out of memory");
          if (!w.read_into(d.p + d.n, len) || !w.skip(pad))
            return fail("connection closed inside a DATA frame");
          d.n += len;
          s.recv_unacked += h.length;
          if (h.flags & kFlagEndStream) {
            s.closed = true;
          } else if (s.recv_unacked >= kOurStreamWindow / 2) {
            if (!w.send_window_update(s.id, s.recv_unacked)) return fail("send failed");
            s.recv_unacked = 0;
          }
        }
        if (recv_unacked >= (256u << 20)) {
          if (!w.send_window_update(0, (uint32_t)recv_unacked)) return fail("send failed");
          recv_unacked = 0;
        }
        return true;
      }
      case kHeaders:
      case kContinuation: {
        if (h.stream == 0 || h.length > (64u << 10) || header_block.size() + h.length > (256u << 10))
          return fail("protocol: bad header frame");
        if (!w.fill(h.length)) return fail("connection closed inside a header frame");
        const uint8_t *p = w.peek();
        size_t n = h.length;
        if (h.type == kHeaders) {
          if (!strip_headers_prefix(h.flags, p, n)) return fail("protocol: bad HEADERS padding");
          header_block.assign(p, p + n);
          cont_flags = h.flags;
        } else {
          if (cont_stream != h.stream) return fail("protocol: stray CONTINUATION");
          header_block.insert(header_block.end(), p, p + n);
        }
        w.consume(h.length);
        if (h.flags & kFlagEndHeaders) {
          cont_stream = 0;
          const bool ok = headers_complete(h.stream, cont_flags);
          header_block.clear();
          return ok;
        }
        cont_stream = h.stream;
        return true;
      }
      case kSettings: {
        if (h.stream != 0 || h.length > 1024 || (h.flags & kFlagAck ? h.length != 0 : h.length % 6 != 0))
          return fail("protocol: bad SETTINGS frame");
        if (h.flags & kFlagAck) return true;
        if (!w.fill(h.length)) return fail("connection closed inside SETTINGS");
        const uint8_t *p = w.peek();
        for (uint32_t i = 0; i + 6 <= h.length; i += 6) {
          const uint16_t id = (uint16_t)(p[i] << 8 | p[i + 1]);
          const uint32_t v = be32(p + i + 2);
          if (id == kSetInitialWindowSize && v <= 0x7fffffffu) {
            const int64_t delta = (int64_t)v - (int64_t)peer_initial_window;
            peer_initial_window = v;
            for (auto &kv : streams) kv.second->send_window += delta;
          } else if (id == kSetMaxFrameSize && v >= kDefaultMaxFrame && v <= kOurMaxFrame) {
            peer_max_frame = v;
          }
        }
        w.consume(h.length);
        settings_seen = true;
        return w.send_settings_ack() ? true : fail("send failed");
      }
      case kWindowUpdate: {
        if (h.length != 4 || !w.fill(4)) return fail("protocol: bad WINDOW_UPDATE");
        const uint32_t inc = be32(w.peek()) & 0x7fffffffu;
        w.consume(4);
        if (h.stream == 0) {
          send_window += inc;
        } else {
          auto it = streams.find(h.stream);
          if (it != streams.end()) it->second->send_window += inc;
        }
        return true;
      }
      case kPing: {
        if (h.length != 8 || !w.fill(8)) return fail("protocol: bad PING");
        uint8_t opaque[8];
        memcpy(opaque, w.peek(), 8);
        w.consume(8);
        if (h.flags & kFlagAck) return true;
        return w.send_ping_ack(opaque) ? true : fail("send failed");
      }
      case kRstStream: {
        if (h.length != 4 || !w.fill(4)) return fail("protocol: bad RST_STREAM");
        const uint32_t code = be32(w.peek());
        w.consume(4);
        auto it = streams.find(h.stream);
        if (it != streams.end() && !it->second->closed) {
          CStream &s = *it->second;
          s.closed = true;
          if (s.grpc_status < 0) {
            s.grpc_status = code == kErrCancel ? 1 /*CANCELLED*/ : 13 /*INTERNAL*/;
            char m[64];
            snprintf(m, sizeof(m), "stream reset by the peer (HTTP/2 error %u)", code);
            s.grpc_message = m;
          }
        }
        return true;
      }
      case kGoaway: {
        if (h.length < 8 || !w.fill(8)) return fail("protocol: bad GOAWAY");
        goaway = true;
        goaway_last = be32(w.peek()) & 0x7fffffffu;
        w.consume(8);
        if (!w.skip(h.length - 8)) return fail("connection closed inside GOAWAY");
        for (auto &kv : streams) {
          if (kv.first > goaway_last && !kv.second->closed) {
            kv.second->closed = true;
            kv.second->grpc_status = 14;  // UNAVAILABLE
            kv.second->grpc_message = "the server is going away";
          }
        }
        return true;
      }
      default:
        return w.skip(h.length) ? true : fail("connection closed");
    }
  }

  // HEADERS + the one request message + END_STREAM; reads frames while flow control blocks
  bool send_request(CStream &s, const char *path, const uint8_t *req, size_t n) {
    std::vector<uint8_t> block;
    encode_indexed(block, 3);  // :method POST
    encode_indexed(block, 6);  // :scheme http
    encode_literal_static_name(block, 4, path, strlen(path));
    encode_literal_static_name(block, 1, "localhost", 9);
    encode_literal_static_name(block, 31, "application/grpc", 16);
    encode_literal(block, "te", 2, "trailers", 8);
    encode_literal_static_name(block, 58, "gpushare-b200", 13);
    uint8_t hf[9], prefix[5];
    put_frame_header(hf, (uint32_t)block.size(), kHeaders, kFlagEndHeaders, s.id);
    prefix[0] = 0;
    put_be32(prefix + 1, (uint32_t)n);
    gs_iovec segs[2] = {{prefix, 5}, {req, n}};
    size_t seg = 0, off = 0;
    uint64_t remaining = 5 + n;
    bool headers_out = false;
    while (remaining) {
      int64_t allow = (int64_t)remaining;
      if (allow > send_window) allow = send_window;
      if (allow > s.send_window) allow = s.send_window;
      if (allow > (int64_t)peer_max_frame) allow = peer_max_frame;
      if (allow <= 0) {
        if (!headers_out) {
          if (!w.write_all(hf, 9) || !w.write_all(block.data(), block.size())) return fail("send failed");
          headers_out = true;
        }
        if (!pump_one()) return false;  // wait for WINDOW_UPDATE / SETTINGS
        if (s.closed) return true;      // the server answered early (e.g. message too large)
        continue;
      }
      struct iovec iov[6];
      int k = 0;
      uint8_t df[9];
      if (!headers_out) {
        iov[k++] = {hf, 9};
        iov[k++] = {block.data(), block.size()};
        headers_out = true;
      }
      put_frame_header(df, (uint32_t)allow, kData, remaining == (uint64_t)allow ? kFlagEndStream : 0, s.id);
      iov[k++] = {df, 9};
      uint64_t need = (uint64_t)allow;
      while (need) {
        const gs_iovec &g = segs[seg];
        const size_t take = g.len - off < need ? g.len - off : (size_t)need;
        if (take) iov[k++] = {(void *)(g.base + off), take};
        off += take;
        need -= take;
        if (off == g.len) {
          ++seg;
          off = 0;
        }
      }
      if (!w.writev_all(iov, k)) return fail("send failed");
      remaining -= (uint64_t)allow;
      send_window -= allow;
      s.send_window -= allow;
    }
    return true;
  }

  CStream *open_stream() {
    if (broken || goaway) return nullptr;
    std::unique_ptr<CStream> s(new CStream());
    s->id = next_id;
    next_id += 2;
    s->send_window = peer_initial_window;
    if (spare.cap) s->data.swap(spare);
    CStream *raw = s.get();
    streams[raw->id] = std::move(s);
    return raw;
  }
};

namespace {

thread_local std::string g_last_error;

// Copies the next complete message of the stream to the caller. GS_OK, GS_E_NOSPACE (the
// message stays queued, *len = its size), or 1 = no complete message buffered yet.
int take_message(CStream &s, uint8_t *out, size_t cap, size_t *len, const uint8_t **borrow = nullptr) {
  RBuf &d = s.data;
  d.compact();
  if (d.n - d.rpos < 5) return 1;
  if (d.p[d.rpos] & 1) return GS_E_WIRE;  // compressed message: never negotiated
  const size_t mlen = be32(d.p + d.rpos + 1);
  if (d.n - d.rpos - 5 < mlen) return 1;
  *len = mlen;
  if (borrow) {  // the bytes stay where they were received until the next call on the stream
    *borrow = d.p + d.rpos + 5;
    d.rpos += 5 + mlen;
    return GS_OK;
  }
  if (mlen > cap || (!out && mlen)) return GS_E_NOSPACE;
  if (mlen) memcpy(out, d.p + d.rpos + 5, mlen);
  d.rpos += 5 + mlen;
  d.compact();
  return GS_OK;
}

void write_message(const std::string &m, char *buf, size_t cap) {
  if (!buf || !cap) return;
  const size_t n = m.size() < cap - 1 ? m.size() : cap - 1;
  memcpy(buf, m.data(), n);
  buf[n] = 0;
}

void put_string_field(std::vector<uint8_t> &o, uint32_t field, const char *s) {
  const size_t n = strlen(s);
  o.push_back((uint8_t)(field << 3 | 2));
  size_t v = n;
  while (v >= 0x80) {
    o.push_back((uint8_t)(v | 0x80));
    v >>= 7;
  }
  o.push_back((uint8_t)v);
  o.insert(o.end(), s, s + n);
}

}  // namespace

extern "C" {

const char *gs_grpc_last_error(void) { return g_last_error.c_str(); }

int gs_grpc_client_connect(const char *unix_path, uint32_t timeout_ms, gs_grpc_client **out) {
  if (!out) return GS_E_INVAL;
  *out = nullptr;
  struct sockaddr_un sa;
  socklen_t salen;
  if (!unix_path || !fill_sockaddr(unix_path, sa, salen)) return GS_E_INVAL;
  const int64_t deadline = now_ms() + (timeout_ms ? timeout_ms : 5000);
  int fd = -1;
  // grpc.Dial(..., WithBlock, WithTimeout): keep trying until the socket accepts
  for (;;) {
    fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) return GS_E_STATE;
    if (connect(fd, (struct sockaddr *)&sa, salen) == 0) break;
    const int e = errno;
    close(fd);
    fd = -1;
    if (now_ms() >= deadline) {
      g_last_error = std::string("dial ") + unix_path + ": " + strerror(e);
      return GS_E_TIMEOUT;
    }
    struct timespec ts = {0, 20 * 1000 * 1000};
    nanosleep(&ts, nullptr);
  }
  std::unique_ptr<gs_grpc_client> c(new (std::nothrow) gs_grpc_client());
  if (!c) {
    close(fd);
    return GS_E_INVAL;
  }
  c->w.fd = fd;
  set_io_timeouts(fd, 30);
  set_big_buffers(fd, 4 << 20);
  const uint16_t ids[3] = {kSetEnablePush, kSetInitialWindowSize, kSetMaxFrameSize};
  const uint32_t vals[3] = {0, kOurStreamWindow, kOurMaxFrame};
  bool ok = c->w.write_all((const uint8_t *)kPreface, 24) && c->w.send_settings(ids, vals, 3) &&
            c->w.send_window_update(0, kOurConnWindow - kDefaultWindow);
  while (ok && !c->settings_seen) {  // the server's preface: its SETTINGS frame
    if (!c->wait_readable(deadline)) {
      c->fail("no SETTINGS from the server before the deadline");
      ok = false;
      break;
    }
    ok = c->pump_one();
  }
  if (!ok) {
    g_last_error = c->last_error.empty() ? "handshake failed" : c->last_error;
    close(fd);
    return GS_E_TIMEOUT;
  }
  *out = c.release();
  return GS_OK;
}

void gs_grpc_client_close(gs_grpc_client *c) {
  if (!c) return;
  if (c->w.fd >= 0) {
    if (!c->broken) c->w.send_goaway(0, kErrNo);
    close(c->w.fd);
  }
  delete c;
}

int gs_grpc_client_unary(gs_grpc_client *c, const char *method_path, const uint8_t *req,
                         size_t req_len, uint8_t *resp, size_t resp_cap, size_t *resp_len,
                         int32_t *grpc_status, char *message, size_t message_cap) {
  if (!c || !method_path || (!req && req_len) || !resp_len || !grpc_status) return GS_E_INVAL;
  *resp_len = 0;
  *grpc_status = -1;
  if (message && message_cap) message[0] = 0;
  CStream *s = c->open_stream();
  if (!s) {
    g_last_error = c->last_error.empty() ? "connection is closed" : c->last_error;
    return GS_E_STATE;
  }
  const uint32_t id = s->id;
  bool ok = c->send_request(*s, method_path, req, req_len);
  while (ok && !s->closed) ok = c->pump_one();
  int rc = GS_OK;
  if (!ok) {
    g_last_error = c->last_error;
    rc = GS_E_STATE;
  } else {
    // grpc-status travels in the trailers; without one the call failed in transit
    *grpc_status = s->grpc_status >= 0 ? s->grpc_status : (s->http_status == 200 ? 2 : 14);
    write_message(s->grpc_message, message, message_cap);
    if (*grpc_status == 0) {
      const int t = take_message(*s, resp, resp_cap, resp_len);
      if (t == GS_E_NOSPACE) rc = GS_E_NOSPACE;
      else if (t != GS_OK) rc = GS_E_WIRE;  // OK status without a complete message
    }
  }
  c->drop_stream(id);
  return rc;
}

int gs_grpc_client_stream_open(gs_grpc_client *c, const char *method_path, const uint8_t *req,
                               size_t req_len, uint32_t *stream) {
  if (!c || !method_path || (!req && req_len) || !stream) return GS_E_INVAL;
  CStream *s = c->open_stream();
  if (!s) return GS_E_STATE;
  if (!c->send_request(*s, method_path, req, req_len)) {
    g_last_error = c->last_error;
    c->drop_stream(s->id);
    return GS_E_STATE;
  }
  *stream = s->id;
  return GS_OK;
}

static int stream_next(gs_grpc_client *c, uint32_t stream, uint8_t *resp, size_t resp_cap,
                       size_t *resp_len, int32_t *grpc_status, uint32_t timeout_ms,
                       const uint8_t **borrow) {
  if (!c || !resp_len || !grpc_status) return GS_E_INVAL;
  auto it = c->streams.find(stream);
  if (it == c->streams.end()) return GS_E_NOTFOUND;
  CStream &s = *it->second;
  *grpc_status = -1;
  const int64_t deadline = now_ms() + timeout_ms;
  for (;;) {
    const int t = take_message(s, resp, resp_cap, resp_len, borrow);
    if (t != 1) return t;
    if (s.closed) {  // no further message: the stream's status is the answer
      *grpc_status = s.grpc_status >= 0 ? s.grpc_status : 2;
      return GS_E_STATE;
    }
    if (c->broken) {
      g_last_error = c->last_error;
      return GS_E_STATE;
    }
    if (!c->wait_readable(deadline)) return GS_E_TIMEOUT;
    if (!c->pump_one()) {
      g_last_error = c->last_error;
      if (s.closed) continue;
      return GS_E_STATE;
    }
  }
}

int gs_grpc_client_stream_next(gs_grpc_client *c, uint32_t stream, uint8_t *resp, size_t resp_cap,
                               size_t *resp_len, int32_t *grpc_status, uint32_t timeout_ms) {
  return stream_next(c, stream, resp, resp_cap, resp_len, grpc_status, timeout_ms, nullptr);
}

int gs_grpc_client_stream_next_borrow(gs_grpc_client *c, uint32_t stream, const uint8_t **msg,
                                      size_t *len, int32_t *grpc_status, uint32_t timeout_ms) {
  if (!msg) return GS_E_INVAL;
  return stream_next(c, stream, nullptr, 0, len, grpc_status, timeout_ms, msg);
}

int gs_grpc_client_stream_cancel(gs_grpc_client *c, uint32_t stream) {
  if (!c) return GS_E_INVAL;
  auto it = c->streams.find(stream);
  if (it == c->streams.end()) return GS_E_NOTFOUND;
  if (!it->second->closed && !c->broken) c->w.send_rst(stream, kErrCancel);
  c->drop_stream(stream);
  return GS_OK;
}

// Register (server.go:158-177): dial kubelet's socket (5 s, blocking), send
// RegisterRequest{version: "v1beta1", endpoint, resource_name} -- options unset -- and fail
// on any non-OK status.
int gs_register(const char *kubelet_socket, const char *endpoint, const char *resource_name,
                uint32_t timeout_ms) {
  if (!kubelet_socket || !endpoint || !resource_name) return GS_E_INVAL;
  gs_grpc_client *c = nullptr;
  int rc = gs_grpc_client_connect(kubelet_socket, timeout_ms ? timeout_ms : 5000, &c);
  if (rc != GS_OK) return rc;
  std::vector<uint8_t> req;
  put_string_field(req, 1, "v1beta1");  // pluginapi.Version
  put_string_field(req, 2, endpoint);
  put_string_field(req, 3, resource_name);
  uint8_t resp[64];
  size_t rlen = 0;
  int32_t status = -1;
  char msg[256];
  rc = gs_grpc_client_unary(c, "/v1beta1.Registration/Register", req.data(), req.size(), resp,
                            sizeof(resp), &rlen, &status, msg, sizeof(msg));
  gs_grpc_client_close(c);
  if (rc == GS_E_NOSPACE) rc = GS_OK;  // an Empty reply never is; tolerate a chatty kubelet
  if (rc != GS_OK) return rc;
  if (status != 0) {
    char num[32];
    snprintf(num, sizeof(num), "%d", status);
    g_last_error = std::string("Register: grpc status ") + num + ": " + msg;
    return GS_E_STATE;
  }
  return GS_OK;
}

}  // extern "C"

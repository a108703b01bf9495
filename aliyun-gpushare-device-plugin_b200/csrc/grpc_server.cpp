// grpc_server.cpp -- the kubelet-facing gRPC server of the plugin, native, behind the C ABI.
//
// Replaces the transport half of pkg/gpu/nvidia/server.go in the reference:
//   Start   114-142  remove a stale socket, listen on the unix socket, serve, self-dial (5 s)
//   Stop    145-155  close(m.stop), stop the server, remove the socket
//   Serve   232-249  Start + Register (the Register call is grpc_client.cpp)
// and routes the five v1beta1.DevicePlugin methods (server.go:93-95,180-193,199-201,
// allocate.go:43) to the gs_plugin_* entry points, so kubelet's bytes reach the Allocate
// decision without passing through another runtime.
//
// One thread serves every connection (kubelet opens one): Allocate is serialised by the
// reference's own lock anyway (allocate.go:60), and ListAndWatch sends borrow the device
// table's pre-encoded per-GPU segments (gs_law_payload), which allows one sender. A health
// change while a list is still going out is sent as a fresh list once that one is finished.
#include <fcntl.h>
#include <poll.h>
#include <stdio.h>
#include <stdlib.h>
#include <sys/eventfd.h>
#include <sys/stat.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "gpushare.h"
#include "h2conn.h"
#include "hpack.h"
#include "internal.h"

using namespace gs_h2;

namespace {

inline uint64_t mono_ns() {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}
constexpr uint64_t kSendSlice = 4u << 20;  // bytes of a device list per turn of the serving loop
constexpr size_t kMaxStreams = 128;  // advertised in our SETTINGS and enforced
constexpr size_t kMaxConns = 64;     // kubelet opens one; a bound on descriptors and buffers
enum Method { kUnknown, kOptions, kListAndWatch, kAllocate, kPreStart, kPreferred };
const char kServicePrefix[] = "/v1beta1.DevicePlugin/";

// gRPC status codes used here (google.golang.org/grpc/codes)
enum { kOk = 0, kUnknownCode = 2, kResourceExhausted = 8, kUnimplemented = 12, kInternal = 13 };

// growable byte buffer that is never zero-filled (requests are written once by recv)
struct Buf {
  uint8_t *p = nullptr;
  size_t n = 0, cap = 0;
  ~Buf() { free(p); }
  Buf() = default;
  Buf(const Buf &) = delete;
  Buf &operator=(const Buf &) = delete;
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
};

struct Outgoing {
  bool active = false;
  std::vector<gs_iovec> segs;  // borrowed or pointing into owned
  size_t seg = 0, off = 0;
  uint64_t remaining = 0;
  uint8_t prefix[5];             // gRPC message prefix: compressed flag + be32 length
  std::vector<uint8_t> owned;    // unary response bytes
  bool borrowed = false;         // segs point into the device table
  bool trailers_after = false;   // unary: grpc-status 0 + END_STREAM once the data is out
};

struct Stream {
  uint32_t id = 0;
  Method method = kUnknown;
  std::string path;
  std::string encoding;        // grpc-encoding of the request, if any
  std::unique_ptr<Buf> req;
  bool bad_content_type = false;
  int64_t send_window = kDefaultWindow;
  uint32_t recv_unacked = 0;
  bool headers_sent = false;
  bool law = false;            // an open ListAndWatch stream
  uint64_t law_gen = 0;        // generation of the list last queued on it
  bool law_sent_once = false;
  Outgoing out;
};

struct Conn {
  uint64_t serial = 0;  // names the connection for answers that come back from the worker
  Wire w;
  Decoder hpack;
  bool preface_seen = false;
  bool dead = false;
  int64_t send_window = kDefaultWindow;
  uint32_t peer_initial_window = kDefaultWindow;
  uint32_t peer_max_frame = kDefaultMaxFrame;
  uint32_t last_stream = 0;
  uint64_t recv_unacked = 0;
  std::unordered_map<uint32_t, Stream> streams;
  std::vector<std::unique_ptr<Buf>> free_bufs;
  std::vector<uint8_t> header_block;
  uint32_t cont_stream = 0;  // stream whose header block awaits CONTINUATION frames
  uint8_t cont_flags = 0;    // flags of the HEADERS frame that opened it
  std::vector<Header> headers;
};

// An Allocate handed to the caller's callback (gs_server_config.allocate). The callback may
// spend seconds on the network (lister ladders, PATCH), so it runs on a worker thread: the
// serving thread keeps answering ListAndWatch and other connections meanwhile. One worker,
// so callbacks stay serialised exactly as m.Lock() serialises Allocate (allocate.go:60).
struct AllocJob {
  uint64_t conn_serial;
  uint32_t stream;
  std::unique_ptr<Buf> req;  // gRPC-framed request: 5-byte prefix + AllocateRequest
};
struct AllocDone {
  uint64_t conn_serial;
  uint32_t stream;
  int rc;
  std::vector<uint8_t> resp;
  std::unique_ptr<Buf> req;  // handed back for reuse
};

}  // namespace

struct gs_server {
  gs_plugin *plugin = nullptr;
  gs_table *table = nullptr;
  std::string socket_path;
  gs_server_allocate_fn allocate_fn = nullptr;
  gs_server_decision_fn decision_fn = nullptr;
  void *user = nullptr;
  size_t max_request = 64u << 20;

  int listen_fd = -1;
  int wake_fd = -1;
  std::thread thread;
  std::atomic<bool> stop_requested{false};
  std::vector<std::unique_ptr<Conn>> conns;

  // the borrowed ListAndWatch payload (one snapshot shared by every stream sending it)
  std::vector<gs_iovec> law_iov;
  uint64_t law_iov_gen = 0;
  uint64_t law_iov_bytes = 0;
  bool law_iov_valid = false;
  uint32_t law_borrowers = 0;  // streams with a borrowed send in progress

  std::vector<uint8_t> resp_headers;   // :status 200, content-type application/grpc
  std::vector<uint8_t> ok_trailers;    // grpc-status 0
  std::vector<uint8_t> resp;           // Allocate response scratch
  std::vector<struct iovec> iov_scratch;
  uint64_t next_conn_serial = 1;

  std::thread worker;
  std::mutex job_mu;
  std::condition_variable job_cv;
  std::deque<AllocJob> jobs;
  std::deque<AllocDone> done;
  bool worker_stop = false;
  void worker_main();
  void deliver_done();

  // stats (read from other threads)
  std::atomic<uint64_t> n_connections{0}, n_requests{0}, n_allocate{0}, n_law_messages{0},
      n_bytes_in{0}, n_bytes_out{0}, n_protocol_errors{0}, ns_recv_body{0}, ns_allocate{0};

  void run();
  void accept_one();
  void serve_conn(Conn &c);
  bool handle_frame(Conn &c, const FrameHeader &h);
  bool on_headers_complete(Conn &c, uint32_t stream_id, uint8_t flags);
  bool on_request_complete(Conn &c, Stream &s);
  bool flush(Conn &c, Stream &s);
  bool flush_all(Conn &c);
  void continue_sends();
  bool send_error(Conn &c, Stream &s, int code, const std::string &msg);
  bool send_unary(Conn &c, Stream &s, const uint8_t *msg, size_t n);
  void close_stream(Conn &c, uint32_t id);
  void queue_law(Conn &c, Stream &s);
  void pump_law();
  void drop_conn(Conn &c);
};

namespace {

std::string percent_encode(const std::string &s) {  // grpc-message (gRPC over HTTP/2 spec)
  static const char hexd[] = "0123456789ABCDEF";
  std::string o;
  for (unsigned char ch : s) {
    if (ch >= 0x20 && ch <= 0x7e && ch != '%') {
      o.push_back((char)ch);
    } else {
      o.push_back('%');
      o.push_back(hexd[ch >> 4]);
      o.push_back(hexd[ch & 15]);
    }
  }
  return o;
}

}  // namespace

// ---- sending ---------------------------------------------------------------------------

void gs_server::close_stream(Conn &c, uint32_t id) {
  auto it = c.streams.find(id);
  if (it == c.streams.end()) return;
  Stream &s = it->second;
  if (s.out.borrowed) {
    s.out.borrowed = false;
    --law_borrowers;
  }
  if (s.req) {
    s.req->n = 0;
    if (c.free_bufs.size() < 4) c.free_bufs.push_back(std::move(s.req));
  }
  c.streams.erase(it);
}

// Sends what the flow-control windows allow of the stream's queued message: response
// HEADERS first if not yet sent, DATA frames gathered into one sendmsg per 32 frames, the
// OK trailers when a unary message is out. false = the connection is gone. A finished
// unary stream is closed here; callers must not touch `s` after a flush of one.
bool gs_server::flush(Conn &c, Stream &s) {
  Outgoing &o = s.out;
  constexpr int kMaxFrames = 32;
  // A device list can be tens of MB: it goes out in slices, and between slices the serving
  // loop looks at its other streams and connections (continue_sends), so an Allocate never
  // waits for a whole list to leave.
  uint64_t budget = o.borrowed ? kSendSlice : ~0ull;
  for (;;) {
    std::vector<struct iovec> &iov = iov_scratch;
    iov.clear();
    uint8_t fh[kMaxFrames + 2][9];
    int nfh = 0;
    bool finished = false;
    if (!s.headers_sent) {
      put_frame_header(fh[nfh], (uint32_t)resp_headers.size(), kHeaders, kFlagEndHeaders, s.id);
      iov.push_back({fh[nfh++], 9});
      iov.push_back({resp_headers.data(), resp_headers.size()});
      s.headers_sent = true;
    }
    while (o.active && o.remaining && nfh < kMaxFrames && budget) {
      int64_t allow = (int64_t)o.remaining;
      if ((uint64_t)allow > budget) allow = (int64_t)budget;
      if (allow > c.send_window) allow = c.send_window;
      if (allow > s.send_window) allow = s.send_window;
      if (allow > (int64_t)c.peer_max_frame) allow = c.peer_max_frame;
      if (allow <= 0) break;
      put_frame_header(fh[nfh], (uint32_t)allow, kData, 0, s.id);
      iov.push_back({fh[nfh++], 9});
      uint64_t need = (uint64_t)allow;
      while (need) {
        const gs_iovec &g = o.segs[o.seg];
        const size_t k = g.len - o.off < need ? g.len - o.off : (size_t)need;
        if (k) iov.push_back({(void *)(g.base + o.off), k});
        o.off += k;
        need -= k;
        if (o.off == g.len) {
          ++o.seg;
          o.off = 0;
        }
      }
      o.remaining -= (uint64_t)allow;
      budget -= (uint64_t)allow;
      c.send_window -= allow;
      s.send_window -= allow;
    }
    if (o.active && o.remaining == 0) {
      o.active = false;
      if (o.trailers_after) {
        put_frame_header(fh[nfh], (uint32_t)ok_trailers.size(), kHeaders,
                         kFlagEndHeaders | kFlagEndStream, s.id);
        iov.push_back({fh[nfh++], 9});
        iov.push_back({ok_trailers.data(), ok_trailers.size()});
        finished = true;
      }
    }
    if (!iov.empty() && !c.w.writev_all(iov.data(), (int)iov.size())) return false;
    if (!o.active && o.borrowed) {  // only after the bytes left: the segments may move now
      o.borrowed = false;
      --law_borrowers;
    }
    if (finished) {
      close_stream(c, s.id);
      return true;
    }
    if (!(o.active && o.remaining && nfh >= kMaxFrames && budget)) return true;
  }
}

// true when a stream of the connection has queued bytes the peer's windows would take now
static bool sendable(const Conn &c) {
  if (c.dead || c.send_window <= 0) return false;
  for (const auto &kv : c.streams)
    if (kv.second.out.active && kv.second.out.remaining && kv.second.send_window > 0) return true;
  return false;
}

// one more slice of every suspended send
void gs_server::continue_sends() {
  for (auto &cp : conns) {
    Conn &c = *cp;
    if (sendable(c) && !flush_all(c)) c.dead = true;
  }
}

bool gs_server::flush_all(Conn &c) {
  // window opened: every stream with queued data may move (ids first: flush can erase)
  std::vector<uint32_t> ids;
  for (auto &kv : c.streams)
    if (kv.second.out.active) ids.push_back(kv.first);
  for (uint32_t id : ids) {
    auto it = c.streams.find(id);
    if (it != c.streams.end() && !flush(c, it->second)) return false;
  }
  return true;
}

bool gs_server::send_unary(Conn &c, Stream &s, const uint8_t *msg, size_t n) {
  Outgoing &o = s.out;
  o.owned.assign(msg, msg + n);
  o.prefix[0] = 0;
  put_be32(o.prefix + 1, (uint32_t)n);
  o.segs.clear();
  o.segs.push_back({o.prefix, 5});
  if (n) o.segs.push_back({o.owned.data(), n});
  o.seg = o.off = 0;
  o.remaining = 5 + n;
  o.active = true;
  o.borrowed = false;
  o.trailers_after = true;
  return flush(c, s);
}

// Ends the stream with a status: "Trailers-Only" when nothing was sent yet (gRPC over
// HTTP/2: one HEADERS frame with END_STREAM), plain trailers otherwise.
bool gs_server::send_error(Conn &c, Stream &s, int code, const std::string &msg) {
  std::vector<uint8_t> block;
  if (!s.headers_sent) block = resp_headers;
  char num[16];
  const int nn = snprintf(num, sizeof(num), "%d", code);
  encode_literal(block, "grpc-status", 11, num, (size_t)nn);
  if (!msg.empty()) {
    const std::string m = percent_encode(msg);
    encode_literal(block, "grpc-message", 12, m.data(), m.size());
  }
  uint8_t fh[9];
  put_frame_header(fh, (uint32_t)block.size(), kHeaders, kFlagEndHeaders | kFlagEndStream, s.id);
  struct iovec iov[2] = {{fh, 9}, {block.data(), block.size()}};
  const bool ok = c.w.writev_all(iov, 2);
  close_stream(c, s.id);
  return ok;
}

// ---- ListAndWatch (server.go:180-193) -----------------------------------------------------

void gs_server::queue_law(Conn &c, Stream &s) {
  if (s.out.active) return;
  const uint64_t gen = gs_law_generation(table);  // read before the snapshot: a change racing
                                                  // with it is sent again, never missed
  if (!law_iov_valid || law_iov_gen != gen) {
    if (law_borrowers) return;  // someone is still sending the old segments: pump_law retries
    uint32_t n = gs_table_gpu_count(table);
    law_iov.resize(n);
    if (gs_law_payload(table, law_iov.data(), n, &n) != GS_OK) return;
    law_iov.resize(n);
    law_iov_bytes = 0;
    for (auto &v : law_iov) law_iov_bytes += v.len;
    law_iov_gen = gen;
    law_iov_valid = true;
  }
  Outgoing &o = s.out;
  o.prefix[0] = 0;
  put_be32(o.prefix + 1, (uint32_t)law_iov_bytes);
  o.segs.clear();
  o.segs.push_back({o.prefix, 5});
  for (auto &v : law_iov)
    if (v.len) o.segs.push_back(v);
  o.seg = o.off = 0;
  o.remaining = 5 + law_iov_bytes;
  o.active = true;
  o.borrowed = true;
  o.trailers_after = false;
  ++law_borrowers;
  s.law_gen = law_iov_gen;
  s.law_sent_once = true;
  n_law_messages.fetch_add(1, std::memory_order_relaxed);
  if (!flush(c, s)) c.dead = true;
}

// (Re)sends the list on every open ListAndWatch stream that has not seen the current
// health generation; ends the streams once the plugin is stopped (<-m.stop, server.go:185).
void gs_server::pump_law() {
  const bool stopped = gs_plugin_stopped(plugin) != 0;
  const uint64_t gen = gs_law_generation(table);
  for (auto &cp : conns) {
    Conn &c = *cp;
    if (c.dead) continue;
    std::vector<uint32_t> ids;
    for (auto &kv : c.streams)
      if (kv.second.law) ids.push_back(kv.first);
    for (uint32_t id : ids) {
      auto it = c.streams.find(id);
      if (it == c.streams.end()) continue;
      Stream &s = it->second;
      if (stopped) {
        if (s.out.active) {  // mid-message: the message cannot be completed honestly
          c.w.send_rst(s.id, kErrCancel);
          close_stream(c, s.id);
        } else if (!s.headers_sent) {
          if (!send_error(c, s, kOk, "")) c.dead = true;
        } else {
          uint8_t fh[9];
          put_frame_header(fh, (uint32_t)ok_trailers.size(), kHeaders,
                           kFlagEndHeaders | kFlagEndStream, s.id);
          struct iovec iov[2] = {{fh, 9}, {ok_trailers.data(), ok_trailers.size()}};
          if (!c.w.writev_all(iov, 2)) c.dead = true;
          close_stream(c, s.id);
        }
      } else if (!s.out.active && (!s.law_sent_once || s.law_gen != gen)) {
        queue_law(c, s);
      }
      if (c.dead) break;
    }
  }
}

// ---- receiving ---------------------------------------------------------------------------

bool gs_server::on_headers_complete(Conn &c, uint32_t id, uint8_t flags) {
  c.headers.clear();
  if (!c.hpack.decode(c.header_block.data(), c.header_block.size(), c.headers)) {
    n_protocol_errors.fetch_add(1, std::memory_order_relaxed);
    c.w.send_goaway(c.last_stream, kErrCompression);
    return false;
  }
  auto it = c.streams.find(id);
  if (it != c.streams.end()) {
    // a second header block on an open stream: request trailers (gRPC clients send none)
    if (!(flags & kFlagEndStream)) {
      c.w.send_goaway(c.last_stream, kErrProtocol);
      return false;
    }
    if (!it->second.req) {  // the request was complete already (half-closed): RFC 7540 5.1
      c.w.send_rst(id, kErrStreamClosed);
      close_stream(c, id);
      return true;
    }
    return on_request_complete(c, it->second);
  }
  if (id <= c.last_stream) return true;  // headers of a stream already closed: ignore
  c.last_stream = id;
  if (c.streams.size() >= kMaxStreams)   // SETTINGS_MAX_CONCURRENT_STREAMS, RFC 7540 5.1.2
    return c.w.send_rst(id, kErrRefused);
  Stream &s = c.streams[id];
  s.id = id;
  s.send_window = c.peer_initial_window;
  bool is_post = false, has_ct = false;
  for (const Header &h : c.headers) {
    if (h.name == ":path") {
      s.path = h.value;
    } else if (h.name == ":method") {
      is_post = h.value == "POST";
    } else if (h.name == "content-type") {
      has_ct = h.value.compare(0, 16, "application/grpc") == 0;
    } else if (h.name == "grpc-encoding") {
      s.encoding = h.value;
    }
  }
  s.bad_content_type = !is_post || !has_ct;
  const size_t plen = sizeof(kServicePrefix) - 1;
  const std::string &path = s.path;
  if (path.compare(0, plen, kServicePrefix) == 0) {
    const std::string m = path.substr(plen);
    if (m == "Allocate") s.method = kAllocate;
    else if (m == "ListAndWatch") s.method = kListAndWatch;
    else if (m == "GetDevicePluginOptions") s.method = kOptions;
    else if (m == "PreStartContainer") s.method = kPreStart;
    else if (m == "GetPreferredAllocation") s.method = kPreferred;
  }
  if (!c.free_bufs.empty()) {
    s.req = std::move(c.free_bufs.back());
    c.free_bufs.pop_back();
  } else {
    s.req.reset(new Buf());
  }
  s.req->n = 0;
  if (flags & kFlagEndStream) return on_request_complete(c, s);
  return true;
}

// The request is half-closed: run the method. Every gRPC failure is a status on the stream,
// never a dropped connection.
bool gs_server::on_request_complete(Conn &c, Stream &s) {
  n_requests.fetch_add(1, std::memory_order_relaxed);
  if (s.bad_content_type) {  // not gRPC at all: HTTP 415, as grpc-go answers
    std::vector<uint8_t> block;
    encode_literal_static_name(block, 8, "415", 3);
    uint8_t fh[9];
    put_frame_header(fh, (uint32_t)block.size(), kHeaders, kFlagEndHeaders | kFlagEndStream, s.id);
    struct iovec iov[2] = {{fh, 9}, {block.data(), block.size()}};
    const bool ok = c.w.writev_all(iov, 2);
    close_stream(c, s.id);
    return ok;
  }
  const std::string &path = s.path;
  if (s.method == kUnknown) {
    // grpc-go's wording (server.go handleStream): unknown service / unknown method
    std::string msg = "malformed method name: \"" + path + "\"";
    const size_t slash = path.rfind('/');
    if (!path.empty() && path[0] == '/' && slash != std::string::npos && slash > 0) {
      const std::string service = path.substr(1, slash - 1), method = path.substr(slash + 1);
      if (service == "v1beta1.DevicePlugin") msg = "unknown method " + method + " for service " + service;
      else msg = "unknown service " + service;
    }
    return send_error(c, s, kUnimplemented, msg);
  }
  const uint8_t *body = s.req->p;
  const size_t n = s.req->n;
  if (n < 5) return send_error(c, s, kInternal, "grpc: the request carried no message");
  const uint32_t mlen = be32(body + 1);
  if (body[0] & 1)
    return send_error(c, s, kUnimplemented,
                      "grpc: Decompressor is not installed for grpc-encoding \"" + s.encoding + "\"");
  if ((size_t)mlen + 5 != n) return send_error(c, s, kInternal, "grpc: malformed message framing");
  const uint8_t *msg = body + 5;
  size_t out_len = 0;
  switch (s.method) {
    case kOptions: {
      if (gs_plugin_get_device_plugin_options(plugin, resp.data(), resp.size(), &out_len) != GS_OK)
        return send_error(c, s, kInternal, "GetDevicePluginOptions failed");
      return send_unary(c, s, resp.data(), out_len);
    }
    case kPreStart: {
      if (gs_plugin_pre_start_container(plugin, msg, mlen, resp.data(), resp.size(), &out_len) != GS_OK)
        return send_error(c, s, kInternal, "PreStartContainer failed");
      return send_unary(c, s, resp.data(), out_len);
    }
    case kPreferred:
      // server.go:37-40 panics ("implement me"); never advertised, so kubelet never calls it
      return send_error(c, s, kUnimplemented, "implement me");
    case kAllocate: {
      n_allocate.fetch_add(1, std::memory_order_relaxed);
      if (allocate_fn) {  // answered by deliver_done once the callback returns
        AllocJob j{c.serial, s.id, std::move(s.req)};
        {
          std::lock_guard<std::mutex> lk(job_mu);
          jobs.push_back(std::move(j));
        }
        job_cv.notify_one();
        return true;
      }
      int rc = GS_OK;
      const uint64_t t0 = mono_ns();
      for (int attempt = 0; attempt < 4; ++attempt) {
        gs_alloc_result res;
        rc = gs_plugin_allocate(plugin, msg, mlen, resp.data(), resp.size(), &res);
        out_len = res.resp_len;
        if (rc == GS_OK && decision_fn) decision_fn(user, &res);
        if (rc != GS_E_NOSPACE) break;
        resp.resize(out_len > resp.size() * 2 ? out_len : resp.size() * 2);
      }
      ns_allocate.fetch_add(mono_ns() - t0, std::memory_order_relaxed);
      if (rc == GS_E_WIRE)  // what grpc-go answers when proto.Unmarshal of the request fails
        return send_error(c, s, kInternal, "grpc: error unmarshalling request: malformed AllocateRequest");
      if (rc != GS_OK) return send_error(c, s, kUnknownCode, std::string("Allocate: ") + gs_strerror(rc));
      return send_unary(c, s, resp.data(), out_len);
    }
    case kListAndWatch: {
      s.law = true;
      s.req->n = 0;
      if (c.free_bufs.size() < 4) c.free_bufs.push_back(std::move(s.req));
      s.req.reset();
      if (gs_plugin_stopped(plugin)) return send_error(c, s, kOk, "");
      queue_law(c, s);
      return !c.dead;
    }
    default:
      break;
  }
  return send_error(c, s, kInternal, "unreachable");
}

bool gs_server::handle_frame(Conn &c, const FrameHeader &h) {
  Wire &w = c.w;
  if (c.cont_stream && h.type != kContinuation) {  // RFC 7540 6.10: nothing may interleave
    w.send_goaway(c.last_stream, kErrProtocol);
    return false;
  }
  switch (h.type) {
    case kData: {
      if (h.stream == 0) {
        w.send_goaway(c.last_stream, kErrProtocol);
        return false;
      }
      size_t len = h.length, pad = 0;
      c.recv_unacked += h.length;
      if (h.flags & kFlagPadded) {
        if (len < 1 || !w.fill(1)) return false;
        pad = w.peek()[0];
        w.consume(1);
        --len;
        if (pad > len) {
          w.send_goaway(c.last_stream, kErrProtocol);
          return false;
        }
        len -= pad;
      }
      auto it = c.streams.find(h.stream);
      if (it == c.streams.end() || !it->second.req) {
        // late frames of a stream we already answered (or never saw): discard
        if (!w.skip(len + pad)) return false;
        if (h.stream > c.last_stream) {
          w.send_goaway(c.last_stream, kErrProtocol);
          return false;
        }
      } else {
        Stream &s = it->second;
        if (s.req->n + len > max_request + 5) {
          if (!w.skip(len + pad)) return false;
          char m[96];
          snprintf(m, sizeof(m), "grpc: received message larger than max (%zu vs. %zu)",
                   s.req->n + len - 5, max_request);
          const uint32_t sid = s.id;
          if (!send_error(c, s, kResourceExhausted, m)) return false;
          // status first, then tell the client to stop sending the rest of the body
          if (!(h.flags & kFlagEndStream)) w.send_rst(sid, kErrNo);
        } else {
          if (!s.req->reserve(s.req->n + len)) return false;
          const uint64_t t0 = mono_ns();
          if (!w.read_into(s.req->p + s.req->n, len) || !w.skip(pad)) return false;
          ns_recv_body.fetch_add(mono_ns() - t0, std::memory_order_relaxed);
          s.req->n += len;
          s.recv_unacked += h.length;
          if (h.flags & kFlagEndStream) {
            if (!on_request_complete(c, s)) return false;
          } else if (s.recv_unacked >= kOurStreamWindow / 2) {
            if (!w.send_window_update(s.id, s.recv_unacked)) return false;
            s.recv_unacked = 0;
          }
        }
      }
      if (c.recv_unacked >= (256u << 20)) {
        if (!w.send_window_update(0, (uint32_t)c.recv_unacked)) return false;
        c.recv_unacked = 0;
      }
      return true;
    }
    case kHeaders:
    case kContinuation: {
      if (h.stream == 0 || !(h.stream & 1) || h.length > (64u << 10) ||
          c.header_block.size() + h.length > (256u << 10)) {
        w.send_goaway(c.last_stream, h.length > (64u << 10) ? kErrFrameSize : kErrProtocol);
        return false;
      }
      if (!w.fill(h.length)) return false;
      const uint8_t *p = w.peek();
      size_t n = h.length;
      if (h.type == kHeaders) {
        if (c.cont_stream) {  // a header block is still open on another frame sequence
          w.send_goaway(c.last_stream, kErrProtocol);
          return false;
        }
        if (!strip_headers_prefix(h.flags, p, n)) {
          w.send_goaway(c.last_stream, kErrProtocol);
          return false;
        }
        c.header_block.assign(p, p + n);
        c.cont_flags = h.flags;
      } else {
        if (c.cont_stream != h.stream) {
          w.send_goaway(c.last_stream, kErrProtocol);
          return false;
        }
        c.header_block.insert(c.header_block.end(), p, p + n);
      }
      w.consume(h.length);
      if (h.flags & kFlagEndHeaders) {
        c.cont_stream = 0;
        const bool ok = on_headers_complete(c, h.stream, c.cont_flags);
        c.header_block.clear();
        return ok;
      }
      c.cont_stream = h.stream;
      return true;
    }
    case kSettings: {
      if (h.stream != 0 || (h.flags & kFlagAck ? h.length != 0 : h.length % 6 != 0) ||
          h.length > 1024) {
        w.send_goaway(c.last_stream, kErrFrameSize);
        return false;
      }
      if (h.flags & kFlagAck) return true;
      if (!w.fill(h.length)) return false;
      const uint8_t *p = w.peek();
      for (uint32_t i = 0; i + 6 <= h.length; i += 6) {
        const uint16_t id = (uint16_t)(p[i] << 8 | p[i + 1]);
        const uint32_t v = be32(p + i + 2);
        if (id == kSetInitialWindowSize) {
          if (v > 0x7fffffffu) {
            w.send_goaway(c.last_stream, kErrFlowControl);
            return false;
          }
          const int64_t delta = (int64_t)v - (int64_t)c.peer_initial_window;
          c.peer_initial_window = v;
          for (auto &kv : c.streams) kv.second.send_window += delta;
        } else if (id == kSetMaxFrameSize) {
          if (v < kDefaultMaxFrame || v > kOurMaxFrame) {
            w.send_goaway(c.last_stream, kErrProtocol);
            return false;
          }
          c.peer_max_frame = v;
        }
      }
      w.consume(h.length);
      if (!w.send_settings_ack()) return false;
      return flush_all(c);
    }
    case kWindowUpdate: {
      if (h.length != 4) {
        w.send_goaway(c.last_stream, kErrFrameSize);
        return false;
      }
      if (!w.fill(4)) return false;
      const uint32_t inc = be32(w.peek()) & 0x7fffffffu;
      w.consume(4);
      if (h.stream == 0) {
        if (inc == 0 || c.send_window + inc > 0x7fffffffLL) {
          w.send_goaway(c.last_stream, inc ? kErrFlowControl : kErrProtocol);
          return false;
        }
        c.send_window += inc;
        return flush_all(c);
      }
      auto it = c.streams.find(h.stream);
      if (it == c.streams.end()) return true;  // closed stream: allowed for a short while
      if (inc == 0 || it->second.send_window + inc > 0x7fffffffLL) {
        w.send_rst(h.stream, inc ? kErrFlowControl : kErrProtocol);
        close_stream(c, h.stream);
        return true;
      }
      it->second.send_window += inc;
      return it->second.out.active ? flush(c, it->second) : true;
    }
    case kPing: {
      if (h.length != 8 || h.stream != 0) {
        w.send_goaway(c.last_stream, h.length != 8 ? kErrFrameSize : kErrProtocol);
        return false;
      }
      if (!w.fill(8)) return false;
      uint8_t opaque[8];
      memcpy(opaque, w.peek(), 8);
      w.consume(8);
      return (h.flags & kFlagAck) ? true : w.send_ping_ack(opaque);
    }
    case kRstStream: {
      if (h.length != 4 || h.stream == 0) {
        w.send_goaway(c.last_stream, h.length != 4 ? kErrFrameSize : kErrProtocol);
        return false;
      }
      if (!w.skip(4)) return false;
      close_stream(c, h.stream);  // a cancelled ListAndWatch: kubelet went away
      return true;
    }
    case kGoaway:
      // the peer will open no new streams; open ones finish, then it closes the socket
      return w.skip(h.length);
    case kPushPromise:
      w.send_goaway(c.last_stream, kErrProtocol);
      return false;
    default:  // PRIORITY and unknown extension frames are ignored (RFC 7540 4.1)
      return w.skip(h.length);
  }
}

// ---- the Allocate callback worker --------------------------------------------------------------

void gs_server::worker_main() {
  std::vector<uint8_t> out(64 << 10);
  for (;;) {
    AllocJob j;
    {
      std::unique_lock<std::mutex> lk(job_mu);
      job_cv.wait(lk, [&] { return worker_stop || !jobs.empty(); });
      if (jobs.empty()) return;  // stop requested and nothing left
      j = std::move(jobs.front());
      jobs.pop_front();
    }
    const uint8_t *msg = j.req->p + 5;
    const size_t mlen = j.req->n - 5;
    size_t out_len = 0;
    int rc = GS_OK;
    for (int attempt = 0; attempt < 4; ++attempt) {
      rc = allocate_fn(user, msg, mlen, out.data(), out.size(), &out_len);
      if (rc != GS_E_NOSPACE) break;
      out.resize(out_len > out.size() * 2 ? out_len : out.size() * 2);
    }
    AllocDone d{j.conn_serial, j.stream, rc, {}, std::move(j.req)};
    if (rc == GS_OK) d.resp.assign(out.begin(), out.begin() + (long)out_len);
    {
      std::lock_guard<std::mutex> lk(job_mu);
      done.push_back(std::move(d));
    }
    const uint64_t one = 1;
    if (write(wake_fd, &one, sizeof(one)) < 0) {
    }
  }
}

// Serving thread: send the answers the worker finished. A connection or stream that went
// away meanwhile (kubelet gave up, RST_STREAM) simply has nobody to answer.
void gs_server::deliver_done() {
  std::deque<AllocDone> ready;
  {
    std::lock_guard<std::mutex> lk(job_mu);
    ready.swap(done);
  }
  for (AllocDone &d : ready) {
    for (auto &cp : conns) {
      Conn &c = *cp;
      if (c.serial != d.conn_serial || c.dead) continue;
      auto it = c.streams.find(d.stream);
      if (it == c.streams.end()) break;
      Stream &s = it->second;
      if (d.req) {
        d.req->n = 0;
        s.req = std::move(d.req);  // close_stream returns it to the pool
      }
      bool ok;
      if (d.rc == GS_E_WIRE)  // what grpc-go answers when proto.Unmarshal of the request fails
        ok = send_error(c, s, kInternal, "grpc: error unmarshalling request: malformed AllocateRequest");
      else if (d.rc != GS_OK)
        ok = send_error(c, s, kUnknownCode, std::string("Allocate: ") + gs_strerror(d.rc));
      else
        ok = send_unary(c, s, d.resp.data(), d.resp.size());
      if (!ok) c.dead = true;
      break;
    }
  }
}

// ---- the serving thread --------------------------------------------------------------------

void gs_server::drop_conn(Conn &c) {
  std::vector<uint32_t> ids;
  for (auto &kv : c.streams) ids.push_back(kv.first);
  for (uint32_t id : ids) close_stream(c, id);  // gives borrowed segments back
  n_bytes_in.fetch_add(c.w.bytes_in, std::memory_order_relaxed);
  n_bytes_out.fetch_add(c.w.bytes_out, std::memory_order_relaxed);
  if (c.w.fd >= 0) close(c.w.fd);
  c.w.fd = -1;
  c.dead = true;
}

void gs_server::accept_one() {
  const int fd = accept4(listen_fd, nullptr, nullptr, SOCK_CLOEXEC);
  if (fd < 0) {
    if (errno == EMFILE || errno == ENFILE || errno == ENOMEM) usleep(10000);  // do not spin on poll
    return;
  }
  if (conns.size() >= kMaxConns) {
    close(fd);
    return;
  }
  set_io_timeouts(fd, 10);
  set_big_buffers(fd, 4 << 20);
  std::unique_ptr<Conn> c(new Conn());
  c->w.fd = fd;
  c->serial = next_conn_serial++;
  // server connection preface (RFC 7540 3.5): SETTINGS, then open the connection window
  const uint16_t ids[3] = {kSetMaxConcurrentStreams, kSetInitialWindowSize, kSetMaxFrameSize};
  const uint32_t vals[3] = {(uint32_t)kMaxStreams, kOurStreamWindow, kOurMaxFrame};
  if (!c->w.send_settings(ids, vals, 3) ||
      !c->w.send_window_update(0, kOurConnWindow - kDefaultWindow)) {
    close(fd);
    return;
  }
  n_connections.fetch_add(1, std::memory_order_relaxed);
  conns.push_back(std::move(c));
}

void gs_server::serve_conn(Conn &c) {
  if (!c.preface_seen) {
    if (!c.w.fill(24) || memcmp(c.w.peek(), kPreface, 24) != 0) {
      c.dead = true;
      return;
    }
    c.w.consume(24);
    c.preface_seen = true;
    if (!c.w.buffered()) return;
  }
  do {
    FrameHeader h;
    if (!c.w.read_frame_header(h) || !handle_frame(c, h)) {
      c.dead = true;
      return;
    }
  } while (!c.dead && c.w.buffered());
}

void gs_server::run() {
  std::vector<struct pollfd> pfds;
  while (!stop_requested.load(std::memory_order_acquire)) {
    pfds.clear();
    pfds.push_back({wake_fd, POLLIN, 0});
    pfds.push_back({listen_fd, POLLIN, 0});
    for (auto &c : conns) pfds.push_back({c->w.fd, POLLIN, 0});
    bool pending = false;
    for (auto &c : conns) pending = pending || sendable(*c);
    const int r = poll(pfds.data(), (nfds_t)pfds.size(), pending ? 0 : 200);
    if (r < 0 && errno != EINTR) break;
    if (r <= 0) {
      if (pending) continue_sends();
      pump_law();  // also: a health change nobody announced with gs_plugin_notify_health
      continue;
    }
    const size_t n_conns = conns.size();  // accept_one may append
    if (pfds[0].revents & POLLIN) {
      uint64_t v;
      while (read(wake_fd, &v, sizeof(v)) > 0) {
      }
      deliver_done();
    }
    for (size_t i = 0; i < n_conns; ++i) {
      if (pfds[2 + i].revents & (POLLIN | POLLHUP | POLLERR)) serve_conn(*conns[i]);
    }
    if (pfds[1].revents & POLLIN) accept_one();
    if (pending) continue_sends();
    // health changed, a list finished (so a deferred one may start), or the plugin stopped
    pump_law();
    for (size_t i = 0; i < conns.size();) {
      if (conns[i]->dead) {
        drop_conn(*conns[i]);
        conns.erase(conns.begin() + (long)i);
      } else {
        ++i;
      }
    }
  }
  // Stop (server.go:145-155): end the watch streams, say goodbye, close
  pump_law();
  for (auto &c : conns) {
    if (!c->dead) c->w.send_goaway(c->last_stream, kErrNo);
    drop_conn(*c);
  }
  conns.clear();
}

// ---- C ABI ----------------------------------------------------------------------------------

extern "C" {

int gs_server_start(gs_plugin *plugin, const gs_server_config *cfg, gs_server **out) {
  if (!out) return GS_E_INVAL;
  *out = nullptr;
  if (!plugin || !cfg || !cfg->socket_path) return GS_E_INVAL;
  struct sockaddr_un sa;
  socklen_t salen;
  if (!fill_sockaddr(cfg->socket_path, sa, salen)) return GS_E_INVAL;
  std::unique_ptr<gs_server> s(new (std::nothrow) gs_server());
  if (!s) return GS_E_INVAL;
  s->plugin = plugin;
  s->table = gs_internal::plugin_table(plugin);
  s->socket_path = cfg->socket_path;
  s->allocate_fn = cfg->allocate;
  s->decision_fn = cfg->on_decision;
  s->user = cfg->user;
  if (cfg->max_request_bytes) s->max_request = cfg->max_request_bytes;
  s->resp.resize(64 << 10);
  s->iov_scratch.reserve(256);
  encode_indexed(s->resp_headers, 8);                                         // :status 200
  encode_literal_static_name(s->resp_headers, 31, "application/grpc", 16);    // content-type
  encode_literal(s->ok_trailers, "grpc-status", 11, "0", 1);

  unlink(cfg->socket_path);  // cleanup(), server.go:203-209: a stale socket from a previous run
  s->listen_fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (s->listen_fd < 0) return GS_E_STATE;
  if (bind(s->listen_fd, (struct sockaddr *)&sa, salen) != 0 || listen(s->listen_fd, 16) != 0) {
    close(s->listen_fd);
    return GS_E_STATE;
  }
  s->wake_fd = eventfd(0, EFD_NONBLOCK | EFD_CLOEXEC);
  if (s->wake_fd < 0) {
    close(s->listen_fd);
    unlink(cfg->socket_path);
    return GS_E_STATE;
  }
  gs_internal::plugin_add_wake_fd(plugin, s->wake_fd);
  gs_server *raw = s.get();
  if (s->allocate_fn) s->worker = std::thread([raw] { raw->worker_main(); });
  s->thread = std::thread([raw] { raw->run(); });

  // server.go:131: dial our own socket (blocking, 5 s) to know that the server answers
  gs_grpc_client *probe = nullptr;
  const int rc = gs_grpc_client_connect(cfg->socket_path, 5000, &probe);
  if (probe) gs_grpc_client_close(probe);
  if (rc != GS_OK) {
    gs_server *dead = s.release();
    gs_server_stop(dead);
    return GS_E_TIMEOUT;
  }
  *out = s.release();
  return GS_OK;
}

int gs_server_stop(gs_server *s) {
  if (!s) return GS_E_INVAL;
  gs_plugin_stop(s->plugin);  // close(m.stop), server.go:152
  s->stop_requested.store(true, std::memory_order_release);
  const uint64_t one = 1;
  if (write(s->wake_fd, &one, sizeof(one)) < 0) {
  }
  if (s->thread.joinable()) s->thread.join();
  {
    std::lock_guard<std::mutex> lk(s->job_mu);
    s->worker_stop = true;
    s->jobs.clear();  // their streams are gone with the connections
  }
  s->job_cv.notify_all();
  if (s->worker.joinable()) s->worker.join();
  gs_internal::plugin_remove_wake_fd(s->plugin, s->wake_fd);
  close(s->wake_fd);
  close(s->listen_fd);
  unlink(s->socket_path.c_str());  // cleanup(), server.go:154
  delete s;
  return GS_OK;
}

int gs_server_get_stats(const gs_server *s, gs_server_stats *out) {
  if (!s || !out) return GS_E_INVAL;
  out->connections = s->n_connections.load(std::memory_order_relaxed);
  out->requests = s->n_requests.load(std::memory_order_relaxed);
  out->allocate_calls = s->n_allocate.load(std::memory_order_relaxed);
  out->law_messages = s->n_law_messages.load(std::memory_order_relaxed);
  out->protocol_errors = s->n_protocol_errors.load(std::memory_order_relaxed);
  out->bytes_in = s->n_bytes_in.load(std::memory_order_relaxed);    // of closed connections
  out->bytes_out = s->n_bytes_out.load(std::memory_order_relaxed);
  out->recv_body_ns = s->ns_recv_body.load(std::memory_order_relaxed);
  out->allocate_ns = s->ns_allocate.load(std::memory_order_relaxed);
  return GS_OK;
}

}  // extern "C"

// fuzz_transport.cpp -- the native gRPC transport under sanitizers.
//
// Built by tests/test_native_transport.py twice: with ASan+UBSan (argv[1] = "fuzz") and with
// TSan (argv[1] = "race"), from csrc/*.cpp directly (no CUDA needed).
//
//   fuzz: a valid client session (preface, SETTINGS, HEADERS, DATA of an Allocate call) is
//         mutated -- bytes flipped, frames truncated, duplicated, reordered, random frames
//         spliced in -- and thrown at a live gs_server over its unix socket; then the same for
//         the client, against a fake server that answers with mutated server sessions. After
//         every batch a well-formed call must still succeed.
//   race: Allocate callers, a ListAndWatch reader, a health marker and Stop, concurrently.
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <atomic>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "gpushare.h"
#include "h2conn.h"
#include "hpack.h"

using namespace gs_h2;

static std::string g_dir;

#define CHECK(x)                                                      \
  do {                                                                \
    if (!(x)) {                                                       \
      fprintf(stderr, "CHECK failed: %s (line %d)\n", #x, __LINE__);  \
      exit(1);                                                        \
    }                                                                 \
  } while (0)

static std::vector<uint8_t> allocate_request(const char *uuid, int n) {
  std::vector<uint8_t> inner;
  for (int j = 0; j < n; ++j) {
    char id[128];
    const int k = snprintf(id, sizeof(id), "%s-_-%d", uuid, j);
    inner.push_back(0x0a);
    inner.push_back((uint8_t)k);
    inner.insert(inner.end(), id, id + k);
  }
  std::vector<uint8_t> out;
  out.push_back(0x0a);
  size_t v = inner.size();
  while (v >= 0x80) out.push_back((uint8_t)(v | 0x80)), v >>= 7;
  out.push_back((uint8_t)v);
  out.insert(out.end(), inner.begin(), inner.end());
  return out;
}

static void frame(std::vector<uint8_t> &o, uint8_t type, uint8_t flags, uint32_t stream,
                  const std::vector<uint8_t> &payload) {
  uint8_t h[9];
  put_frame_header(h, (uint32_t)payload.size(), type, flags, stream);
  o.insert(o.end(), h, h + 9);
  o.insert(o.end(), payload.begin(), payload.end());
}

// preface + SETTINGS + one complete call on stream `id`
static std::vector<uint8_t> client_session(const char *path, const std::vector<uint8_t> &msg,
                                           uint32_t id, bool with_preface) {
  std::vector<uint8_t> o;
  if (with_preface) {
    o.insert(o.end(), kPreface, kPreface + 24);
    frame(o, kSettings, 0, 0, {});
  }
  std::vector<uint8_t> block;
  encode_indexed(block, 3);
  encode_indexed(block, 6);
  encode_literal_static_name(block, 4, path, strlen(path));
  encode_literal_static_name(block, 31, "application/grpc", 16);
  encode_literal(block, "te", 2, "trailers", 8);
  frame(o, kHeaders, kFlagEndHeaders, id, block);
  std::vector<uint8_t> body(5);
  body[0] = 0;
  put_be32(body.data() + 1, (uint32_t)msg.size());
  body.insert(body.end(), msg.begin(), msg.end());
  frame(o, kData, kFlagEndStream, id, body);
  return o;
}

static int dial(const std::string &path) {
  struct sockaddr_un sa;
  socklen_t len;
  CHECK(fill_sockaddr(path.c_str(), sa, len));
  const int fd = socket(AF_UNIX, SOCK_STREAM, 0);
  CHECK(fd >= 0);
  CHECK(connect(fd, (struct sockaddr *)&sa, len) == 0);
  set_io_timeouts(fd, 5);
  return fd;
}

// send everything, half-close, read until the server closes
static size_t throw_at_server(const std::string &sock, const std::vector<uint8_t> &bytes) {
  const int fd = dial(sock);
  size_t off = 0;
  while (off < bytes.size()) {
    const ssize_t w = send(fd, bytes.data() + off, bytes.size() - off, MSG_NOSIGNAL);
    if (w <= 0) break;  // the server hung up on us already: fine
    off += (size_t)w;
  }
  shutdown(fd, SHUT_WR);
  size_t got = 0;
  uint8_t buf[65536];
  for (;;) {
    const ssize_t r = recv(fd, buf, sizeof(buf), 0);
    if (r <= 0) break;
    got += (size_t)r;
    if (got > (64u << 20)) break;
  }
  close(fd);
  return got;
}

static void mutate(std::vector<uint8_t> &b, std::mt19937 &rng) {
  const int kind = (int)(rng() % 9);
  if (b.empty()) return;
  switch (kind) {
    case 0:  // flip a few bytes anywhere
      for (int i = 0, n = 1 + (int)(rng() % 6); i < n; ++i) b[rng() % b.size()] ^= (uint8_t)(1u << (rng() % 8));
      break;
    case 1:  // truncate
      b.resize(rng() % b.size());
      break;
    case 2: {  // overwrite a run with random bytes
      const size_t at = rng() % b.size(), n = 1 + rng() % 24;
      for (size_t i = at; i < b.size() && i < at + n; ++i) b[i] = (uint8_t)rng();
      break;
    }
    case 3: {  // splice a random frame after the preface
      std::vector<uint8_t> f, payload(rng() % 40);
      for (auto &x : payload) x = (uint8_t)rng();
      frame(f, (uint8_t)(rng() % 11), (uint8_t)rng(), (uint32_t)(rng() % 8), payload);
      const size_t at = b.size() > 24 ? 24 + rng() % (b.size() - 24) : b.size();
      b.insert(b.begin() + (long)at, f.begin(), f.end());
      break;
    }
    case 4: {  // lie about a frame length
      if (b.size() > 40) {
        const size_t at = 24 + rng() % 16;
        b[at] = (uint8_t)rng();
      }
      break;
    }
    case 5: {  // duplicate the tail (a second HEADERS/DATA for the same stream)
      const size_t from = b.size() > 33 ? 33 : 0;
      std::vector<uint8_t> tail(b.begin() + (long)from, b.end());
      b.insert(b.end(), tail.begin(), tail.end());
      break;
    }
    case 6: {  // a late header block (END_STREAM) on a stream whose request is already complete
      std::vector<uint8_t> block;
      encode_literal(block, "x-late", 6, "trailer", 7);
      frame(b, kHeaders, kFlagEndHeaders | kFlagEndStream, (uint32_t)(1 + 2 * (rng() % 3)), block);
      break;
    }
    case 7: {  // DATA / WINDOW_UPDATE / RST_STREAM for streams in every state
      std::vector<uint8_t> payload(4 + rng() % 8);
      for (auto &x : payload) x = (uint8_t)rng();
      const uint8_t types[3] = {kData, kWindowUpdate, kRstStream};
      frame(b, types[rng() % 3], (uint8_t)(rng() % 2), (uint32_t)(rng() % 8), payload);
      break;
    }
    default:  // append garbage
      for (int i = 0, n = (int)(rng() % 64); i < n; ++i) b.push_back((uint8_t)rng());
  }
}

static void good_call(const std::string &sock, const std::vector<uint8_t> &req) {
  gs_grpc_client *c = nullptr;
  CHECK(gs_grpc_client_connect(sock.c_str(), 5000, &c) == GS_OK);
  uint8_t resp[4096];
  size_t n = 0;
  int32_t st = -1;
  char msg[128];
  CHECK(gs_grpc_client_unary(c, "/v1beta1.DevicePlugin/Allocate", req.data(), req.size(), resp,
                             sizeof(resp), &n, &st, msg, sizeof(msg)) == GS_OK);
  CHECK(st == 0 && n > 100);
  uint32_t sid = 0;
  CHECK(gs_grpc_client_stream_open(c, "/v1beta1.DevicePlugin/ListAndWatch", nullptr, 0, &sid) == GS_OK);
  const uint8_t *p = nullptr;
  CHECK(gs_grpc_client_stream_next_borrow(c, sid, &p, &n, &st, 5000) == GS_OK);
  CHECK(n > 1000 && p[0] == 0x0a);
  gs_grpc_client_close(c);
}

// a fake server: accepts one connection, writes `bytes`, half-closes, drains, closes
struct FakeServer {
  int lfd = -1;
  std::string path;
  std::thread th;
  std::vector<uint8_t> bytes;
  void start(const std::string &p, std::vector<uint8_t> b) {
    path = p;
    bytes = std::move(b);
    unlink(path.c_str());
    struct sockaddr_un sa;
    socklen_t len;
    CHECK(fill_sockaddr(path.c_str(), sa, len));
    lfd = socket(AF_UNIX, SOCK_STREAM, 0);
    CHECK(lfd >= 0 && bind(lfd, (struct sockaddr *)&sa, len) == 0 && listen(lfd, 4) == 0);
    th = std::thread([this] {
      const int fd = accept(lfd, nullptr, nullptr);
      if (fd < 0) return;
      set_io_timeouts(fd, 2);
      size_t off = 0;
      while (off < bytes.size()) {
        const ssize_t w = send(fd, bytes.data() + off, bytes.size() - off, MSG_NOSIGNAL);
        if (w <= 0) break;
        off += (size_t)w;
      }
      shutdown(fd, SHUT_WR);
      uint8_t buf[4096];
      while (recv(fd, buf, sizeof(buf), 0) > 0) {
      }
      close(fd);
    });
  }
  void stop() {
    if (th.joinable()) th.join();
    close(lfd);
    unlink(path.c_str());
  }
};

static std::vector<uint8_t> server_session(uint32_t id) {
  std::vector<uint8_t> o, block, trailers, body;
  frame(o, kSettings, 0, 0, {});
  encode_indexed(block, 8);
  encode_literal_static_name(block, 31, "application/grpc", 16);
  frame(o, kHeaders, kFlagEndHeaders, id, block);
  body = {0, 0, 0, 0, 3, 'a', 'b', 'c'};
  frame(o, kData, 0, id, body);
  encode_literal(trailers, "grpc-status", 11, "0", 1);
  frame(o, kHeaders, kFlagEndHeaders | kFlagEndStream, id, trailers);
  return o;
}

static gs_table *make_table(const char *uuid) {
  gs_table *t = nullptr;
  const char *uuids[1] = {uuid};
  const uint32_t lens[1] = {(uint32_t)strlen(uuid)};
  CHECK(gs_table_build(uuids, lens, 1, 64, GS_UNIT_GIB, &t) == GS_OK);
  return t;
}

static int run_fuzz(int iterations) {
  const char *uuid = "GPU-0f9b9cf4-2b1c-4a3e-9c1d-5e2f6a7b8c9d";
  gs_table *t = make_table(uuid);
  gs_plugin *p = nullptr;
  CHECK(gs_plugin_new(t, "node-a", 6, 0, &p) == GS_OK);
  const std::string sock = g_dir + "/srv.sock";
  gs_server_config cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.socket_path = sock.c_str();
  cfg.max_request_bytes = 1 << 20;
  gs_server *srv = nullptr;
  CHECK(gs_server_start(p, &cfg, &srv) == GS_OK);
  const std::vector<uint8_t> req = allocate_request(uuid, 4);
  good_call(sock, req);
  std::mt19937 rng(20261106);
  const std::vector<uint8_t> base = client_session("/v1beta1.DevicePlugin/Allocate", req, 1, true);
  CHECK(throw_at_server(sock, base) > 100);  // unmutated: the server answers the call
  for (int i = 0; i < iterations; ++i) {
    std::vector<uint8_t> b = base;
    if (rng() % 4 == 0) {  // a second and third call on the same connection
      std::vector<uint8_t> more = client_session("/v1beta1.DevicePlugin/ListAndWatch", {}, 3, false);
      b.insert(b.end(), more.begin(), more.end());
      more = client_session("/v1beta1.DevicePlugin/GetDevicePluginOptions", {}, 5, false);
      b.insert(b.end(), more.begin(), more.end());
    }
    for (int k = 0, n = 1 + (int)(rng() % 3); k < n; ++k) mutate(b, rng);
    throw_at_server(sock, b);
    if (i % 100 == 99) {
      gs_law_mark_unhealthy_gpu(t, 0);
      gs_plugin_notify_health(p);
      good_call(sock, req);
      gs_law_reset_health(t);
    }
  }
  gs_server_stats st;
  CHECK(gs_server_get_stats(srv, &st) == GS_OK);
  printf("server fuzz: %d sessions, %llu connections, %llu requests, %llu compression errors\n", iterations,
         (unsigned long long)st.connections, (unsigned long long)st.requests,
         (unsigned long long)st.protocol_errors);
  CHECK(gs_server_stop(srv) == GS_OK);
  gs_plugin_free(p);
  gs_table_free(t);

  // the client against a lying server
  const std::vector<uint8_t> sbase = server_session(1);
  int ok_calls = 0;
  for (int i = 0; i < iterations / 4; ++i) {
    std::vector<uint8_t> b = sbase;
    if (i) mutate(b, rng);
    FakeServer fs;
    fs.start(g_dir + "/fake.sock", b);
    gs_grpc_client *c = nullptr;
    if (gs_grpc_client_connect(fs.path.c_str(), 2000, &c) == GS_OK) {
      uint8_t resp[64];
      size_t n = 0;
      int32_t st2 = -1;
      char msg[64];
      const int rc = gs_grpc_client_unary(c, "/x/y", (const uint8_t *)"q", 1, resp, sizeof(resp), &n, &st2,
                                          msg, sizeof(msg));
      if (rc == GS_OK && st2 == 0 && n == 3 && !memcmp(resp, "abc", 3)) ++ok_calls;
      gs_grpc_client_close(c);
    }
    fs.stop();
  }
  printf("client fuzz: %d sessions, %d still complete calls\n", iterations / 4, ok_calls);
  CHECK(ok_calls >= 1);
  return 0;
}

// the caller-supplied Allocate (gs_server_config.allocate): runs on the server's worker thread
static int allocate_via_callback(void *user, const uint8_t *req, size_t req_len, uint8_t *resp, size_t cap,
                                 size_t *resp_len) {
  gs_alloc_result res;
  const int rc = gs_plugin_allocate((gs_plugin *)user, req, req_len, resp, cap, &res);
  *resp_len = res.resp_len;
  usleep(200);  // a little "network"
  return rc;
}

static int run_race(int seconds, bool with_callback) {
  const char *uuid = "GPU-0f9b9cf4-2b1c-4a3e-9c1d-5e2f6a7b8c9d";
  gs_table *t = make_table(uuid);
  gs_plugin *p = nullptr;
  CHECK(gs_plugin_new(t, "node-a", 6, 0, &p) == GS_OK);
  const std::string sock = g_dir + "/race.sock";
  gs_server_config cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.socket_path = sock.c_str();
  if (with_callback) {
    cfg.allocate = allocate_via_callback;
    cfg.user = p;
  }
  gs_server *srv = nullptr;
  CHECK(gs_server_start(p, &cfg, &srv) == GS_OK);
  const std::vector<uint8_t> req = allocate_request(uuid, 8);
  std::atomic<bool> stop{false};
  std::atomic<long> calls{0}, lists{0};
  std::vector<std::thread> th;
  for (int k = 0; k < 2; ++k)
    th.emplace_back([&] {
      gs_grpc_client *c = nullptr;
      CHECK(gs_grpc_client_connect(sock.c_str(), 5000, &c) == GS_OK);
      uint8_t resp[4096];
      size_t n;
      int32_t st;
      char msg[64];
      while (!stop.load()) {
        CHECK(gs_grpc_client_unary(c, "/v1beta1.DevicePlugin/Allocate", req.data(), req.size(), resp,
                                   sizeof(resp), &n, &st, msg, sizeof(msg)) == GS_OK && st == 0);
        ++calls;
      }
      gs_grpc_client_close(c);
    });
  th.emplace_back([&] {
    gs_grpc_client *c = nullptr;
    CHECK(gs_grpc_client_connect(sock.c_str(), 5000, &c) == GS_OK);
    uint32_t sid = 0;
    CHECK(gs_grpc_client_stream_open(c, "/v1beta1.DevicePlugin/ListAndWatch", nullptr, 0, &sid) == GS_OK);
    const uint8_t *m;
    size_t n;
    int32_t st;
    for (;;) {
      const int rc = gs_grpc_client_stream_next_borrow(c, sid, &m, &n, &st, 200);
      if (rc == GS_OK) ++lists;
      else if (rc == GS_E_TIMEOUT) continue;
      else break;  // the stream ended: Stop
    }
    CHECK(st == 0);  // ListAndWatch returns nil on m.stop
    gs_grpc_client_close(c);
  });
  th.emplace_back([&] {
    while (!stop.load()) {
      gs_law_mark_unhealthy_gpu(t, 0);
      gs_plugin_notify_health(p);
      usleep(2000);
      gs_law_reset_health(t);
      gs_plugin_notify_health(p);
      usleep(2000);
    }
  });
  sleep((unsigned)seconds);
  stop.store(true);
  th[0].join();
  th[1].join();
  th[3].join();
  CHECK(gs_server_stop(srv) == GS_OK);  // ends the watch stream with OK
  th[2].join();
  printf("race%s: %ld allocate calls, %ld lists\n", with_callback ? " (callback worker)" : "", calls.load(),
         lists.load());
  CHECK(calls.load() > 10 && lists.load() > 2);
  gs_plugin_free(p);
  gs_table_free(t);
  return 0;
}

int main(int argc, char **argv) {
  char tmpl[] = "/tmp/gsfuzzXXXXXX";
  CHECK(mkdtemp(tmpl) != nullptr);
  g_dir = tmpl;
  const std::string mode = argc > 1 ? argv[1] : "fuzz";
  const int n = argc > 2 ? atoi(argv[2]) : 0;
  int rc;
  if (mode == "race") rc = run_race(n > 0 ? n : 2, false) || run_race(n > 0 ? n : 2, true);
  else rc = run_fuzz(n > 0 ? n : 1500);
  rmdir(tmpl);
  if (rc == 0) printf("TRANSPORT %s OK\n", mode == "race" ? "RACE" : "FUZZ");
  return rc;
}

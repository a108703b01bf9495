// h2conn.h -- HTTP/2 framing (RFC 7540) over a blocking unix-domain stream socket, shared
// by the gRPC server (grpc_server.cpp) and client (grpc_client.cpp).
//
// Stands where google.golang.org/grpc's http2 transport stands in the reference
// (server.go:120-129 grpc.NewServer/Serve, server.go:158-177,252-265 grpc.Dial). Only what
// gRPC over a unix socket needs: no TLS, no push, no priorities (parsed and ignored).
//
// I/O model: sockets are blocking with send/receive timeouts; one thread owns a connection.
// Reads go through a small buffer so that frame headers cost no syscall of their own, and
// DATA payloads are received straight into the message buffer of their stream (one copy,
// kernel -> message). Writes are gathered with writev so that a response (HEADERS + DATA +
// trailers) or a run of DATA frames borrowed from the device table is one syscall.
#pragma once
#include <errno.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <sys/uio.h>
#include <sys/un.h>
#include <unistd.h>

#include <vector>

namespace gs_h2 {

enum FrameType : uint8_t {
  kData = 0,
  kHeaders = 1,
  kPriority = 2,
  kRstStream = 3,
  kSettings = 4,
  kPushPromise = 5,
  kPing = 6,
  kGoaway = 7,
  kWindowUpdate = 8,
  kContinuation = 9,
};
enum : uint8_t {
  kFlagEndStream = 0x1,
  kFlagAck = 0x1,
  kFlagEndHeaders = 0x4,
  kFlagPadded = 0x8,
  kFlagPriority = 0x20,
};
enum : uint16_t {
  kSetHeaderTableSize = 1,
  kSetEnablePush = 2,
  kSetMaxConcurrentStreams = 3,
  kSetInitialWindowSize = 4,
  kSetMaxFrameSize = 5,
  kSetMaxHeaderListSize = 6,
};
enum : uint32_t {  // RFC 7540 section 7
  kErrNo = 0,
  kErrProtocol = 1,
  kErrInternal = 2,
  kErrFlowControl = 3,
  kErrStreamClosed = 5,
  kErrFrameSize = 6,
  kErrRefused = 7,
  kErrCancel = 8,
  kErrCompression = 9,
};
constexpr uint32_t kDefaultWindow = 65535;
constexpr uint32_t kDefaultMaxFrame = 16384;
// what this endpoint advertises: windows and frames large enough that a 2 MB
// AllocateRequest (40 960 ids at MiB granularity) arrives without a flow-control stall
constexpr uint32_t kOurMaxFrame = (1u << 24) - 1;
constexpr uint32_t kOurStreamWindow = 64u << 20;
constexpr uint32_t kOurConnWindow = 0x7fffffffu;
static const char kPreface[] = "PRI * HTTP/2.0\r\n\r\nSM\r\n\r\n";  // 24 bytes

struct FrameHeader {
  uint32_t length;
  uint8_t type;
  uint8_t flags;
  uint32_t stream;
};

inline void put_frame_header(uint8_t *p, uint32_t length, uint8_t type, uint8_t flags,
                             uint32_t stream) {
  p[0] = (uint8_t)(length >> 16);
  p[1] = (uint8_t)(length >> 8);
  p[2] = (uint8_t)length;
  p[3] = type;
  p[4] = flags;
  p[5] = (uint8_t)((stream >> 24) & 0x7f);
  p[6] = (uint8_t)(stream >> 16);
  p[7] = (uint8_t)(stream >> 8);
  p[8] = (uint8_t)stream;
}
inline uint32_t be32(const uint8_t *p) {
  return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3];
}
inline void put_be32(uint8_t *p, uint32_t v) {
  p[0] = (uint8_t)(v >> 24);
  p[1] = (uint8_t)(v >> 16);
  p[2] = (uint8_t)(v >> 8);
  p[3] = (uint8_t)v;
}

inline void set_io_timeouts(int fd, int seconds) {
  struct timeval tv;
  tv.tv_sec = seconds;
  tv.tv_usec = 0;
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof(tv));
}

// Asks for large socket buffers: a whole AllocateRequest / a long run of list frames in
// flight saves a sleep and a wake-up per 208 KB (net.core.wmem_default). The FORCE forms
// need CAP_NET_ADMIN (the DaemonSet runs privileged); without it the plain forms are
// clamped to net.core.[rw]mem_max and that is fine too.
inline void set_big_buffers(int fd, int bytes) {
  if (setsockopt(fd, SOL_SOCKET, SO_SNDBUFFORCE, &bytes, sizeof(bytes)) != 0)
    setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &bytes, sizeof(bytes));
  if (setsockopt(fd, SOL_SOCKET, SO_RCVBUFFORCE, &bytes, sizeof(bytes)) != 0)
    setsockopt(fd, SOL_SOCKET, SO_RCVBUF, &bytes, sizeof(bytes));
}

inline bool fill_sockaddr(const char *path, struct sockaddr_un &sa, socklen_t &len) {
  const size_t n = strlen(path);
  if (n == 0 || n >= sizeof(sa.sun_path)) return false;
  memset(&sa, 0, sizeof(sa));
  sa.sun_family = AF_UNIX;
  memcpy(sa.sun_path, path, n);
  len = (socklen_t)(offsetof(struct sockaddr_un, sun_path) + n + 1);
  return true;
}

// Buffered reader + gathered writer over one socket.
struct Wire {
  int fd = -1;
  std::vector<uint8_t> rbuf;
  size_t rpos = 0, rend = 0;
  uint64_t bytes_in = 0, bytes_out = 0;

  static constexpr size_t kReadAhead = 16 << 10;
  explicit Wire(size_t cap = 256 << 10) : rbuf(cap) {}

  size_t buffered() const { return rend - rpos; }

  // makes at least `need` bytes (<= capacity) available at rbuf[rpos]
  bool fill(size_t need) {
    if (buffered() >= need) return true;
    if (rpos && rpos + need > rbuf.size()) {
      memmove(rbuf.data(), rbuf.data() + rpos, buffered());
      rend -= rpos;
      rpos = 0;
    } else if (rpos == rend) {
      rpos = rend = 0;
    }
    while (buffered() < need) {
      // read ahead, but not far: what is read here and turns out to be a DATA payload is copied a
      // second time by read_into, while the payload's remainder goes straight to its destination.
      // 16 KiB takes a small request whole in one syscall and costs a 2 MB one ~1 us of copying
      // instead of the ~20 us a buffer-full (256 KiB) of its body did.
      size_t room = rbuf.size() - rend;
      const size_t want = need - buffered() > kReadAhead ? need - buffered() : kReadAhead;
      if (room > want) room = want;
      const ssize_t r = recv(fd, rbuf.data() + rend, room, 0);
      if (r > 0) {
        rend += (size_t)r;
        bytes_in += (uint64_t)r;
      } else if (r < 0 && errno == EINTR) {
        continue;
      } else {
        return false;  // EOF, timeout or error
      }
    }
    return true;
  }
  const uint8_t *peek() const { return rbuf.data() + rpos; }
  void consume(size_t n) { rpos += n; }

  // n bytes into dst: buffered bytes first, the rest straight from the socket
  bool read_into(uint8_t *dst, size_t n) {
    const size_t have = buffered() < n ? buffered() : n;
    if (have) {
      memcpy(dst, rbuf.data() + rpos, have);
      rpos += have;
      dst += have;
      n -= have;
    }
    while (n) {
      const ssize_t r = recv(fd, dst, n, n >= (64 << 10) ? MSG_WAITALL : 0);
      if (r > 0) {
        dst += r;
        n -= (size_t)r;
        bytes_in += (uint64_t)r;
      } else if (r < 0 && errno == EINTR) {
        continue;
      } else {
        return false;
      }
    }
    return true;
  }
  bool skip(size_t n) {
    while (n) {
      if (!fill(1)) return false;
      const size_t k = buffered() < n ? buffered() : n;
      rpos += k;
      n -= k;
    }
    return true;
  }
  bool read_frame_header(FrameHeader &h) {
    if (!fill(9)) return false;
    const uint8_t *p = peek();
    h.length = (uint32_t)p[0] << 16 | (uint32_t)p[1] << 8 | p[2];
    h.type = p[3];
    h.flags = p[4];
    h.stream = be32(p + 5) & 0x7fffffffu;
    consume(9);
    return true;
  }

  bool writev_all(struct iovec *iov, int n) {
    while (n > 0) {
      struct msghdr m;
      memset(&m, 0, sizeof(m));
      m.msg_iov = iov;
      m.msg_iovlen = (size_t)(n > 512 ? 512 : n);
      ssize_t w = sendmsg(fd, &m, MSG_NOSIGNAL);
      if (w < 0) {
        if (errno == EINTR) continue;
        return false;
      }
      bytes_out += (uint64_t)w;
      while (n > 0 && (size_t)w >= iov->iov_len) {
        w -= (ssize_t)iov->iov_len;
        ++iov;
        --n;
      }
      if (n > 0 && w) {
        iov->iov_base = (uint8_t *)iov->iov_base + w;
        iov->iov_len -= (size_t)w;
      }
    }
    return true;
  }
  bool write_all(const uint8_t *p, size_t n) {
    struct iovec v;
    v.iov_base = (void *)p;
    v.iov_len = n;
    return writev_all(&v, 1);
  }

  // small control frames
  bool send_settings(const uint16_t *ids, const uint32_t *vals, int n) {
    uint8_t b[9 + 6 * 8];
    put_frame_header(b, (uint32_t)(6 * n), kSettings, 0, 0);
    for (int i = 0; i < n; ++i) {
      b[9 + 6 * i] = (uint8_t)(ids[i] >> 8);
      b[10 + 6 * i] = (uint8_t)ids[i];
      put_be32(b + 11 + 6 * i, vals[i]);
    }
    return write_all(b, (size_t)(9 + 6 * n));
  }
  bool send_settings_ack() {
    uint8_t b[9];
    put_frame_header(b, 0, kSettings, kFlagAck, 0);
    return write_all(b, 9);
  }
  bool send_window_update(uint32_t stream, uint32_t inc) {
    uint8_t b[13];
    put_frame_header(b, 4, kWindowUpdate, 0, stream);
    put_be32(b + 9, inc & 0x7fffffffu);
    return write_all(b, 13);
  }
  bool send_ping_ack(const uint8_t *opaque8) {
    uint8_t b[17];
    put_frame_header(b, 8, kPing, kFlagAck, 0);
    memcpy(b + 9, opaque8, 8);
    return write_all(b, 17);
  }
  bool send_rst(uint32_t stream, uint32_t code) {
    uint8_t b[13];
    put_frame_header(b, 4, kRstStream, 0, stream);
    put_be32(b + 9, code);
    return write_all(b, 13);
  }
  bool send_goaway(uint32_t last_stream, uint32_t code) {
    uint8_t b[17];
    put_frame_header(b, 8, kGoaway, 0, 0);
    put_be32(b + 9, last_stream);
    put_be32(b + 13, code);
    return write_all(b, 17);
  }
};

// Strips the padding / priority prefix of a HEADERS payload in place; false = malformed.
inline bool strip_headers_prefix(uint8_t flags, const uint8_t *&p, size_t &n) {
  size_t pad = 0;
  if (flags & kFlagPadded) {
    if (n < 1) return false;
    pad = p[0];
    ++p;
    --n;
  }
  if (flags & kFlagPriority) {
    if (n < 5) return false;
    p += 5;
    n -= 5;
  }
  if (pad > n) return false;
  n -= pad;
  return true;
}

}  // namespace gs_h2

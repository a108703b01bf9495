// http_client.h -- one-shot HTTP/1.1 requests for the native daemon, plain or over TLS
// (OpenSSL), standing where net/http + client-go's rest transport stand in the reference
// (pkg/kubelet/client/client.go:38-117, podmanager.go:32-53). One connection per request,
// "Connection: close", whole body read; Content-Length and chunked bodies.
#pragma once
#include <errno.h>
#include <netdb.h>
#include <openssl/err.h>
#include <openssl/pem.h>
#include <openssl/ssl.h>
#include <openssl/x509v3.h>
#include <stdio.h>
#include <string.h>
#include <strings.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <unistd.h>

#include <string>
#include <utility>
#include <vector>

namespace gsd {

struct HttpResponse {
  int status = 0;
  std::string body;
};

struct TlsOptions {
  std::string ca_file, cert_file, key_file;
  std::string ca_pem, cert_pem, key_pem;  // the same three inline (kubeconfig *-data fields, decoded)
  bool insecure = false;  // skip verification (the kubelet client always does, main.go:41)
};

struct Url {
  bool tls = false;
  std::string host;
  int port = 0;
};

inline bool parse_url(const std::string &u, Url &out) {
  size_t i;
  if (u.compare(0, 7, "http://") == 0) {
    out.tls = false;
    i = 7;
  } else if (u.compare(0, 8, "https://") == 0) {
    out.tls = true;
    i = 8;
  } else {
    return false;
  }
  std::string rest = u.substr(i);
  const size_t slash = rest.find('/');
  if (slash != std::string::npos) rest.resize(slash);
  out.port = out.tls ? 443 : 80;
  if (!rest.empty() && rest[0] == '[') {  // [v6]:port
    const size_t close = rest.find(']');
    if (close == std::string::npos) return false;
    out.host = rest.substr(1, close - 1);
    if (close + 1 < rest.size() && rest[close + 1] == ':') out.port = atoi(rest.c_str() + close + 2);
  } else {
    const size_t colon = rest.rfind(':');
    if (colon != std::string::npos) {
      out.host = rest.substr(0, colon);
      out.port = atoi(rest.c_str() + colon + 1);
    } else {
      out.host = rest;
    }
  }
  return !out.host.empty() && out.port > 0 && out.port < 65536;
}

inline std::string url_escape(const std::string &s) {  // url.QueryEscape-like for path/query parts
  static const char hexd[] = "0123456789ABCDEF";
  std::string o;
  for (unsigned char c : s) {
    if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == '-' || c == '_' ||
        c == '.' || c == '~') {
      o.push_back((char)c);
    } else {
      o.push_back('%');
      o.push_back(hexd[c >> 4]);
      o.push_back(hexd[c & 15]);
    }
  }
  return o;
}

namespace detail {

struct Stream {
  int fd = -1;
  SSL_CTX *ctx = nullptr;
  SSL *ssl = nullptr;
  ~Stream() {
    if (ssl) {
      SSL_shutdown(ssl);
      SSL_free(ssl);
    }
    if (ctx) SSL_CTX_free(ctx);
    if (fd >= 0) close(fd);
  }
  bool write_all(const char *p, size_t n) {
    while (n) {
      ssize_t w = ssl ? SSL_write(ssl, p, (int)(n > (1 << 30) ? (1 << 30) : n)) : send(fd, p, n, MSG_NOSIGNAL);
      if (w <= 0) {
        if (!ssl && errno == EINTR) continue;
        return false;
      }
      p += w;
      n -= (size_t)w;
    }
    return true;
  }
  // >0 bytes, 0 = orderly end, <0 = error/timeout
  ssize_t read_some(char *p, size_t n) {
    for (;;) {
      if (ssl) {
        const int r = SSL_read(ssl, p, (int)n);
        if (r > 0) return r;
        const int e = SSL_get_error(ssl, r);
        if (e == SSL_ERROR_ZERO_RETURN) return 0;
        // a peer that closes without close_notify after a complete body: treat as the end
        if (e == SSL_ERROR_SYSCALL && errno == 0) return 0;
        if (e == SSL_ERROR_SSL) {
          const unsigned long q = ERR_peek_last_error();
          if (ERR_GET_REASON(q) == SSL_R_UNEXPECTED_EOF_WHILE_READING) return 0;
        }
        return -1;
      }
      const ssize_t r = recv(fd, p, n, 0);
      if (r < 0 && errno == EINTR) continue;
      return r;
    }
  }
};

inline std::string ssl_error() {
  char buf[256];
  const unsigned long e = ERR_get_error();
  if (!e) return "TLS failure";
  ERR_error_string_n(e, buf, sizeof(buf));
  return buf;
}

inline bool dechunk(const std::string &in, std::string &out) {
  size_t i = 0;
  for (;;) {
    const size_t eol = in.find("\r\n", i);
    if (eol == std::string::npos) return false;
    const unsigned long n = strtoul(in.c_str() + i, nullptr, 16);
    i = eol + 2;
    if (n == 0) return true;
    if (i + n > in.size()) return false;
    out.append(in, i, n);
    i += n + 2;
  }
}

}  // namespace detail

// false + err: the request did not complete (what net/http returns as an error). Any HTTP
// status is a completed request.
inline bool http_request(const std::string &base_url, const std::string &method, const std::string &path,
                         const std::vector<std::pair<std::string, std::string>> &headers,
                         const std::string &body, const TlsOptions &tls, double timeout_s, HttpResponse &out,
                         std::string &err) {
  Url u;
  if (!parse_url(base_url, u)) {
    err = "bad URL " + base_url;
    return false;
  }
  struct addrinfo hints, *res = nullptr;
  memset(&hints, 0, sizeof(hints));
  hints.ai_socktype = SOCK_STREAM;
  char port[16];
  snprintf(port, sizeof(port), "%d", u.port);
  const int gai = getaddrinfo(u.host.c_str(), port, &hints, &res);
  if (gai != 0) {
    err = "dial tcp: lookup " + u.host + ": " + gai_strerror(gai);
    return false;
  }
  detail::Stream s;
  std::string last = "no address";
  struct timeval tv;
  tv.tv_sec = (long)timeout_s;
  tv.tv_usec = (long)((timeout_s - (double)tv.tv_sec) * 1e6);
  if (tv.tv_sec == 0 && tv.tv_usec == 0) tv.tv_sec = 10;
  for (struct addrinfo *a = res; a; a = a->ai_next) {
    const int fd = socket(a->ai_family, a->ai_socktype | SOCK_CLOEXEC, a->ai_protocol);
    if (fd < 0) continue;
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof(tv));
    if (connect(fd, a->ai_addr, a->ai_addrlen) == 0) {
      s.fd = fd;
      break;
    }
    last = std::string("dial tcp ") + u.host + ":" + port + ": " + strerror(errno);
    close(fd);
  }
  freeaddrinfo(res);
  if (s.fd < 0) {
    err = last;
    return false;
  }
  if (u.tls) {
    s.ctx = SSL_CTX_new(TLS_client_method());
    if (!s.ctx) {
      err = detail::ssl_error();
      return false;
    }
    // Verification is off only when the configuration says so (insecure-skip-tls-verify, or the
    // kubelet client, main.go:41). A kubeconfig that names no CA is verified against the system's
    // roots, as client-go does -- never silently left unauthenticated with a bearer token on it.
    const bool verify = !tls.insecure;
    if (verify) {
      if (tls.ca_file.empty() && tls.ca_pem.empty() && SSL_CTX_set_default_verify_paths(s.ctx) != 1) {
        err = "unable to load the system's CA certificates: " + detail::ssl_error();
        return false;
      }
      if (!tls.ca_file.empty() && SSL_CTX_load_verify_locations(s.ctx, tls.ca_file.c_str(), nullptr) != 1) {
        err = "unable to load CA file " + tls.ca_file + ": " + detail::ssl_error();
        return false;
      }
      if (!tls.ca_pem.empty()) {
        BIO *bio = BIO_new_mem_buf(tls.ca_pem.data(), (int)tls.ca_pem.size());
        int added = 0;
        while (X509 *x = PEM_read_bio_X509(bio, nullptr, nullptr, nullptr)) {
          added += X509_STORE_add_cert(SSL_CTX_get_cert_store(s.ctx), x) == 1;
          X509_free(x);
        }
        BIO_free(bio);
        ERR_clear_error();  // the read that found the end of the bundle
        if (!added) {
          err = "certificate-authority-data holds no certificate";
          return false;
        }
      }
      SSL_CTX_set_verify(s.ctx, SSL_VERIFY_PEER, nullptr);
    } else {
      SSL_CTX_set_verify(s.ctx, SSL_VERIFY_NONE, nullptr);
    }
    if (!tls.cert_file.empty() && !tls.key_file.empty()) {
      if (SSL_CTX_use_certificate_chain_file(s.ctx, tls.cert_file.c_str()) != 1 ||
          SSL_CTX_use_PrivateKey_file(s.ctx, tls.key_file.c_str(), SSL_FILETYPE_PEM) != 1) {
        err = "unable to load client certificate: " + detail::ssl_error();
        return false;
      }
    } else if (!tls.cert_pem.empty() && !tls.key_pem.empty()) {
      BIO *cb = BIO_new_mem_buf(tls.cert_pem.data(), (int)tls.cert_pem.size());
      BIO *kb = BIO_new_mem_buf(tls.key_pem.data(), (int)tls.key_pem.size());
      X509 *x = PEM_read_bio_X509(cb, nullptr, nullptr, nullptr);
      EVP_PKEY *k = PEM_read_bio_PrivateKey(kb, nullptr, nullptr, nullptr);
      const bool ok = x && k && SSL_CTX_use_certificate(s.ctx, x) == 1 && SSL_CTX_use_PrivateKey(s.ctx, k) == 1;
      if (x) X509_free(x);
      if (k) EVP_PKEY_free(k);
      BIO_free(cb);
      BIO_free(kb);
      if (!ok) {
        err = "unable to load client-certificate-data / client-key-data: " + detail::ssl_error();
        return false;
      }
    }
    s.ssl = SSL_new(s.ctx);
    if (!s.ssl) {
      err = detail::ssl_error();
      return false;
    }
    SSL_set_fd(s.ssl, s.fd);
    SSL_set_tlsext_host_name(s.ssl, u.host.c_str());
    if (verify) {
      X509_VERIFY_PARAM *vp = SSL_get0_param(s.ssl);
      if (X509_VERIFY_PARAM_set1_ip_asc(vp, u.host.c_str()) != 1) SSL_set1_host(s.ssl, u.host.c_str());
    }
    if (SSL_connect(s.ssl) != 1) {
      err = "tls: " + detail::ssl_error();
      return false;
    }
  }
  std::string req = method + " " + path + " HTTP/1.1\r\nHost: " + u.host + ":" + port +
                    "\r\nConnection: close\r\nUser-Agent: gpushare-device-plugin-v2\r\n";
  for (const auto &h : headers) req += h.first + ": " + h.second + "\r\n";
  if (!body.empty() || method == "PATCH" || method == "POST" || method == "PUT") {
    char cl[48];
    snprintf(cl, sizeof(cl), "Content-Length: %zu\r\n", body.size());
    req += cl;
  }
  req += "\r\n";
  req += body;
  if (!s.write_all(req.data(), req.size())) {
    err = std::string("write: ") + (u.tls ? detail::ssl_error() : strerror(errno));
    return false;
  }
  std::string raw;
  char buf[65536];
  size_t header_end = std::string::npos;
  long content_length = -1;
  bool chunked = false;
  for (;;) {
    if (header_end != std::string::npos && content_length >= 0 &&
        raw.size() >= header_end + 4 + (size_t)content_length)
      break;
    if (header_end != std::string::npos && chunked && raw.size() >= header_end + 9 &&
        raw.compare(raw.size() - 5, 5, "0\r\n\r\n") == 0) {
      std::string probe;  // looks like the last chunk: believe it only if the framing agrees
      if (detail::dechunk(raw.substr(header_end + 4), probe)) break;
    }
    const ssize_t r = s.read_some(buf, sizeof(buf));
    if (r < 0) {
      err = "read: timeout or connection error";
      return false;
    }
    if (r == 0) break;
    raw.append(buf, (size_t)r);
    if (raw.size() > (512u << 20)) {
      err = "response too large";
      return false;
    }
    if (header_end == std::string::npos) {
      header_end = raw.find("\r\n\r\n");
      if (header_end != std::string::npos) {
        size_t i = raw.find("\r\n") + 2;
        while (i < header_end) {
          const size_t eol = raw.find("\r\n", i);
          const std::string line = raw.substr(i, eol - i);
          if (strncasecmp(line.c_str(), "content-length:", 15) == 0) content_length = atol(line.c_str() + 15);
          if (strncasecmp(line.c_str(), "transfer-encoding:", 18) == 0 &&
              strcasestr(line.c_str() + 18, "chunked"))
            chunked = true;
          i = eol + 2;
        }
      }
    }
  }
  if (header_end == std::string::npos || raw.compare(0, 5, "HTTP/") != 0) {
    err = "malformed HTTP response";
    return false;
  }
  const size_t sp = raw.find(' ');
  out.status = sp == std::string::npos ? 0 : atoi(raw.c_str() + sp + 1);
  const std::string payload = raw.substr(header_end + 4);
  out.body.clear();
  if (chunked) {
    if (!detail::dechunk(payload, out.body)) {
      err = "malformed chunked body";
      return false;
    }
  } else if (content_length >= 0) {
    if (payload.size() < (size_t)content_length) {
      err = "unexpected EOF";
      return false;
    }
    out.body = payload.substr(0, (size_t)content_length);
  } else {
    out.body = payload;
  }
  return true;
}

}  // namespace gsd

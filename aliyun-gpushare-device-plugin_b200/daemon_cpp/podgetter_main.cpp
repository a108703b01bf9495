// podgetter -- print what kubelet's /pods answers (cmd/podgetter/main.go:19-57), natively.
// A debugging aid: the reference's four flags, kubelet at 127.0.0.1:10250 (never verified:
// Insecure: true there), the service-account token when no credentials are given. Prints the
// body as kubelet sent it (the reference prints Go's %v of the decoded list).
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>

#include <string>

#include "k8s.h"

using namespace gsd;

int main(int argc, char **argv) {
  signal(SIGPIPE, SIG_IGN);
  std::string cert, key, token, url = "https://127.0.0.1:10250";
  long timeout = 10;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    if (a.size() < 2 || a[0] != '-') break;
    a = a.substr(a[1] == '-' ? 2 : 1);
    std::string val;
    const size_t eq = a.find('=');
    if (eq != std::string::npos) {
      val = a.substr(eq + 1);
      a.resize(eq);
    } else if (i + 1 < argc) {
      val = argv[++i];
    } else {
      fprintf(stderr, "flag needs an argument: -%s\n", a.c_str());
      return 2;
    }
    if (a == "client-cert") cert = val;
    else if (a == "client-key") key = val;
    else if (a == "token") token = val;
    else if (a == "timeout") timeout = atol(val.c_str());
    else if (a == "kubelet-url") url = val;  // addition: where kubelet listens
    else {
      fprintf(stderr, "flag provided but not defined: -%s\n", a.c_str());
      return 2;
    }
  }
  if (cert.empty() && key.empty() && token.empty()) {
    if (!read_file(std::string(kSaDir) + "/token", token)) {
      fprintf(stderr, "in cluster mode, find token failed, error: open %s/token: no such file or directory\n", kSaDir);
      return 2;  // the reference panics here
    }
  }
  Transport t;
  t.kubelet_url = url;
  t.kubelet_token = token;
  t.kubelet_tls.insecure = true;
  t.kubelet_tls.cert_file = cert;
  t.kubelet_tls.key_file = key;
  t.timeout_s = (double)timeout;
  std::vector<std::pair<std::string, std::string>> h;
  if (!token.empty()) h.push_back({"Authorization", "Bearer " + token});
  HttpResponse r;
  std::string err;
  if (!http_request(url, "GET", "/pods/", h, "", t.kubelet_tls, t.timeout_s, r, err)) {
    printf("%s\n", err.c_str());
    return 0;  // the reference prints the error and returns
  }
  fwrite(r.body.data(), 1, r.body.size(), stdout);
  if (r.body.empty() || r.body.back() != '\n') putchar('\n');
  return 0;
}

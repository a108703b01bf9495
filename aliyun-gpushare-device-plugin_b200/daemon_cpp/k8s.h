// k8s.h -- the Kubernetes I/O around Allocate for the native daemon,
// with the reference's retry ladders and error behaviour:
//   kubeInit                         pkg/gpu/nvidia/podmanager.go:32-60
//   getPodList                       podmanager.go:187-208  (kubelet /pods, Pending only)
//   getPodListsByQueryKubelet        podmanager.go:210-225  (8 retries x 100 ms, then apiserver)
//   getPodListsByListAPIServer       podmanager.go:227-245  (3 retries x 1 s)
//   KubeletClient.GetNodeRunningPods pkg/kubelet/client/client.go:119-134
//   the PATCH with one retry         allocate.go:132-152
//   disableCGPUIsolationOrNot        podmanager.go:62-75
//   patchGPUCount                    podmanager.go:160-185
// The same logic lives in Python in k8s_io.py; tests drive both against one fake cluster.
#pragma once
#include <stdlib.h>
#include <time.h>

#include <fstream>
#include <functional>
#include <sstream>
#include <string>
#include <vector>

#include "gpushare.h"
#include "http_client.h"
#include "json.h"

namespace gsd {

static const char kOptimisticLockErrorMsg[] =
    "the object has been modified; please apply your changes to the latest version and try again";  // const.go:15
static const char kSaDir[] = "/var/run/secrets/kubernetes.io/serviceaccount";
constexpr int kRetries = 8;  // podmanager.go:29

inline bool read_file(const std::string &path, std::string &out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::ostringstream ss;
  ss << f.rdbuf();
  out = ss.str();
  return true;
}
inline std::string trim(std::string s) {
  while (!s.empty() && (s.back() == '\n' || s.back() == '\r' || s.back() == ' ' || s.back() == '\t')) s.pop_back();
  size_t i = 0;
  while (i < s.size() && (s[i] == ' ' || s[i] == '\t')) ++i;
  return s.substr(i);
}

inline std::string base64_decode(const std::string &in) {
  std::string out;
  uint32_t acc = 0;
  int bits = 0;
  for (unsigned char ch : in) {
    int v;
    if (ch >= 'A' && ch <= 'Z') v = ch - 'A';
    else if (ch >= 'a' && ch <= 'z') v = ch - 'a' + 26;
    else if (ch >= '0' && ch <= '9') v = ch - '0' + 52;
    else if (ch == '+' || ch == '-') v = 62;
    else if (ch == '/' || ch == '_') v = 63;
    else continue;  // padding, whitespace
    acc = acc << 6 | (uint32_t)v;
    bits += 6;
    if (bits >= 8) {
      bits -= 8;
      out.push_back((char)(acc >> bits & 0xff));
    }
  }
  return out;
}

// Where the apiserver is and how to authenticate (kubeInit, podmanager.go:32-53).
struct ApiConfig {
  std::string host, token;
  TlsOptions tls;
  bool valid = false;

  static bool in_cluster(ApiConfig &c, std::string &err) {
    const char *h = getenv("KUBERNETES_SERVICE_HOST"), *p = getenv("KUBERNETES_SERVICE_PORT");
    if (!h || !*h || !p || !*p) {
      err = "unable to load in-cluster configuration, KUBERNETES_SERVICE_HOST and KUBERNETES_SERVICE_PORT "
            "must be defined";
      return false;
    }
    std::string tok;
    if (!read_file(std::string(kSaDir) + "/token", tok)) {
      err = std::string("open ") + kSaDir + "/token: no such file or directory";
      return false;
    }
    std::string host = h;
    if (host.find(':') != std::string::npos) host = "[" + host + "]";
    c.host = "https://" + host + ":" + p;
    c.token = trim(tok);
    c.tls.ca_file = std::string(kSaDir) + "/ca.crt";
    c.valid = true;
    return true;
  }

  // A kubeconfig with one cluster, one user, one context (what `kubectl config view --minify
  // --raw` prints, and what kubeadm writes as kubelet.conf, the file the reference DaemonSet
  // points KUBECONFIG at, device-plugin-ds.yaml:41-47): the keys are taken where they first
  // appear, block style. Inline *-data fields (base64 PEM) are read as well as the file forms.
  static bool from_kubeconfig(const std::string &path, ApiConfig &c, std::string &err) {
    std::string text;
    if (!read_file(path, text)) {
      err = "cannot read " + path;
      return false;
    }
    auto first = [&](const char *key) -> std::string {
      std::istringstream in(text);
      std::string line;
      const std::string k = std::string(key) + ":";
      while (std::getline(in, line)) {
        std::string t = trim(line);
        if (t.compare(0, 2, "- ") == 0) t = trim(t.substr(2));
        if (t.compare(0, k.size(), k) == 0) {
          std::string v = trim(t.substr(k.size()));
          if (v.size() >= 2 && ((v.front() == '"' && v.back() == '"') || (v.front() == '\'' && v.back() == '\'')))
            v = v.substr(1, v.size() - 2);
          return v;
        }
      }
      return "";
    };
    c.host = first("server");
    while (!c.host.empty() && c.host.back() == '/') c.host.pop_back();
    c.token = first("token");
    c.tls.ca_file = first("certificate-authority");
    c.tls.cert_file = first("client-certificate");
    c.tls.key_file = first("client-key");
    c.tls.insecure = first("insecure-skip-tls-verify") == "true";
    c.tls.ca_pem = base64_decode(first("certificate-authority-data"));
    c.tls.cert_pem = base64_decode(first("client-certificate-data"));
    c.tls.key_pem = base64_decode(first("client-key-data"));
    if (c.host.empty()) {
      err = "no cluster server in " + path;
      return false;
    }
    c.valid = true;
    return true;
  }

  static bool load(ApiConfig &c, std::string &err) {  // KUBECONFIG if that file exists
    const char *kc = getenv("KUBECONFIG");
    std::string probe;
    if (kc && *kc && read_file(kc, probe)) return from_kubeconfig(kc, c, err);
    return in_cluster(c, err);
  }
};

// One pod as gs_allocate wants it, owning its strings (api.Pod.from_k8s).
struct Pod {
  std::string uid, ns, name, node;
  uint64_t limit = 0;
  bool has_idx = false, has_time = false, has_assigned = false;
  std::string idx, time, assigned;
};

// Limits go through resource.Quantity.Value() and are summed with Go's wrapping uint add
// (podutils.go:122-131); false when a quantity does not parse (the Go decode of the PodList
// would have failed).
inline bool pod_from_json(const JVal &p, Pod &out, std::string &err) {
  auto str_at = [&](std::initializer_list<const char *> path) -> std::string {
    const JVal *v = p.at(path);
    return v && v->is_str() ? v->s : std::string();
  };
  out.uid = str_at({"metadata", "uid"});
  out.ns = str_at({"metadata", "namespace"});
  out.name = str_at({"metadata", "name"});
  out.node = str_at({"spec", "nodeName"});
  const JVal *ann = p.at({"metadata", "annotations"});
  auto ann_get = [&](const char *k, bool &has, std::string &dst) {
    const JVal *v = ann ? ann->get(k) : nullptr;
    has = v && v->is_str();
    if (has) dst = v->s;
  };
  ann_get(GS_ENV_RESOURCE_INDEX, out.has_idx, out.idx);
  ann_get(GS_ENV_ASSUME_TIME, out.has_time, out.time);
  ann_get(GS_ENV_ASSIGNED_FLAG, out.has_assigned, out.assigned);
  out.limit = 0;
  const JVal *cs = p.at({"spec", "containers"});
  if (cs && cs->type == JVal::Arr) {
    for (const JVal &c : cs->arr) {
      const JVal *q = c.at({"resources", "limits", GS_RESOURCE_NAME});
      if (!q) continue;
      const std::string t = q->text();
      int64_t v = 0;
      if (gs_quantity_value(t.data(), t.size(), &v) != GS_OK) {
        err = "quantities must match the regular expression: bad " GS_RESOURCE_NAME " limit \"" + t + "\"";
        return false;
      }
      out.limit += (uint64_t)v;
    }
  }
  return true;
}

inline void pod_views(const std::vector<Pod> &pods, std::vector<gs_pod_view> &views) {
  views.resize(pods.size());
  for (size_t i = 0; i < pods.size(); ++i) {
    const Pod &p = pods[i];
    gs_pod_view &v = views[i];
    v.uid = gs_str{p.uid.data(), p.uid.size()};
    v.ns = gs_str{p.ns.data(), p.ns.size()};
    v.name = gs_str{p.name.data(), p.name.size()};
    v.node_name = gs_str{p.node.data(), p.node.size()};
    v.gpu_mem_limit = p.limit;
    v.ann_idx = p.has_idx ? gs_str{p.idx.data(), p.idx.size()} : gs_str{nullptr, 0};
    v.ann_assume_time = p.has_time ? gs_str{p.time.data(), p.time.size()} : gs_str{nullptr, 0};
    v.ann_assigned = p.has_assigned ? gs_str{p.assigned.data(), p.assigned.size()} : gs_str{nullptr, 0};
  }
}

// The four network calls of the hot path plus the two node calls of start-up.
struct Transport {
  ApiConfig api;
  std::string kubelet_url, kubelet_token;
  TlsOptions kubelet_tls;  // insecure always (main.go:41)
  double timeout_s = 10.0;

  // err: client-go's err.Error() -- for a Status object its .message, which is what
  // allocate.go:140 compares with the optimistic-lock text
  bool api_call(const std::string &method, const std::string &path, const std::string &body,
                const char *content_type, JVal &out, std::string &err) const {
    if (!api.valid) {
      err = "no apiserver configured";
      return false;
    }
    std::vector<std::pair<std::string, std::string>> h = {{"Accept", "application/json"}};
    if (!api.token.empty()) h.push_back({"Authorization", "Bearer " + api.token});
    if (content_type) h.push_back({"Content-Type", content_type});
    HttpResponse r;
    if (!http_request(api.host, method, path, h, body, api.tls, timeout_s, r, err)) return false;
    JVal doc;
    const bool parsed = r.body.empty() ? true : json_parse(r.body, doc);
    if (r.status >= 400) {
      const JVal *m = parsed ? doc.get("message") : nullptr;
      err = m && m->is_str() && !m->s.empty() ? m->s : r.body;
      return false;
    }
    if (!parsed) {
      err = "invalid JSON from the apiserver";
      return false;
    }
    out = std::move(doc);
    return true;
  }
  // client.go:119-134: GET https://host:port/pods/, body decoded whatever the status
  bool kubelet_pods(JVal &out, std::string &err) const {
    std::vector<std::pair<std::string, std::string>> h;
    if (!kubelet_token.empty()) h.push_back({"Authorization", "Bearer " + kubelet_token});
    HttpResponse r;
    if (!http_request(kubelet_url, "GET", "/pods/", h, "", kubelet_tls, timeout_s, r, err)) return false;
    if (!json_parse(r.body, out)) {
      err = "invalid /pods/ body";
      return false;
    }
    return true;
  }
  bool api_list(const std::string &field_selector, JVal &out, std::string &err) const {
    return api_call("GET", "/api/v1/pods?fieldSelector=" + url_escape(field_selector), "", nullptr, out, err);
  }
  bool api_patch(const std::string &ns, const std::string &name, const std::string &body, std::string &err) const {
    JVal ignored;
    return api_call("PATCH", "/api/v1/namespaces/" + url_escape(ns) + "/pods/" + url_escape(name), body,
                    "application/strategic-merge-patch+json", ignored, err);
  }
  bool get_node(const std::string &name, JVal &out, std::string &err) const {
    return api_call("GET", "/api/v1/nodes/" + url_escape(name), "", nullptr, out, err);
  }
  bool patch_node_status(const std::string &name, const std::string &body, std::string &err) const {
    JVal ignored;
    return api_call("PATCH", "/api/v1/nodes/" + url_escape(name) + "/status", body,
                    "application/strategic-merge-patch+json", ignored, err);
  }
};

inline void sleep_s(double s) {
  struct timespec ts;
  ts.tv_sec = (time_t)s;
  ts.tv_nsec = (long)((s - (double)ts.tv_sec) * 1e9);
  nanosleep(&ts, nullptr);
}

// getPendingPodsInNode's input side: the pod list through the reference's ladders.
struct PodLister {
  const Transport *t = nullptr;
  std::string node_name;
  bool query_kubelet = false;
  std::function<void(double)> sleep = sleep_s;

  bool convert(const JVal *items, bool pending_only, std::vector<Pod> &out, std::string &err) const {
    out.clear();
    if (!items || items->type != JVal::Arr) return true;
    for (const JVal &p : items->arr) {
      if (pending_only) {
        const JVal *ph = p.at({"status", "phase"});
        if (!ph || !ph->is_str() || ph->s != "Pending") continue;
      }
      out.emplace_back();
      if (!pod_from_json(p, out.back(), err)) return false;
    }
    return true;
  }
  bool get_pod_list(std::vector<Pod> &out, std::string &err) const {  // podmanager.go:187-208
    JVal doc;
    if (!t->kubelet_pods(doc, err)) return false;
    if (!convert(doc.get("items"), true, out, err)) return false;
    if (out.empty()) {
      err = "not found pending pod";
      return false;
    }
    return true;
  }
  bool by_apiserver(std::vector<Pod> &out, std::string &err) const {  // podmanager.go:227-245
    const std::string selector = "spec.nodeName=" + node_name + ",status.phase=Pending";
    JVal doc;
    std::string e;
    bool ok = t->api_list(selector, doc, e);
    for (int i = 0; i < 3 && !ok; ++i) {
      ok = t->api_list(selector, doc, e);
      sleep(1.0);
    }
    if (!ok) {
      err = "failed to get Pods assigned to node " + node_name;
      return false;
    }
    return convert(doc.get("items"), false, out, err);
  }
  bool by_kubelet(std::vector<Pod> &out, std::string &err) const {  // podmanager.go:210-225
    std::string e;
    bool ok = get_pod_list(out, e);
    for (int i = 0; i < kRetries && !ok; ++i) {
      ok = get_pod_list(out, e);
      sleep(0.1);
    }
    if (!ok) return by_apiserver(out, err);
    return true;
  }
  bool list(std::vector<Pod> &out, std::string &err) const {
    return query_kubelet ? by_kubelet(out, err) : by_apiserver(out, err);
  }
};

// allocate.go:136-152: one retry, only when the error text IS the optimistic-lock message
inline bool patch_pod_assigned(const Transport &t, const std::string &ns, const std::string &name,
                               const std::string &body, std::string &err) {
  if (t.api_patch(ns, name, body, err)) return true;
  if (err == kOptimisticLockErrorMsg) return t.api_patch(ns, name, body, err);
  return false;
}

// podmanager.go:62-75
inline bool disable_cgpu_isolation_or_not(const Transport &t, const std::string &node, bool &disable,
                                          std::string &err) {
  JVal doc;
  if (!t.get_node(node, doc, err)) return false;
  const JVal *v = doc.at({"metadata", "labels", GS_NODE_LABEL_DISABLE_CGPU});
  disable = v && v->is_str() && v->s == "true";
  return true;
}

// podmanager.go:160-185: set status.capacity/allocatable["aliyun.com/gpu-count"] unless
// capacity already says so. The reference builds a two-way strategic-merge patch of the whole
// node; its only difference is these two keys, so the patch is exactly this body.
inline bool patch_gpu_count(const Transport &t, const std::string &node, int gpu_count, std::string &err) {
  JVal doc;
  if (!t.get_node(node, doc, err)) return false;
  const JVal *cap = doc.at({"status", "capacity", GS_RESOURCE_COUNT});
  if (cap) {
    const std::string s = cap->text();
    int64_t v = 0;
    if (gs_quantity_value(s.data(), s.size(), &v) == GS_OK && v == gpu_count) return true;
  }
  char body[256];
  snprintf(body, sizeof(body),
           "{\"status\":{\"allocatable\":{\"" GS_RESOURCE_COUNT "\":\"%d\"},\"capacity\":{\"" GS_RESOURCE_COUNT
           "\":\"%d\"}}}",
           gpu_count, gpu_count);
  return t.patch_node_status(node, body, err);
}

}  // namespace gsd

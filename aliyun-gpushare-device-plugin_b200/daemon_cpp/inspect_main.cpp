// kubectl-inspect-gpushare-v2 -- per-node / per-GPU allocation report as a native binary
// (SURVEY 8(f) row N3). Restates cmd/inspect of the reference: main.go:33-79 (flags: -d,
// optional node name), nodeinfo.go:47-272 (node/device bookkeeping, the allocation
// annotation), podinfo.go:58-135 (pod listing with 5 retries x 100 ms, active-pod filter),
// display.go:15-255 (the two tables, laid out by Go's text/tabwriter with padding 2).
// The same program exists in Python (inspect_gpushare.py); tests compare the two outputs
// byte for byte on generated clusters.
//
// Deliberate differences from the reference, as in the Python version: the MiB/GiB guess flips
// at 1024 units per GPU instead of 100 (a 179-GiB B200 is not "MiB"), --unit overrides it;
// nodes keep the apiserver's order and devices go by index with pending (-1) last, where Go
// iterates maps in random order.
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <algorithm>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "gpushare.h"
#include "k8s.h"

using namespace gsd;

namespace {

const char kAllocationFlag[] = "scheduler.framework.gpushare.allocation";  // main.go:25
constexpr int kListRetries = 5;                                            // podinfo.go:25

// ---- text/tabwriter: NewWriter(out, 0, 0, 2, ' ', 0) -------------------------------------------
size_t runes(const std::string &s) {
  size_t n = 0;
  for (unsigned char c : s) n += (c & 0xc0) != 0x80;
  return n;
}

struct TabWriter {
  std::vector<std::vector<std::string>> lines;
  std::vector<std::string> out;
  size_t padding = 2;

  void write_lines(const std::vector<size_t> &widths, size_t a, size_t b) {
    for (size_t i = a; i < b; ++i) {
      const auto &cells = lines[i];
      std::string s;
      for (size_t j = 0; j < cells.size(); ++j) {
        s += cells[j];
        if (j < widths.size() && j + 1 < cells.size()) {
          const size_t w = runes(cells[j]);
          if (widths[j] > w) s.append(widths[j] - w, ' ');
        }
      }
      out[i] = s;
    }
  }
  // column blocks: consecutive lines that have a cell in this column share its width
  void format(std::vector<size_t> widths, size_t line0, size_t line1) {
    const size_t column = widths.size();
    size_t cur = line0;
    while (cur < line1) {
      if (column + 1 >= lines[cur].size()) {
        ++cur;
        continue;
      }
      write_lines(widths, line0, cur);
      line0 = cur;
      size_t width = 0;
      while (cur < line1 && column + 1 < lines[cur].size()) {
        width = std::max(width, runes(lines[cur][column]) + padding);
        ++cur;
      }
      std::vector<size_t> deeper = widths;
      deeper.push_back(width);
      format(deeper, line0, cur);
      line0 = cur;
    }
    write_lines(widths, line0, line1);
  }
  std::string run(const std::string &text) {
    lines.clear();
    size_t start = 0;
    for (;;) {
      const size_t nl = text.find('\n', start);
      const std::string line = text.substr(start, nl == std::string::npos ? std::string::npos : nl - start);
      std::vector<std::string> cells;
      size_t c0 = 0;
      for (;;) {
        const size_t tab = line.find('\t', c0);
        cells.push_back(line.substr(c0, tab == std::string::npos ? std::string::npos : tab - c0));
        if (tab == std::string::npos) break;
        c0 = tab + 1;
      }
      lines.push_back(std::move(cells));
      if (nl == std::string::npos) break;
      start = nl + 1;
    }
    out.assign(lines.size(), "");
    format({}, 0, lines.size());
    std::string joined;
    for (size_t i = 0; i < out.size(); ++i) {
      if (i) joined.push_back('\n');
      joined += out[i];
    }
    return joined;
  }
};

// ---- model ---------------------------------------------------------------------------------------
[[noreturn]] void die(const std::string &msg) {
  fprintf(stderr, "%s\n", msg.c_str());
  exit(1);
}

int64_t quantity(const JVal &q) {
  const std::string t = q.text();
  int64_t v = 0;
  if (gs_quantity_value(t.data(), t.size(), &v) != GS_OK) die("unable to parse quantity \"" + t + "\"");
  return v;
}

int64_t gpu_memory_in_pod(const JVal &pod) {  // podinfo.go:125-135 / display.go:247-255
  int64_t total = 0;
  const JVal *cs = pod.at({"spec", "containers"});
  if (cs && cs->type == JVal::Arr)
    for (const JVal &c : cs->arr) {
      const JVal *q = c.at({"resources", "limits", GS_RESOURCE_NAME});
      if (q) total += quantity(*q);
    }
  return total;
}

using Allocation = std::vector<std::pair<int64_t, int64_t>>;  // gpu index -> memory, first-seen order

bool whole_number(const JVal &v, int64_t &out) {
  if (v.type != JVal::Num) return false;
  char *end = nullptr;
  out = strtoll(v.s.c_str(), &end, 10);
  return end && *end == 0 && !v.s.empty();
}

// nodeinfo.go:245-272: {gpu index: memory} from the scheduler-framework annotation, a JSON
// map[container index]map[gpu id]memory; anything else reads as "no allocation"
Allocation get_allocation(const JVal &pod) {
  Allocation out;
  const JVal *a = pod.at({"metadata", "annotations", kAllocationFlag});
  if (!a || !a->is_str()) return out;
  JVal doc;
  if (!json_parse(a->s, doc) || doc.type != JVal::Obj) return {};
  for (const auto &container : doc.obj) {
    int64_t ignored;
    if (gs_go_atoi(container.first.data(), container.first.size(), &ignored) != GS_OK) return {};
    if (container.second.type != JVal::Obj) return {};
    for (const auto &kv : container.second.obj) {
      int64_t gid, mem;
      if (gs_go_atoi(kv.first.data(), kv.first.size(), &gid) != GS_OK || !whole_number(kv.second, mem)) return {};
      auto it = std::find_if(out.begin(), out.end(), [&](const std::pair<int64_t, int64_t> &p) { return p.first == gid; });
      if (it == out.end()) out.emplace_back(gid, mem);
      else it->second += mem;
    }
  }
  return out;
}

struct DeviceInfo {
  int64_t idx = 0, used = 0, total = 0;
  std::vector<const JVal *> pods;
  std::string str() const {  // nodeinfo.go:23-28
    char b[64];
    if (idx == -1) snprintf(b, sizeof(b), "%lld", (long long)used);
    else snprintf(b, sizeof(b), "%lld/%lld", (long long)used, (long long)total);
    return b;
  }
};

struct NodeInfo {
  const JVal *node = nullptr;
  std::vector<const JVal *> pods;
  int64_t gpu_count = 0, gpu_total = 0;
  std::map<int64_t, DeviceInfo> devs;

  explicit NodeInfo(const JVal *n) : node(n) {
    const JVal *c = n->at({"status", "allocatable", GS_RESOURCE_COUNT});
    const JVal *m = n->at({"status", "allocatable", GS_RESOURCE_NAME});
    gpu_count = c ? quantity(*c) : 0;
    gpu_total = m ? quantity(*m) : 0;
    const int64_t per = gpu_count ? gpu_total / gpu_count : 0;
    for (int64_t i = 0; i < gpu_count; ++i) {
      devs[i].idx = i;
      devs[i].total = per;
    }
  }
  std::string name() const {
    const JVal *v = node->at({"metadata", "name"});
    return v && v->is_str() ? v->s : "";
  }
  bool has_pending() const { return devs.count(-1) != 0; }
  Allocation device_info(const JVal &pod) const {  // nodeinfo.go:169-197
    Allocation a = get_allocation(pod);
    if (!a.empty()) return a;
    int64_t dev = -1;
    const JVal *idx = pod.at({"metadata", "annotations", GS_ENV_RESOURCE_INDEX});
    if (idx && idx->is_str() && gs_go_atoi(idx->s.data(), idx->s.size(), &dev) != GS_OK) dev = -1;
    return {{dev, gpu_memory_in_pod(pod)}};
  }
  void build_device_info() {  // nodeinfo.go:143-167
    const int64_t per = gpu_count > 0 ? gpu_total / gpu_count : 0;
    for (const JVal *pod : pods) {
      if (gpu_memory_in_pod(*pod) <= 0) continue;
      for (const auto &kv : device_info(*pod)) {
        if (!devs.count(kv.first)) {
          devs[kv.first].idx = kv.first;
          devs[kv.first].total = per;
        }
        devs[kv.first].used += kv.second;
        devs[kv.first].pods.push_back(pod);
      }
    }
  }
  std::vector<int64_t> dev_order() const {  // by index, pending last
    std::vector<int64_t> o;
    for (const auto &kv : devs)
      if (kv.first != -1) o.push_back(kv.first);
    if (has_pending()) o.push_back(-1);
    return o;
  }
};

std::string address(const JVal &node) {
  const JVal *as = node.at({"status", "addresses"});
  if (as && as->type == JVal::Arr)
    for (const JVal &a : as->arr) {
      const JVal *t = a.get("type");
      if (t && t->is_str() && t->s == "InternalIP") {
        const JVal *v = a.get("address");
        return v && v->is_str() ? v->s : "unknown";
      }
    }
  return "unknown";
}

std::string str_field(const JVal &doc, std::initializer_list<const char *> path) {
  const JVal *v = doc.at(path);
  return v && v->is_str() ? v->s : "";
}

std::string num(long long v) {
  char b[32];
  snprintf(b, sizeof(b), "%lld", v);
  return b;
}

// ---- display ---------------------------------------------------------------------------------------
std::string display_summary(std::vector<NodeInfo> &infos, const std::string &unit) {  // display.go:141-245
  bool has_pending = false;
  int64_t max_gpus = 0;
  for (auto &i : infos) {
    has_pending = has_pending || i.has_pending();
    max_gpus = std::max(max_gpus, i.gpu_count);
  }
  std::string text = "NAME\tIPADDRESS\t";
  for (int64_t i = 0; i < max_gpus; ++i) text += "GPU" + num(i) + "(Allocated/Total)\t";
  if (has_pending) text += "PENDING(Allocated)\t";
  text += "GPU Memory(" + unit + ")\n";
  long long used_cluster = 0, total_cluster = 0;
  size_t line_len = 0;
  for (auto &info : infos) {
    if (info.gpu_total <= 0) continue;
    long long used = 0;
    std::string line = info.name() + "\t" + address(*info.node) + "\t";
    for (int64_t i = 0; i < max_gpus; ++i) {
      if (info.devs.count(i)) {
        line += info.devs[i].str() + "\t";
        used += info.devs[i].used;
      } else {
        line += "0/0\t";
      }
    }
    std::string pending;
    if (info.has_pending()) {
      pending = num(info.devs[-1].used);
      used += info.devs[-1].used;
    }
    if (has_pending) line += pending + "\t";
    line += num(used) + "/" + num(info.gpu_total) + "\n";
    text += line;
    if (line_len == 0) line_len = runes(line) + 20;
    used_cluster += used;
    total_cluster += info.gpu_total;
  }
  text += std::string(line_len, '-') + "\n";
  text += "Allocated/Total GPU Memory In Cluster:\n";
  const long long usage = total_cluster > 0 ? (long long)((double)used_cluster / (double)total_cluster * 100) : 0;
  text += num(used_cluster) + "/" + num(total_cluster) + " (" + num(usage) + "%)\t\n";
  return TabWriter().run(text);
}

std::string display_details(std::vector<NodeInfo> &infos) {  // display.go:15-129
  std::string text;
  long long used_cluster = 0, total_cluster = 0;
  size_t line_len = 0;
  for (auto &info : infos) {
    if (info.gpu_total <= 0) continue;
    text += "\nNAME:\t" + info.name() + "\nIPADDRESS:\t" + address(*info.node) + "\n\n";
    text += "NAME\tNAMESPACE\t";
    for (int64_t i = 0; i < info.gpu_count; ++i) text += "GPU" + num(i) + "(Allocated)\t";
    if (info.has_pending()) text += "Pending(Allocated)\t";
    text += "\n";
    long long used = 0;
    std::string body;
    std::set<std::string> seen;
    bool seen_no_uid = false;
    for (int64_t i : info.dev_order()) {
      DeviceInfo &dev = info.devs[i];
      used += dev.used;
      for (const JVal *pod : dev.pods) {
        const JVal *uidv = pod->at({"metadata", "uid"});
        const bool has_uid = uidv && uidv->is_str();
        if (has_uid ? seen.count(uidv->s) != 0 : seen_no_uid) continue;
        body += str_field(*pod, {"metadata", "name"}) + "\t" + str_field(*pod, {"metadata", "namespace"}) + "\t";
        const int64_t count = info.gpu_count + (info.has_pending() ? 1 : 0);
        const Allocation allocation = get_allocation(*pod);
        for (int64_t k = 0; k < count; ++k) {
          if (!allocation.empty()) {
            long long v = 0;
            for (const auto &kv : allocation)
              if (kv.first == k) v = kv.second;
            body += num(v) + "\t";
          } else if (k == i || (i == -1 && k == info.gpu_count)) {
            body += num(gpu_memory_in_pod(*pod)) + "\t";
          } else {
            body += "0\t";
          }
        }
        body += "\n";
        if (has_uid) seen.insert(uidv->s);
        else seen_no_uid = true;
      }
    }
    if (line_len == 0) line_len = runes(body) + 10;
    text += body;
    const long long usage = (long long)((double)used / (double)info.gpu_total * 100);
    text += "Allocated :\t" + num(used) + " (" + num(usage) + "%)\t\n";
    text += "Total :\t" + num(info.gpu_total) + " \t\n";
    text += std::string(line_len, '-') + "\n";
    used_cluster += used;
    total_cluster += info.gpu_total;
  }
  text += "\n\nAllocated/Total GPU Memory In Cluster:\t";
  const long long usage = total_cluster > 0 ? (long long)((double)used_cluster / (double)total_cluster * 100) : 0;
  text += num(used_cluster) + "/" + num(total_cluster) + " (" + num(usage) + "%)\t\n";
  return TabWriter().run(text);
}

// podinfo.go:58-94: first try plus up to 5 more, 100 ms after each retry
bool retry_get(const Transport &t, const std::string &path, JVal &out) {
  std::string err;
  if (t.api_call("GET", path, "", nullptr, out, err)) return true;
  for (int i = 0; i < kListRetries; ++i) {
    const bool ok = t.api_call("GET", path, "", nullptr, out, err);
    sleep_s(0.1);
    if (ok) return true;
  }
  return false;
}

}  // namespace

int main(int argc, char **argv) {
  signal(SIGPIPE, SIG_IGN);  // a TLS peer that hangs up is an error message, not a death
  bool details = false;
  std::string unit, node_name;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    if (a == "-d") details = true;
    else if (a == "--unit" && i + 1 < argc) unit = argv[++i];
    else if (a.compare(0, 7, "--unit=") == 0) unit = a.substr(7);
    else if (!a.empty() && a[0] == '-') die("usage: kubectl-inspect-gpushare-v2 [-d] [--unit GiB|MiB] [node]");
    else node_name = a;
  }
  if (!unit.empty() && unit != "GiB" && unit != "MiB") die("--unit must be GiB or MiB");
  const char *kc = getenv("KUBECONFIG");
  const char *home = getenv("HOME");
  const std::string path = kc && *kc ? kc : std::string(home ? home : "") + "/.kube/config";
  struct stat st;
  if (stat(path.c_str(), &st) != 0)  // podinfo.go:30-36
    die("kubeconfig " + path + " failed to find, please set KUBECONFIG env");
  Transport t;
  std::string err;
  if (!ApiConfig::from_kubeconfig(path, t.api, err)) die(err);
  t.timeout_s = 15.0;

  JVal pods_doc, nodes_doc, one_node;
  std::vector<const JVal *> nodes;
  if (!node_name.empty()) {
    if (!t.get_node(node_name, one_node, err)) {
      printf("Failed due to %s", err.c_str());
      return 1;
    }
    nodes.push_back(&one_node);
    if (!retry_get(t, "/api/v1/pods?fieldSelector=spec.nodeName%3D" + node_name, pods_doc)) {
      printf("Failed due to failed to get Pods in node %s", node_name.c_str());
      return 1;
    }
  } else {
    if (!t.api_call("GET", "/api/v1/nodes", "", nullptr, nodes_doc, err)) {
      printf("Failed due to %s", err.c_str());
      return 1;
    }
    const JVal *items = nodes_doc.get("items");
    if (items && items->type == JVal::Arr)
      for (const JVal &n : items->arr) {  // isGPUSharingNode, nodeinfo.go:214-222
        const JVal *m = n.at({"status", "allocatable", GS_RESOURCE_NAME});
        if (m && quantity(*m) > 0) nodes.push_back(&n);
      }
    if (!retry_get(t, "/api/v1/pods", pods_doc)) {
      printf("Failed due to failed to get Pods");
      return 1;
    }
  }
  std::vector<const JVal *> pods;  // podinfo.go:96-107: drop finished pods
  const JVal *items = pods_doc.get("items");
  if (items && items->type == JVal::Arr)
    for (const JVal &p : items->arr) {
      const std::string phase = str_field(p, {"status", "phase"});
      if (phase != "Succeeded" && phase != "Failed") pods.push_back(&p);
    }

  // buildAllNodeInfos (nodeinfo.go:47-60,96-135)
  std::vector<NodeInfo> infos;
  std::map<std::string, size_t> by_name;
  for (const JVal *n : nodes) {
    const std::string name = str_field(*n, {"metadata", "name"});
    if (!by_name.count(name)) {
      by_name[name] = infos.size();
      infos.emplace_back(n);
    }
    NodeInfo &info = infos[by_name[name]];
    for (const JVal *p : pods) {
      const JVal *nn = p->at({"spec", "nodeName"});
      if (nn && nn->is_str() && nn->s == name) info.pods.push_back(p);
    }
  }
  for (NodeInfo &info : infos) {
    if (info.gpu_total <= 0) continue;
    if (unit.empty() && info.gpu_count != 0)  // setUnit with the threshold moved from 100 to 1024
      unit = info.gpu_total / info.gpu_count > 1024 ? "MiB" : "GiB";
    info.build_device_info();
  }
  const std::string report = details ? display_details(infos) : display_summary(infos, unit);
  fwrite(report.data(), 1, report.size(), stdout);
  return 0;
}

// pods_from_json.cpp -- test driver for daemon_cpp/json.h + k8s.h: reads a PodList (or a bare
// array) on stdin and prints, one line per pod, the fields the native daemon hands to
// gs_allocate, as a JSON array. tests/test_native_daemon_cluster.py compares the lines with
// what the Python host derives from the same document (api.Pod.from_k8s).
#include <stdio.h>

#include <iostream>
#include <iterator>
#include <string>

#include "k8s.h"

using namespace gsd;

int main() {
  std::string text((std::istreambuf_iterator<char>(std::cin)), std::istreambuf_iterator<char>());
  JVal doc;
  if (!json_parse(text, doc)) {
    printf("PARSE-ERROR\n");
    return 0;
  }
  const JVal *items = doc.type == JVal::Arr ? &doc : doc.get("items");
  if (!items || items->type != JVal::Arr) return 0;
  for (const JVal &p : items->arr) {
    Pod pod;
    std::string err;
    if (!pod_from_json(p, pod, err)) {
      printf("POD-ERROR\n");
      continue;
    }
    auto opt = [](bool has, const std::string &s) { return has ? json_quote(s) : std::string("null"); };
    printf("[%s,%s,%s,%s,%llu,%s,%s,%s]\n", json_quote(pod.uid).c_str(), json_quote(pod.ns).c_str(),
           json_quote(pod.name).c_str(), json_quote(pod.node).c_str(), (unsigned long long)pod.limit,
           opt(pod.has_idx, pod.idx).c_str(), opt(pod.has_time, pod.time).c_str(),
           opt(pod.has_assigned, pod.assigned).c_str());
  }
  return 0;
}

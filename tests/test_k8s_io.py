"""N1: pod source + patcher with the reference's retry ladders (podmanager.go:187-245,
allocate.go:132-152), against the oracle's LadderPodSource and over real HTTP."""
import copy
import http.server
import json
import threading
import urllib.parse

import pytest

from gpushare_b200 import api, k8s_io
from oracle import ref_model as R


class FakeCluster:
    """In-memory kubelet + apiserver. `script` maps a call name to a list of outcomes consumed
    per call: an Exception to raise, or None to answer normally. Records every call."""

    def __init__(self, pods, script=None):
        self.pods = copy.deepcopy(pods)
        self.script = {k: list(v) for k, v in (script or {}).items()}
        self.calls = []

    def _next(self, name):
        q = self.script.get(name) or []
        out = q.pop(0) if q else None
        if isinstance(out, Exception):
            raise out

    def kubelet_pods(self):
        self.calls.append(("kubelet_pods",))
        self._next("kubelet_pods")
        return {"items": copy.deepcopy(self.pods)}

    def api_list(self, field_selector, label_selector):
        self.calls.append(("api_list", field_selector, label_selector))
        self._next("api_list")
        want = dict(kv.split("=") for kv in field_selector.split(","))
        return {"items": [copy.deepcopy(p) for p in self.pods
                          if p["spec"].get("nodeName") == want["spec.nodeName"]
                          and p["status"]["phase"] == want["status.phase"]]}

    def api_patch(self, namespace, name, body):
        self.calls.append(("api_patch", namespace, name, bytes(body)))
        self._next("api_patch")
        ann = json.loads(body)["metadata"]["annotations"]
        for p in self.pods:
            if p["metadata"]["namespace"] == namespace and p["metadata"]["name"] == name:
                p["metadata"].setdefault("annotations", {}).update(ann)


def mkpod(name, limit, idx, assume, phase="Pending", node="node-a", assigned="false"):
    return {"metadata": {"uid": "uid-" + name, "name": name, "namespace": "default",
                         "annotations": {"ALIYUN_COM_GPU_MEM_IDX": idx,
                                         "ALIYUN_COM_GPU_MEM_ASSUME_TIME": assume,
                                         "ALIYUN_COM_GPU_MEM_ASSIGNED": assigned}},
            "spec": {"nodeName": node, "containers": [
                {"name": "c", "resources": {"limits": {"aliyun.com/gpu-mem": limit}}}]},
            "status": {"phase": phase}}


PODS = [mkpod("a", "2", "1", "300"), mkpod("b", "2", "2", "200"), mkpod("run", "2", "3", "100", phase="Running"),
        mkpod("other", "2", "4", "50", node="node-b"), mkpod("big", "4", "5", "10")]
UUIDS = ["GPU-%08x-0000-0000-0000-%012x" % (i, i) for i in range(8)]
E = k8s_io.ApiError
LOCK = k8s_io.OptimisticLockErrorMsg

SCRIPTS = {
    "all-fine": {},
    "kubelet-down-then-apiserver": {"kubelet_pods": [E("refused")] * 9},
    "kubelet-recovers-on-4th": {"kubelet_pods": [E("x"), E("x"), E("x")]},
    "kubelet-and-apiserver-down": {"kubelet_pods": [E("x")] * 9, "api_list": [E("y")] * 4},
    "apiserver-recovers-on-last": {"kubelet_pods": [E("x")] * 9, "api_list": [E("y")] * 3},
    "patch-conflict-once": {"api_patch": [E(LOCK)]},
    "patch-conflict-twice": {"api_patch": [E(LOCK), E(LOCK)]},
    "patch-other-error": {"api_patch": [E("Operation cannot be fulfilled on pods \"b\": " + LOCK)]},
    "patch-fails-then-next-call-fine": {"api_patch": [E("boom")]},
}


@pytest.mark.parametrize("query_kubelet", [False, True])
@pytest.mark.parametrize("name", sorted(SCRIPTS))
def test_ladders_match_oracle(name, query_kubelet):
    reqs = [["x", "y"]] * 3 + [["p", "q", "r", "s"]]
    # oracle
    oc, osleeps = FakeCluster(PODS, SCRIPTS[name]), []
    src = R.LadderPodSource(oc, osleeps.append, "node-a", query_kubelet)
    oracle = R.NvidiaDevicePlugin(R.Globals("GiB"), [(u, 16384) for u in UUIDS], "node-a", src,
                                  clock=lambda: 42)
    # product
    pc, psleeps = FakeCluster(PODS, SCRIPTS[name]), []
    table = api.DeviceTable.from_raw(UUIDS, 16384, "GiB")
    lister = k8s_io.PodLister(pc, "node-a", query_kubelet, sleep=psleeps.append)
    plug = k8s_io.ClusterDevicePlugin(table, "node-a", lister, pc, clock=lambda: 42)
    for r in reqs:
        want = oracle.Allocate([r])
        got = R.decode_allocate_response(plug.Allocate(R.encode_allocate_request([r])).response)
        assert got == want, (name, r)
    assert pc.calls == oc.calls                  # same requests, same order, same bodies
    assert psleeps == osleeps                    # same back-off
    assert pc.pods == oc.pods                    # same annotations landed


def test_ladder_shapes():
    """Hand-checked shapes of the ladders (podmanager.go:212-216, 234-240)."""
    c, sleeps = FakeCluster(PODS, {"kubelet_pods": [E("x")] * 9, "api_list": [E("y")] * 4}), []
    lister = k8s_io.PodLister(c, "node-a", True, sleep=sleeps.append)
    with pytest.raises(k8s_io.ListerError, match="failed to get Pods assigned to node node-a"):
        lister.list()
    assert [x[0] for x in c.calls] == ["kubelet_pods"] * 9 + ["api_list"] * 4
    assert sleeps == [0.1] * 8 + [1.0] * 3
    assert c.calls[-1] == ("api_list", "spec.nodeName=node-a,status.phase=Pending", "")
    # kubelet path keeps Pending pods only and errors when there are none (H11)
    c2 = FakeCluster([mkpod("r", "2", "0", "1", phase="Running")])
    with pytest.raises(k8s_io.ListerError):
        k8s_io.PodLister(c2, "node-a", True, sleep=lambda s: None)._get_pod_list()


class _Handler(http.server.BaseHTTPRequestHandler):
    cluster = None
    log = []

    def log_message(self, *a):
        pass

    def _send(self, code, doc):
        body = json.dumps(doc).encode()
        self.send_response(code)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(body)))
        self.end_headers()
        self.wfile.write(body)

    def do_GET(self):
        u = urllib.parse.urlsplit(self.path)
        self.log.append(("GET", u.path, urllib.parse.parse_qs(u.query), self.headers.get("Authorization")))
        if u.path == "/pods/":
            return self._send(200, self.cluster.kubelet_pods())
        if u.path == "/api/v1/pods":
            q = urllib.parse.parse_qs(u.query)
            return self._send(200, self.cluster.api_list(q["fieldSelector"][0], ""))
        if u.path == "/api/v1/nodes/node-a":
            return self._send(200, {"metadata": {"labels": {"cgpu.disable.isolation": "true"}},
                                    "status": {"capacity": {"aliyun.com/gpu-count": "4"}}})
        self._send(404, {"kind": "Status", "message": "not found"})

    def do_PATCH(self):
        body = self.rfile.read(int(self.headers["Content-Length"]))
        self.log.append(("PATCH", self.path, body, self.headers.get("Content-Type")))
        parts = self.path.split("/")
        if self.path.startswith("/api/v1/nodes/"):
            return self._send(200, {})
        try:
            self.cluster.api_patch(parts[4], parts[6], body)
        except Exception as e:
            return self._send(409, {"kind": "Status", "message": str(e), "code": 409})
        self._send(200, {})


def test_over_real_http():
    cluster = FakeCluster(PODS, {"api_patch": [E(LOCK)]})
    _Handler.cluster, _Handler.log = cluster, []
    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), _Handler)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        base = "http://127.0.0.1:%d" % srv.server_address[1]
        t = k8s_io.HttpTransport(k8s_io.ApiConfig(base, token="tok"), kubelet_url=base, kubelet_token="ktok")
        table = api.DeviceTable.from_raw(UUIDS, 16384, "GiB")
        for query_kubelet in (True, False):
            lister = k8s_io.PodLister(t, "node-a", query_kubelet, sleep=lambda s: None)
            assert sorted(p.name for p in lister.list()) == (["a", "b", "big", "other"] if query_kubelet
                                                             else ["a", "b", "big"])
        plug = k8s_io.ClusterDevicePlugin(table, "node-a", k8s_io.PodLister(t, "node-a"), t, clock=lambda: 7)
        d = plug.Allocate(R.encode_allocate_request([["x", "y"]]))
        assert R.decode_allocate_response(d.response)[0]["ALIYUN_COM_GPU_MEM_IDX"] == "2"   # pod b is older
        patches = [x for x in _Handler.log if x[0] == "PATCH"]
        assert len(patches) == 2                                  # conflict, then the retry
        assert patches[0][1] == "/api/v1/namespaces/default/pods/b"
        assert patches[0][3] == "application/strategic-merge-patch+json"
        assert json.loads(patches[1][2]) == {"metadata": {"annotations": {
            "ALIYUN_COM_GPU_MEM_ASSIGNED": "true", "ALIYUN_COM_GPU_MEM_ASSUME_TIME": "7"}}}
        gets = [x for x in _Handler.log if x[0] == "GET" and x[1] == "/api/v1/pods"]
        assert gets[-1][2] == {"fieldSelector": ["spec.nodeName=node-a,status.phase=Pending"]}
        assert gets[-1][3] == "Bearer tok"
        assert [x[3] for x in _Handler.log if x[1] == "/pods/"] == ["Bearer ktok"]
        # start-up node calls
        assert k8s_io.disable_cgpu_isolation_or_not(t, "node-a") is True
        assert k8s_io.patch_gpu_count(t, "node-a", 4) is False    # capacity already says 4
        assert k8s_io.patch_gpu_count(t, "node-a", 8) is True
        last = _Handler.log[-1]
        assert last[1] == "/api/v1/nodes/node-a/status" and json.loads(last[2]) == {
            "status": {"allocatable": {"aliyun.com/gpu-count": "8"}, "capacity": {"aliyun.com/gpu-count": "8"}}}
    finally:
        srv.shutdown()


def test_kubeconfig_and_in_cluster(tmp_path, monkeypatch):
    kc = tmp_path / "config"
    kc.write_text("""
apiVersion: v1
kind: Config
current-context: c
contexts: [{name: c, context: {cluster: k, user: u}}]
clusters: [{name: k, cluster: {server: "https://10.0.0.1:6443", insecure-skip-tls-verify: true}}]
users: [{name: u, user: {token: abc}}]
""")
    monkeypatch.setenv("KUBECONFIG", str(kc))
    cfg = k8s_io.ApiConfig.load()
    assert (cfg.host, cfg.token, cfg.insecure) == ("https://10.0.0.1:6443", "abc", True)
    monkeypatch.setenv("KUBECONFIG", str(tmp_path / "missing"))
    monkeypatch.delenv("KUBERNETES_SERVICE_HOST", raising=False)
    with pytest.raises(RuntimeError, match="in-cluster"):
        k8s_io.ApiConfig.load()

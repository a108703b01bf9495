"""N3 (kubectl-inspect-gpushare) and N4 (manifests) -- SURVEY 8(f)."""
import os

import yaml

from gpushare_b200 import inspect_gpushare as I
from tests.conftest import ROOT


def node(name, ip, mem, count):
    return {"metadata": {"name": name},
            "status": {"allocatable": {"aliyun.com/gpu-mem": str(mem), "aliyun.com/gpu-count": str(count)},
                       "addresses": [{"type": "Hostname", "address": name},
                                     {"type": "InternalIP", "address": ip}]}}


def pod(name, ns, node_name, limit, idx=None, phase="Running", extra=None):
    ann = dict(extra or {})
    if idx is not None:
        ann["ALIYUN_COM_GPU_MEM_IDX"] = idx
    return {"metadata": {"uid": "u-" + name, "name": name, "namespace": ns, "annotations": ann},
            "spec": {"nodeName": node_name, "containers": [
                {"resources": {"limits": {"aliyun.com/gpu-mem": str(limit)}}}]},
            "status": {"phase": phase}}


NODES = [node("node-a", "10.0.0.5", 358, 2)]
PODS = [pod("train", "default", "node-a", 40, "0"), pod("infer", "team-b", "node-a", 10, "1"),
        pod("pending", "default", "node-a", 5, None, extra={"x": "y"}),
        pod("elsewhere", "default", "node-b", 7, "0"), pod("cpu-only", "default", "node-a", 0, None)]


def test_tabwriter_column_blocks():
    # widths are per block of consecutive lines that HAVE the column; the text after the last
    # tab is never padded; a line without tabs ends every block
    text = "a\tbbb\tc\nxxxx\ty\tz\nno tabs here\nlonger-cell\tq\n"
    assert I.tabwrite(text) == ("a     bbb  c\n"
                                "xxxx  y    z\n"
                                "no tabs here\n"
                                "longer-cell  q\n")
    assert I.tabwrite("x\t\n") == "x  \n"          # one tab-terminated cell, empty tail


def test_summary_table_hand_derived():
    """display.go:141-245 laid out by hand: columns are 8, 11, 23, 23 and 20 wide."""
    infos, unit = I.build_all_node_infos(PODS, NODES)
    assert unit == "GiB"                            # the reference would say MiB (179 > 100)
    got = I.display_summary(infos, unit)
    want = ("NAME" + " " * 4 + "IPADDRESS" + " " * 2 + "GPU0(Allocated/Total)" + " " * 2 +
            "GPU1(Allocated/Total)" + " " * 2 + "PENDING(Allocated)" + " " * 2 + "GPU Memory(GiB)\n" +
            "node-a" + " " * 2 + "10.0.0.5" + " " * 3 + "40/179" + " " * 17 + "10/179" + " " * 17 +
            "5" + " " * 19 + "55/358\n" +
            "-" * 59 + "\n" +
            "Allocated/Total GPU Memory In Cluster:\n" +
            "55/358 (15%)  \n")
    assert got == want


def test_details_table_hand_derived():
    """display.go:15-129: the pod rows form one block with the header row."""
    infos, _ = I.build_all_node_infos(PODS, NODES)
    got = I.display_details(infos)
    # the "Allocated :" and "Total :" lines follow the pod rows directly, so they belong to
    # the same blocks for the first two columns: widths 13 ("Allocated :" + 2) and 11
    rows = ("NAME" + " " * 9 + "NAMESPACE" + " " * 2 + "GPU0(Allocated)" + " " * 2 + "GPU1(Allocated)" + " " * 2 +
            "Pending(Allocated)" + " " * 2 + "\n" +
            "train" + " " * 8 + "default" + " " * 4 + "40" + " " * 15 + "0" + " " * 16 + "0" + " " * 19 + "\n" +
            "infer" + " " * 8 + "team-b" + " " * 5 + "0" + " " * 16 + "10" + " " * 15 + "0" + " " * 19 + "\n" +
            "pending" + " " * 6 + "default" + " " * 4 + "0" + " " * 16 + "0" + " " * 16 + "5" + " " * 19 + "\n" +
            "Allocated :" + " " * 2 + "55 (15%)" + " " * 3 + "\n" +
            "Total :" + " " * 6 + "358 " + " " * 7 + "\n")
    assert got.startswith("\nNAME:       node-a\nIPADDRESS:  10.0.0.5\n\n")
    assert rows in got
    assert got.endswith("Allocated/Total GPU Memory In Cluster:  55/358 (15%)  \n")
    # raw (pre-layout) pod rows incl. tabs and newline: train 22 + infer 21 + pending 23, + 10
    assert "\n" + "-" * 76 + "\n" in got and "-" * 77 not in got


def test_allocation_annotation_and_unit_guess():
    alloc = {"scheduler.framework.gpushare.allocation": '{"0": {"0": 3, "1": 4}, "1": {"1": 2}}'}
    p = pod("multi", "default", "node-a", 9, "0", extra=alloc)
    assert I.get_allocation(p) == {0: 3, 1: 6}
    infos, _ = I.build_all_node_infos([p], NODES)
    assert (infos[0].devs[0].used, infos[0].devs[1].used) == (3, 6)
    assert I.get_allocation(pod("bad", "d", "n", 1, extra={"scheduler.framework.gpushare.allocation": "{"})) == {}
    assert I.guess_unit(8 * 183359, 8) == "MiB" and I.guess_unit(8 * 179, 8) == "GiB"
    assert I.guess_unit(8 * 16, 8) == "GiB" and I.guess_unit(5, 0) == ""
    assert [p["metadata"]["name"] for p in I.filter_active_pods(
        [pod("a", "d", "n", 1, phase="Succeeded"), pod("b", "d", "n", 1, phase="Failed"),
         pod("c", "d", "n", 1, phase="Pending")])] == ["c"]


def test_manifests_keep_the_drop_in_names():
    """device-plugin-ds.yaml / device-plugin-rbac.yaml of the reference, name for name."""
    ds = yaml.safe_load(open(os.path.join(ROOT, "deploy", "gpushare-b200-daemonset.yaml")))
    assert (ds["kind"], ds["metadata"]["name"], ds["metadata"]["namespace"]) == \
        ("DaemonSet", "gpushare-device-plugin-ds", "kube-system")
    labels = {"component": "gpushare-device-plugin", "app": "gpushare", "name": "gpushare-device-plugin-ds"}
    assert ds["spec"]["selector"]["matchLabels"] == labels == ds["spec"]["template"]["metadata"]["labels"]
    spec = ds["spec"]["template"]["spec"]
    assert spec["serviceAccount"] == "gpushare-device-plugin" and spec["hostNetwork"] is True
    assert spec["nodeSelector"] == {"gpushare": "true"}
    c = spec["containers"][0]
    assert c["command"][:4] == ["gpushare-device-plugin-v2", "-logtostderr", "--v=5", "--memory-unit=GiB"]
    env = {e["name"]: e for e in c["env"]}
    assert env["KUBECONFIG"]["value"] == "/etc/kubernetes/kubelet.conf"
    assert env["NODE_NAME"]["valueFrom"]["fieldRef"]["fieldPath"] == "spec.nodeName"
    assert c["volumeMounts"][0]["mountPath"] == spec["volumes"][0]["hostPath"]["path"] == \
        "/var/lib/kubelet/device-plugins"
    assert c["resources"]["limits"] == c["resources"]["requests"]
    # the daemon accepts exactly these arguments
    from gpushare_b200 import daemon
    a = daemon.parse_args(c["command"][1:])
    assert a.memory_unit == "GiB" and a.health_check and a.v == 5
    docs = list(yaml.safe_load_all(open(os.path.join(ROOT, "deploy", "gpushare-b200-rbac.yaml"))))
    kinds = {d["kind"]: d for d in docs}
    assert set(kinds) == {"ClusterRole", "ServiceAccount", "ClusterRoleBinding"}
    assert all(d["metadata"]["name"] == "gpushare-device-plugin" for d in docs)
    rules = {tuple(r["resources"]): set(r["verbs"]) for r in kinds["ClusterRole"]["rules"]}
    assert {"patch", "list", "get"} <= rules[("pods",)] and "patch" in rules[("nodes/status",)]
    assert "get" in rules[("nodes", "nodes/proxy")]
    assert kinds["ClusterRoleBinding"]["subjects"][0]["namespace"] == "kube-system"


def test_inspect_cli_end_to_end_over_http(tmp_path, monkeypatch, capsys):
    """main(): kubeconfig -> apiserver (nodes, pods) -> summary / details / single node."""
    import http.server
    import json
    import threading
    import urllib.parse
    nodes = NODES + [node("cpu-node", "10.0.0.9", 0, 0)]
    pods = PODS + [pod("done", "default", "node-a", 3, "0", phase="Succeeded")]
    seen = []

    class H(http.server.BaseHTTPRequestHandler):
        def log_message(self, *a):
            pass

        def do_GET(self):
            u = urllib.parse.urlsplit(self.path)
            seen.append(self.path)
            if u.path == "/api/v1/nodes":
                doc = {"items": nodes}
            elif u.path == "/api/v1/nodes/node-a":
                doc = nodes[0]
            elif u.path == "/api/v1/pods":
                sel = urllib.parse.parse_qs(u.query).get("fieldSelector", [""])[0]
                want = sel.split("=", 1)[1] if sel else None
                doc = {"items": [p for p in pods if want is None or p["spec"]["nodeName"] == want]}
            else:
                self.send_response(404)
                self.end_headers()
                return
            body = json.dumps(doc).encode()
            self.send_response(200)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), H)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        kc = tmp_path / "config"
        kc.write_text("apiVersion: v1\nkind: Config\ncurrent-context: c\n"
                      "contexts: [{name: c, context: {cluster: k, user: u}}]\n"
                      "clusters: [{name: k, cluster: {server: 'http://127.0.0.1:%d'}}]\n"
                      "users: [{name: u, user: {token: t}}]\n" % srv.server_address[1])
        monkeypatch.setenv("KUBECONFIG", str(kc))
        assert I.main([]) == 0
        out = capsys.readouterr().out
        assert "GPU Memory(GiB)" in out and "node-a" in out and "cpu-node" not in out   # not a sharing node
        assert "55/358" in out                                                          # "done" pod filtered
        assert I.main(["-d", "node-a"]) == 0
        out = capsys.readouterr().out
        assert "NAME:       node-a" in out and "train" in out and "elsewhere" not in out
        assert any("fieldSelector=spec.nodeName%3Dnode-a" in p for p in seen)
        assert I.main(["--unit", "MiB"]) == 0
        assert "GPU Memory(MiB)" in capsys.readouterr().out
    finally:
        srv.shutdown()
    monkeypatch.setenv("KUBECONFIG", str(tmp_path / "missing"))
    monkeypatch.setenv("HOME", str(tmp_path))
    import pytest
    with pytest.raises(SystemExit):
        I.main([])


def test_native_inspect_binary_matches_the_python_cli(tmp_path, monkeypatch, capsys):
    """bin/kubectl-inspect-gpushare-v2 (daemon_cpp/inspect_main.cpp) against main() here, byte
    for byte, on generated clusters: several nodes, pending pods, the scheduler-framework
    allocation annotation, finished pods, non-sharing nodes, every flag."""
    import http.server
    import json
    import random
    import subprocess
    import threading
    import urllib.parse
    exe = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "bin", "kubectl-inspect-gpushare-v2")
    rng = random.Random(7)
    state = {}

    class H(http.server.BaseHTTPRequestHandler):
        def log_message(self, *a):
            pass

        def do_GET(self):
            u = urllib.parse.urlsplit(self.path)
            nodes, pods = state["nodes"], state["pods"]
            if state.get("fail_pods", 0) > 0 and u.path == "/api/v1/pods":
                state["fail_pods"] -= 1
                body = json.dumps({"kind": "Status", "message": "etcdserver: request timed out"}).encode()
                self.send_response(500)
            elif u.path == "/api/v1/nodes":
                body = json.dumps({"items": nodes}).encode()
                self.send_response(200)
            elif u.path.startswith("/api/v1/nodes/"):
                hit = [n for n in nodes if n["metadata"]["name"] == u.path.rsplit("/", 1)[1]]
                body = json.dumps(hit[0] if hit else {"kind": "Status", "message": "nodes not found"}).encode()
                self.send_response(200 if hit else 404)
            elif u.path == "/api/v1/pods":
                sel = urllib.parse.parse_qs(u.query).get("fieldSelector", [""])[0]
                want = sel.split("=", 1)[1] if sel else None
                body = json.dumps({"items": [p for p in pods if want is None or p["spec"].get("nodeName") == want]}).encode()
                self.send_response(200)
            else:
                body = b"{}"
                self.send_response(404)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), H)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    kc = tmp_path / "config"
    kc.write_text("apiVersion: v1\nkind: Config\ncurrent-context: c\nclusters:\n- cluster:\n"
                  "    server: http://127.0.0.1:%d\n  name: k\ncontexts:\n- context:\n    cluster: k\n    user: u\n"
                  "  name: c\nusers:\n- name: u\n  user:\n    token: t\n" % srv.server_address[1])
    monkeypatch.setenv("KUBECONFIG", str(kc))
    env = dict(os.environ, KUBECONFIG=str(kc))

    def both(argv):
        snapshot = dict(state)
        rc = I.main(argv)
        want = capsys.readouterr().out
        state.update(snapshot)
        got = subprocess.run([exe] + argv, env=env, capture_output=True, timeout=60)
        assert got.returncode == rc, got.stderr
        assert got.stdout.decode() == want, (argv, got.stdout.decode(), want)
        return want

    try:
        for trial in range(12):
            n_nodes = rng.randrange(1, 5)
            nodes, pods = [], []
            for k in range(n_nodes):
                count = rng.choice([0, 1, 2, 4, 8])
                per = rng.choice([16, 179, 16384, 183359])
                nodes.append(node("node-%d" % k, "10.0.%d.%d" % (trial, k), count * per, count))
                if rng.random() < 0.2:
                    nodes[-1]["status"]["addresses"] = []
                for j in range(rng.randrange(0, 9)):
                    extra = {}
                    idx = rng.choice([None, "0", "1", "7", "-1", "x", "+2"])
                    if rng.random() < 0.25 and count:
                        extra["scheduler.framework.gpushare.allocation"] = json.dumps(
                            {str(c): {str(rng.randrange(count)): rng.randrange(1, 50)} for c in range(rng.randrange(1, 3))})
                    elif rng.random() < 0.1:
                        extra["scheduler.framework.gpushare.allocation"] = rng.choice(["{", "[]", '{"a": {"0": 1}}'])
                    p = pod("p%d-%d-%d" % (trial, k, j), rng.choice(["default", "team-b", "kube-system"]),
                            "node-%d" % k, rng.choice([0, 1, 2, 10, 40, 1024]), idx,
                            phase=rng.choice(["Running", "Running", "Pending", "Succeeded", "Failed"]), extra=extra)
                    if rng.random() < 0.3:
                        p["spec"]["containers"].append({"resources": {"limits": {"aliyun.com/gpu-mem": "3"}}})
                    pods.append(p)
            state.update(nodes=nodes, pods=pods)
            both([])
            both(["-d"])
            both(["--unit", "MiB"])
            both(["-d", "node-0"])
            both(["node-%d" % (n_nodes - 1)])
        state["fail_pods"] = 3                              # three failures, then the list arrives
        both([])
        state["fail_pods"] = 7                              # more than the 1 + 5 attempts
        out = both(["-d"])
        assert out == "Failed due to failed to get Pods"
        state["fail_pods"] = 0
        assert both(["no-such-node"]) == "Failed due to nodes not found"
    finally:
        srv.shutdown()

"""The native daemon binary (daemon_cpp/) in cluster mode: kubelet /pods, apiserver LIST and
PATCH over real HTTP against the fake cluster of test_k8s_io.py, under the same failure
scripts, compared with the oracle's LadderPodSource call for call (podmanager.go:187-245,
allocate.go:60-152). Sleeps are real here: the 8 x 100 ms and 3 x 1 s ladders are timed."""
import http.server
import json
import os
import re
import signal
import subprocess
import tempfile
import threading
import time
import urllib.parse

import pytest

from gpushare_b200.grpc_harness import KubeletStub
from gpushare_b200.native_transport import NativeClient
from oracle import ref_model as R
from tests.conftest import FAKE_NVML, ROOT
from tests.test_daemon import NATIVE_BIN
from tests.test_k8s_io import PODS, SCRIPTS, UUIDS, FakeCluster

STAMP = re.compile(rb'"ALIYUN_COM_GPU_MEM_ASSUME_TIME":"\d+"')


class Handler(http.server.BaseHTTPRequestHandler):
    cluster = None
    node_gets = 0
    node_patches = []
    labels = {}
    capacity = {}

    def log_message(self, *a):
        pass

    def _send(self, code, doc, raw=None):
        body = raw if raw is not None else json.dumps(doc).encode()
        self.send_response(code)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(body)))
        self.end_headers()
        self.wfile.write(body)

    def do_GET(self):
        cls = type(self)
        u = urllib.parse.urlsplit(self.path)
        try:
            if u.path == "/pods/":
                return self._send(200, cls.cluster.kubelet_pods())
            if u.path == "/api/v1/pods":
                q = urllib.parse.parse_qs(u.query)
                return self._send(200, cls.cluster.api_list(q["fieldSelector"][0], ""))
        except Exception as e:                          # the script says this call fails
            if u.path == "/pods/":
                return self._send(500, None, raw=b"kubelet is sad: " + str(e).encode())
            return self._send(500, {"kind": "Status", "message": str(e), "code": 500})
        if u.path == "/api/v1/nodes/node-a":
            cls.node_gets += 1
            return self._send(200, {"metadata": {"labels": cls.labels}, "status": {"capacity": cls.capacity}})
        self._send(404, {"kind": "Status", "message": "not found"})

    def do_PATCH(self):
        cls = type(self)
        body = self.rfile.read(int(self.headers["Content-Length"]))
        if self.path.startswith("/api/v1/nodes/"):
            cls.node_patches.append((self.path, body, self.headers.get("Content-Type")))
            return self._send(200, {})
        parts = self.path.split("/")
        try:
            cls.cluster.api_patch(parts[4], parts[6], body)
        except Exception as e:
            return self._send(409, {"kind": "Status", "message": str(e), "code": 409})
        self._send(200, {})


def _normalise(calls):
    return [tuple(STAMP.sub(b'"T"', x) if isinstance(x, bytes) else x for x in c) for c in calls]


def _strip_time(pods):
    out = json.loads(json.dumps(pods))
    for p in out:
        ann = p["metadata"].get("annotations", {})
        if ann.get("ALIYUN_COM_GPU_MEM_ASSIGNED") == "true":
            ann["ALIYUN_COM_GPU_MEM_ASSUME_TIME"] = "T"
    return out


class Rig:
    """fake cluster over HTTP + kubelet registration stub + the binary."""

    def __init__(self, d, script, query_kubelet, gpus=8, labels=None, capacity=None, tls=None, ca_line="",
                 user_lines="", client_ca=None):
        self.cluster = FakeCluster(PODS, script)
        self.handler = type("H", (Handler,), {"cluster": self.cluster, "node_gets": 0, "node_patches": [],
                                              "labels": labels or {}, "capacity": capacity or {}})
        self.http = http.server.ThreadingHTTPServer(("127.0.0.1", 0), self.handler)
        if tls:                                        # (cert_file, key_file)
            import ssl
            ctx = ssl.SSLContext(ssl.PROTOCOL_TLS_SERVER)
            ctx.load_cert_chain(*tls)
            if client_ca:                              # mutual TLS: the apiserver wants a client certificate
                ctx.verify_mode = ssl.CERT_REQUIRED
                ctx.load_verify_locations(client_ca)
            self.http.socket = ctx.wrap_socket(self.http.socket, server_side=True)
        threading.Thread(target=self.http.serve_forever, daemon=True).start()
        base = "%s://127.0.0.1:%d" % ("https" if tls else "http", self.http.server_address[1])
        kc = os.path.join(d, "kubeconfig")
        with open(kc, "w") as f:                       # as `kubectl config view --minify` prints it
            f.write("apiVersion: v1\nkind: Config\ncurrent-context: c\nclusters:\n- cluster:\n%s    server: %s\n"
                    "  name: k\ncontexts:\n- context:\n    cluster: k\n    user: u\n  name: c\nusers:\n- name: u\n"
                    "  user:\n    token: tok\n%s" % (ca_line, base, user_lines))
        self.kubelet = KubeletStub(os.path.join(d, "kubelet.sock"))
        self.kubelet.start()
        env = dict(os.environ, NODE_NAME="node-a", FAKE_NVML_GPUS=str(gpus), FAKE_NVML_MEM_MIB="16384",
                   KUBECONFIG=kc, GS_KUBELET_URL=base)
        args = [NATIVE_BIN, "--v=1", "--memory-unit=GiB", "--nvml-lib", FAKE_NVML, "--socket-dir", d,
                "--token", "ktok", "--timeout", "5"]
        if query_kubelet:
            args.append("--query-kubelet")
        self.proc = subprocess.Popen(args, env=env, stderr=subprocess.PIPE, text=True)
        self.sock = os.path.join(d, "aliyungpushare.sock")

    def close(self):
        if self.proc.poll() is None:
            self.proc.send_signal(signal.SIGTERM)
            try:
                self.proc.wait(10)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        self.kubelet.stop()
        self.http.shutdown()


@pytest.mark.parametrize("query_kubelet", [False, True])
@pytest.mark.parametrize("name", sorted(SCRIPTS))
def test_binary_ladders_match_oracle(name, query_kubelet):
    reqs = [["x", "y"]] * 3 + [["p", "q", "r", "s"]]
    oc, osleeps = FakeCluster(PODS, SCRIPTS[name]), []
    src = R.LadderPodSource(oc, osleeps.append, "node-a", query_kubelet)
    oracle = R.NvidiaDevicePlugin(R.Globals("GiB"), [(u, 16384) for u in UUIDS], "node-a", src,
                                  clock=lambda: 42)
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        rig = Rig(d, SCRIPTS[name], query_kubelet, capacity={"aliyun.com/gpu-count": "8"})
        try:
            assert rig.kubelet.wait_registered(30), rig.proc.stderr.read() if rig.proc.poll() is not None else ""
            t0 = time.time()
            with NativeClient(rig.sock, timeout_ms=10000) as cl:
                for r in reqs:
                    want = oracle.Allocate([r])
                    got = R.decode_allocate_response(cl.call("Allocate", R.encode_allocate_request([r])))
                    assert got == want, (name, r)
            elapsed = time.time() - t0
        finally:
            rig.close()
        assert _normalise(rig.cluster.calls) == _normalise(oc.calls)   # same requests, same order, same bodies
        assert _strip_time(rig.cluster.pods) == _strip_time(oc.pods)   # same annotations landed
        assert elapsed >= 0.9 * sum(osleeps), (elapsed, osleeps)       # the back-off really happened
        assert elapsed < sum(osleeps) + 5
        assert rig.proc.returncode == 0
        assert rig.handler.node_gets == 2 and rig.handler.node_patches == []   # capacity already said 8


def test_binary_startup_node_calls_and_cgpu_label():
    """server.go:57-64: gpu-count patched when capacity disagrees; the cGPU label adds
    CGPU_DISABLE=true to successful answers (allocate.go:125-127)."""
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        rig = Rig(d, {}, False, gpus=4, labels={"cgpu.disable.isolation": "true"},
                  capacity={"aliyun.com/gpu-count": "2"})
        try:
            assert rig.kubelet.wait_registered(30), rig.proc.stderr.read() if rig.proc.poll() is not None else ""
            path, body, ctype = rig.handler.node_patches[0]
            assert path == "/api/v1/nodes/node-a/status" and ctype == "application/strategic-merge-patch+json"
            assert json.loads(body) == {"status": {"allocatable": {"aliyun.com/gpu-count": "4"},
                                                   "capacity": {"aliyun.com/gpu-count": "4"}}}
            with NativeClient(rig.sock) as cl:
                envs = R.decode_allocate_response(cl.call("Allocate", R.encode_allocate_request([["x", "y"]])))
            assert envs[0]["CGPU_DISABLE"] == "true" and envs[0]["ALIYUN_COM_GPU_MEM_IDX"] == "2"
            patch = [c for c in rig.cluster.calls if c[0] == "api_patch"][0]
            assert patch[1:3] == ("default", "b")
            stamp = int(json.loads(patch[3])["metadata"]["annotations"]["ALIYUN_COM_GPU_MEM_ASSUME_TIME"])
            assert abs(stamp - time.time_ns()) < 60e9               # time.Now().UnixNano()
        finally:
            rig.close()
        assert rig.proc.returncode == 0


def _self_signed(d, name):
    import datetime
    import ipaddress

    from cryptography import x509
    from cryptography.hazmat.primitives import hashes, serialization
    from cryptography.hazmat.primitives.asymmetric import ec
    from cryptography.x509.oid import NameOID
    key = ec.generate_private_key(ec.SECP256R1())
    subject = x509.Name([x509.NameAttribute(NameOID.COMMON_NAME, name)])
    now = datetime.datetime.now(datetime.timezone.utc)
    cert = (x509.CertificateBuilder().subject_name(subject).issuer_name(subject).public_key(key.public_key())
            .serial_number(x509.random_serial_number()).not_valid_before(now - datetime.timedelta(days=1))
            .not_valid_after(now + datetime.timedelta(days=2))
            .add_extension(x509.SubjectAlternativeName([x509.IPAddress(ipaddress.ip_address("127.0.0.1"))]), False)
            .add_extension(x509.BasicConstraints(ca=True, path_length=None), True)
            .sign(key, hashes.SHA256()))
    cf, kf = os.path.join(d, name + ".crt"), os.path.join(d, name + ".key")
    open(cf, "wb").write(cert.public_bytes(serialization.Encoding.PEM))
    open(kf, "wb").write(key.private_bytes(serialization.Encoding.PEM, serialization.PrivateFormat.PKCS8,
                                           serialization.NoEncryption()))
    return cf, kf


def test_binary_over_tls():
    """The apiserver verified against the kubeconfig's certificate-authority (a wrong CA stops
    the daemon with exit code 1, skip-verify lets it through); the kubelet port is never
    verified (main.go:41 Insecure: true)."""
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        good, other = _self_signed(d, "server"), _self_signed(d, "stranger")
        rig = Rig(d, {}, True, tls=good, ca_line="    certificate-authority: %s\n" % good[0],
                  capacity={"aliyun.com/gpu-count": "8"})
        try:
            assert rig.kubelet.wait_registered(30), rig.proc.stderr.read() if rig.proc.poll() is not None else ""
            with NativeClient(rig.sock) as cl:
                envs = R.decode_allocate_response(cl.call("Allocate", R.encode_allocate_request([["x", "y"]])))
            assert envs[0]["ALIYUN_COM_GPU_MEM_IDX"] == "2"
            assert [c[0] for c in rig.cluster.calls] == ["kubelet_pods", "api_patch"]
        finally:
            rig.close()
        assert rig.proc.returncode == 0
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        good, other = _self_signed(d, "server"), _self_signed(d, "stranger")
        rig = Rig(d, {}, False, tls=good, ca_line="    certificate-authority: %s\n" % other[0])
        try:
            assert rig.proc.wait(30) == 1
            assert "certificate verify failed" in rig.proc.stderr.read()
        finally:
            rig.close()
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        # a kubeconfig that names no CA at all is verified against the system's roots (client-go's
        # behaviour), so a self-signed apiserver is refused rather than trusted by default
        good = _self_signed(d, "server")
        rig = Rig(d, {}, False, tls=good, ca_line="")
        try:
            assert rig.proc.wait(30) == 1
            assert "certificate verify failed" in rig.proc.stderr.read()
        finally:
            rig.close()
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        good = _self_signed(d, "server")
        rig = Rig(d, {}, False, tls=good, ca_line="    insecure-skip-tls-verify: true\n",
                  capacity={"aliyun.com/gpu-count": "8"})
        try:
            assert rig.kubelet.wait_registered(30), rig.proc.stderr.read() if rig.proc.poll() is not None else ""
        finally:
            rig.close()
        assert rig.proc.returncode == 0


def test_binary_reads_inline_kubeconfig_credentials():
    """kubeadm's kubelet.conf (what the reference DaemonSet mounts as KUBECONFIG,
    device-plugin-ds.yaml:41-47) carries its CA and client key pair inline as base64 *-data
    fields: the CA is honoured (a wrong one stops the daemon), the key pair is presented."""
    import base64
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        server, client, other = _self_signed(d, "server"), _self_signed(d, "client"), _self_signed(d, "stranger")
        b64 = lambda path: base64.b64encode(open(path, "rb").read()).decode()
        user = "    client-certificate-data: %s\n    client-key-data: %s\n" % (b64(client[0]), b64(client[1]))
        rig = Rig(d, {}, False, tls=server, ca_line="    certificate-authority-data: %s\n" % b64(server[0]),
                  user_lines=user, client_ca=client[0], capacity={"aliyun.com/gpu-count": "8"})
        try:
            assert rig.kubelet.wait_registered(30), rig.proc.stderr.read() if rig.proc.poll() is not None else ""
            with NativeClient(rig.sock) as cl:
                envs = R.decode_allocate_response(cl.call("Allocate", R.encode_allocate_request([["x", "y"]])))
            assert envs[0]["ALIYUN_COM_GPU_MEM_IDX"] == "2"
            # the Python host reads the same file the same way and gets through the same door
            from gpushare_b200 import k8s_io
            cfg = k8s_io.ApiConfig.from_kubeconfig(os.path.join(d, "kubeconfig"))
            assert "BEGIN CERTIFICATE" in cfg.ca_data and open(cfg.cert_file).read() == open(client[0]).read()
            node = k8s_io.HttpTransport(cfg, timeout=5).get_node("node-a")
            assert node["status"]["capacity"] == {"aliyun.com/gpu-count": "8"}
            os.remove(cfg.cert_file), os.remove(cfg.key_file)
        finally:
            rig.close()
        assert rig.proc.returncode == 0
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        server, other = _self_signed(d, "server"), _self_signed(d, "stranger")
        rig = Rig(d, {}, False, tls=server,
                  ca_line="    certificate-authority-data: %s\n" % base64.b64encode(open(other[0], "rb").read()).decode())
        try:
            assert rig.proc.wait(30) == 1
            assert "certificate verify failed" in rig.proc.stderr.read()
        finally:
            rig.close()


def test_binary_exit_codes():
    """gpumanager.go:71-76: 1 when the plugin cannot be built (apiserver unreachable), 2 when it
    cannot serve (no kubelet to register with); Go's flag package exits 2 on a bad flag."""
    with tempfile.TemporaryDirectory(prefix="gsc") as d:
        kc = os.path.join(d, "kubeconfig")
        open(kc, "w").write("clusters:\n- cluster:\n    server: http://127.0.0.1:9\n")
        env = dict(os.environ, NODE_NAME="node-a", FAKE_NVML_GPUS="1", KUBECONFIG=kc)
        out = subprocess.run([NATIVE_BIN, "--nvml-lib", FAKE_NVML, "--socket-dir", d, "--token", "t"],
                             env=env, capture_output=True, text=True, timeout=60)
        assert out.returncode == 1 and "Failed to get device plugin due to" in out.stderr
        pods = os.path.join(d, "pods.json")
        open(pods, "w").write('{"items": []}')
        out = subprocess.run([NATIVE_BIN, "--nvml-lib", FAKE_NVML, "--socket-dir", d, "--pod-source-file", pods],
                             env=env, capture_output=True, text=True, timeout=60)
        assert out.returncode == 2 and "Failed to start device plugin due to" in out.stderr
        out = subprocess.run([NATIVE_BIN, "--no-such-flag"], env=env, capture_output=True, text=True, timeout=60)
        assert out.returncode == 2 and "flag provided but not defined: -no-such-flag" in out.stderr
        out = subprocess.run([NATIVE_BIN, "--health-check=maybe"], env=env, capture_output=True, text=True, timeout=60)
        assert out.returncode == 2 and "invalid boolean value" in out.stderr


def test_binary_links_only_what_it_should():
    """The daemon links the core library and the system's OpenSSL; nothing of the oracle."""
    out = subprocess.run(["ldd", NATIVE_BIN], capture_output=True, text=True).stdout
    assert "libgpushare_core.so" in out and "libssl" in out and "oracle" not in out
    assert ROOT in out or "not found" not in out


def test_pod_decoding_matches_the_python_host(tmp_path):
    """daemon_cpp/json.h + pod_from_json against api.Pod.from_k8s on generated PodLists:
    escapes, non-ASCII, surrogate pairs, missing and null fields, quantities in every
    accepted form, numbers where strings are usual, wrapping sums."""
    import random

    from gpushare_b200 import _lib as L
    from gpushare_b200 import api
    exe = str(tmp_path / "pods_from_json")
    subprocess.run(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined",
                    "-I" + os.path.join(ROOT, "include"),
                    "-I" + os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "daemon_cpp"),
                    os.path.join(ROOT, "tests", "cpp", "pods_from_json.cpp"),
                    os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "csrc", "gosem.cpp"),
                    os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "csrc", "errors.cpp"),
                    "-o", exe, "-lssl", "-lcrypto"], check=True)
    rng = random.Random(20261106)
    words = ["a", "binpack-1", "kube-system", "node-a", "p\"q", "back\\slash", "tab\there", "nl\nx", "été",
             "中文", "\U0001f600", "", "x" * 300, "ctl\x01"]
    quantities = ["2", "1024", "1Ki", "1500m", "2Gi", "1e3", "0", "+3", ".5", "12E2", "18446744073709551615",
                  "9223372036854775807", 7, 40, 0]

    def pod():
        p = {"metadata": {}, "spec": {}, "status": {"phase": rng.choice(["Pending", "Running"])}}
        for k in ("uid", "namespace", "name"):
            if rng.random() < 0.9:
                p["metadata"][k] = rng.choice(words)
        if rng.random() < 0.8:
            ann = {}
            for k in ("ALIYUN_COM_GPU_MEM_IDX", "ALIYUN_COM_GPU_MEM_ASSUME_TIME", "ALIYUN_COM_GPU_MEM_ASSIGNED", "other"):
                if rng.random() < 0.7:
                    ann[k] = rng.choice(words + ["0", "-1", "true", "false", "1700000000000000000"])
            p["metadata"]["annotations"] = ann if rng.random() < 0.9 else None
        if rng.random() < 0.9:
            p["spec"]["nodeName"] = rng.choice(words)
        cs = []
        for _ in range(rng.randrange(4)):
            c = {"name": rng.choice(words)}
            r = rng.random()
            if r < 0.6:
                c["resources"] = {"limits": {"aliyun.com/gpu-mem": rng.choice(quantities), "cpu": "1"}}
            elif r < 0.7:
                c["resources"] = {"limits": {"cpu": "500m"}}
            elif r < 0.8:
                c["resources"] = {}
            cs.append(c)
        if rng.random() < 0.95:
            p["spec"]["containers"] = cs
        if rng.random() < 0.05:
            del p["spec"]
        return p

    for trial in range(40):
        items = [pod() for _ in range(rng.randrange(1, 12))]
        text = json.dumps({"items": items}, ensure_ascii=rng.random() < 0.5)
        out = subprocess.run([exe], input=text.encode(), capture_output=True, timeout=60)
        assert out.returncode == 0, out.stderr.decode()[-2000:]
        lines = out.stdout.decode().splitlines()
        assert len(lines) == len(items)
        for line, p in zip(lines, items):
            try:
                want = api.Pod.from_k8s(p)
            except L.GpushareError:
                assert line == "POD-ERROR"
                continue
            got = json.loads(line)
            assert got == [want.uid, want.namespace, want.name, want.node_name, want.gpu_mem_limit,
                           want.ann_idx, want.ann_assume_time, want.ann_assigned], (trial, p)
    for bad in (b"", b"{", b'{"items": [}', b'{"a": "\\x"}', b'[1,]', b'{"items": [{"metadata": "\\ud800"}]}'):
        out = subprocess.run([exe], input=bad, capture_output=True, timeout=60)
        assert out.returncode == 0 and (out.stdout in (b"PARSE-ERROR\n", b"") or b"[" in out.stdout)

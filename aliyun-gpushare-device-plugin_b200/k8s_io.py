"""Pod source and patcher: the Kubernetes I/O around Allocate.

Replaces, with the same retry ladders and error behaviour:
    kubeInit                          pkg/gpu/nvidia/podmanager.go:32-60
    getPodList                        podmanager.go:187-208   (kubelet /pods, Pending only)
    getPodListsByQueryKubelet         podmanager.go:210-225   (8 retries x 100 ms, then apiserver)
    getPodListsByListAPIServer        podmanager.go:227-245   (3 retries x 1 s)
    KubeletClient.GetNodeRunningPods  pkg/kubelet/client/client.go:119-134
    the PATCH with one retry          allocate.go:132-152
    disableCGPUIsolationOrNot         podmanager.go:62-75
    patchGPUCount                     podmanager.go:160-185

The decision itself stays in the C ABI (gs_allocate); this module only fetches the pods it
decides over and delivers the patch it produced. Transport is plain `http.client` + `ssl`
(no kubernetes client library in the image).
"""
import http.client
import json
import os
import ssl
import threading
import time
import urllib.parse

from . import _lib as L
from . import api

OptimisticLockErrorMsg = ("the object has been modified; please apply your changes to the "
                          "latest version and try again")          # const.go:15
RETRIES = 8                                                         # podmanager.go:29
SA_DIR = "/var/run/secrets/kubernetes.io/serviceaccount"


class ApiError(Exception):
    """An HTTP-level or Status-object failure; str() is what client-go's err.Error() would be
    for a Status (its .message), which is what allocate.go:140 compares."""

    def __init__(self, message, code=0):
        super().__init__(message)
        self.code = code


class ListerError(Exception):
    pass


# ---------------------------------------------------------------- configuration ---------
class ApiConfig:
    """Where the apiserver is and how to authenticate (kubeInit, podmanager.go:32-53)."""

    def __init__(self, host, token="", ca_file=None, cert_file=None, key_file=None, insecure=False, ca_data=None):
        self.host = host.rstrip("/")
        self.token = token
        self.ca_file, self.cert_file, self.key_file, self.insecure = ca_file, cert_file, key_file, insecure
        self.ca_data = ca_data                    # PEM text of certificate-authority-data

    @classmethod
    def load(cls):
        """KUBECONFIG if that file exists, else the in-cluster service account."""
        path = os.environ.get("KUBECONFIG", "")
        if path and os.path.exists(path):
            return cls.from_kubeconfig(path)
        return cls.in_cluster()

    @classmethod
    def in_cluster(cls):
        host, port = os.environ.get("KUBERNETES_SERVICE_HOST"), os.environ.get("KUBERNETES_SERVICE_PORT")
        if not host or not port:
            raise RuntimeError("unable to load in-cluster configuration, KUBERNETES_SERVICE_HOST and "
                               "KUBERNETES_SERVICE_PORT must be defined")
        with open(os.path.join(SA_DIR, "token")) as f:
            token = f.read().strip()
        if ":" in host:
            host = "[%s]" % host
        return cls("https://%s:%s" % (host, port), token, ca_file=os.path.join(SA_DIR, "ca.crt"))

    @classmethod
    def from_kubeconfig(cls, path):
        import yaml
        with open(path) as f:
            doc = yaml.safe_load(f)
        ctx_name = doc.get("current-context")
        ctx = next(c["context"] for c in doc.get("contexts", []) if c["name"] == ctx_name)
        cluster = next(c["cluster"] for c in doc.get("clusters", []) if c["name"] == ctx["cluster"])
        user = next((u["user"] for u in doc.get("users", []) if u["name"] == ctx.get("user")), {})
        # inline *-data fields (kubeadm's kubelet.conf, which the reference DaemonSet points
        # KUBECONFIG at, device-plugin-ds.yaml:41-47): the CA stays in memory, a client key pair
        # has to be a file for ssl.load_cert_chain
        import base64
        import tempfile

        def as_file(data):
            f = tempfile.NamedTemporaryFile(prefix="gs-kubeconfig-", suffix=".pem", delete=False)
            f.write(base64.b64decode(data))
            f.close()
            os.chmod(f.name, 0o600)
            return f.name
        cert_file, key_file = user.get("client-certificate"), user.get("client-key")
        if not cert_file and user.get("client-certificate-data") and user.get("client-key-data"):
            cert_file, key_file = as_file(user["client-certificate-data"]), as_file(user["client-key-data"])
        ca_data = cluster.get("certificate-authority-data")
        return cls(cluster["server"], user.get("token", ""), ca_file=cluster.get("certificate-authority"),
                   cert_file=cert_file, key_file=key_file,
                   insecure=bool(cluster.get("insecure-skip-tls-verify")),
                   ca_data=base64.b64decode(ca_data).decode() if ca_data else None)


def _connection(url, timeout, ca_file=None, cert_file=None, key_file=None, insecure=False, ca_data=None):
    u = urllib.parse.urlsplit(url)
    if u.scheme == "http":
        return http.client.HTTPConnection(u.hostname, u.port or 80, timeout=timeout), u
    # verification is off only when the configuration says so (insecure-skip-tls-verify; the kubelet
    # client, main.go:41); with no CA named, the system's roots apply, as in client-go
    have_ca = bool(ca_file or ca_data)
    ctx = (ssl.create_default_context(cafile=ca_file, cadata=ca_data) if have_ca
           else ssl.create_default_context())
    if insecure:
        ctx.check_hostname = False
        ctx.verify_mode = ssl.CERT_NONE
    if cert_file and key_file:
        ctx.load_cert_chain(cert_file, key_file)
    return http.client.HTTPSConnection(u.hostname, u.port or 443, timeout=timeout, context=ctx), u


# ---------------------------------------------------------------- transport -------------
class HttpTransport:
    """The four network calls of the hot path plus the two node calls of start-up."""

    def __init__(self, api: ApiConfig = None, kubelet_url=None, kubelet_token="", kubelet_cert=None,
                 kubelet_key=None, timeout=10.0):
        self.api = api
        self.kubelet_url = kubelet_url            # e.g. https://0.0.0.0:10250 (main.go:20-21)
        self.kubelet_token, self.kubelet_cert, self.kubelet_key = kubelet_token, kubelet_cert, kubelet_key
        self.timeout = timeout

    def _api(self, method, path, body=None, content_type=None):
        if self.api is None:
            raise ApiError("no apiserver configured")
        conn, _u = _connection(self.api.host, self.timeout, self.api.ca_file, self.api.cert_file,
                               self.api.key_file, self.api.insecure, getattr(self.api, "ca_data", None))
        headers = {"Accept": "application/json"}
        if self.api.token:
            headers["Authorization"] = "Bearer " + self.api.token
        if content_type:
            headers["Content-Type"] = content_type
        try:
            conn.request(method, path, body=body, headers=headers)
            resp = conn.getresponse()
            data = resp.read()
        except (OSError, http.client.HTTPException) as e:
            raise ApiError(str(e))
        finally:
            conn.close()
        if resp.status >= 400:
            try:
                msg = json.loads(data).get("message") or data.decode(errors="replace")
            except ValueError:
                msg = data.decode(errors="replace")
            raise ApiError(msg, resp.status)
        return json.loads(data) if data else {}

    def kubelet_pods(self):
        """client.go:119-134: GET https://host:port/pods/, body decoded whatever the status."""
        conn, _u = _connection(self.kubelet_url, self.timeout, None, self.kubelet_cert, self.kubelet_key,
                               insecure=True)      # main.go:41 Insecure: true
        headers = {"Authorization": "Bearer " + self.kubelet_token} if self.kubelet_token else {}
        try:
            conn.request("GET", "/pods/", headers=headers)
            data = conn.getresponse().read()
        except (OSError, http.client.HTTPException) as e:
            raise ApiError(str(e))
        finally:
            conn.close()
        try:
            return json.loads(data)
        except ValueError as e:
            raise ApiError("invalid /pods/ body: %s" % e)

    def api_list(self, field_selector, label_selector):
        q = {"fieldSelector": field_selector}
        if label_selector:
            q["labelSelector"] = label_selector
        return self._api("GET", "/api/v1/pods?" + urllib.parse.urlencode(q))

    def api_patch(self, namespace, name, body: bytes):
        self._api("PATCH", "/api/v1/namespaces/%s/pods/%s" % (urllib.parse.quote(namespace),
                                                               urllib.parse.quote(name)),
                  body=body, content_type="application/strategic-merge-patch+json")

    def get_node(self, name):
        return self._api("GET", "/api/v1/nodes/" + urllib.parse.quote(name))

    def patch_node_status(self, name, body: bytes):
        return self._api("PATCH", "/api/v1/nodes/%s/status" % urllib.parse.quote(name), body=body,
                         content_type="application/strategic-merge-patch+json")


# ---------------------------------------------------------------- lister ----------------
class PodLister:
    """getPendingPodsInNode's input side: the pod list through the reference's ladders."""

    def __init__(self, transport, node_name, query_kubelet=False, sleep=time.sleep):
        self.t, self.node_name, self.query_kubelet, self.sleep = transport, node_name, query_kubelet, sleep

    def _get_pod_list(self):
        pod_list = self.t.kubelet_pods()
        items = [p for p in (pod_list.get("items") or [])
                 if ((p.get("status") or {}).get("phase")) == "Pending"]
        if not items:
            raise ListerError("not found pending pod")
        return items

    def _by_kubelet(self):
        err = None
        try:
            return_items = self._get_pod_list()
        except Exception as e:
            return_items, err = None, e
        i = 0
        while i < RETRIES and err is not None:
            try:
                return_items, err = self._get_pod_list(), None
            except Exception as e:
                err = e
            self.sleep(0.1)
            i += 1
        if err is not None:
            return_items = self._by_apiserver()
        return return_items

    def _by_apiserver(self):
        selector = "spec.nodeName=%s,status.phase=Pending" % self.node_name
        err, doc = None, None
        try:
            doc = self.t.api_list(selector, "")
        except Exception as e:
            err = e
        i = 0
        while i < 3 and err is not None:
            try:
                doc, err = self.t.api_list(selector, ""), None
            except Exception as e:
                err = e
            self.sleep(1.0)
            i += 1
        if err is not None:
            raise ListerError("failed to get Pods assigned to node %s" % self.node_name)
        return doc.get("items") or []

    def list_k8s(self):
        return self._by_kubelet() if self.query_kubelet else self._by_apiserver()

    def list(self):
        """-> [api.Pod] for gs_allocate (node filter and de-dup happen inside it)."""
        return [api.Pod.from_k8s(p) for p in self.list_k8s()]


def patch_pod_assigned(transport, namespace, name, body: bytes):
    """allocate.go:136-152: one retry, only when the error text IS the optimistic-lock
    message (exact string comparison, as the reference does)."""
    try:
        transport.api_patch(namespace, name, body)
    except Exception as e:
        if str(e) == OptimisticLockErrorMsg:
            transport.api_patch(namespace, name, body)
        else:
            raise


class ClusterDevicePlugin(api.NvidiaDevicePlugin):
    """NvidiaDevicePlugin whose Allocate runs against a live pod source: list, decide in
    the C ABI, patch -- the whole of allocate.go:60-152 under one lock."""

    def __init__(self, table, node_name, lister: PodLister, transport, disable_cgpu_isolation=False,
                 clock=time.time_ns):
        super().__init__(table, node_name, disable_cgpu_isolation)
        self.lister, self.transport, self.clock = lister, transport, clock
        self.node_name, self.disable_cgpu_isolation = node_name, disable_cgpu_isolation
        self._lock = threading.Lock()                     # m.Lock(), allocate.go:60

    def Allocate(self, request: bytes):
        with self._lock:
            try:
                pods = self.lister.list()
            except Exception:                             # allocate.go:64-67
                return api.allocate_error_response(self.table, self.node_name, request,
                                                   L.GS_ALLOC_ERR_PODS)
            d = api.allocate(self.table, self.node_name, request, pods, self.clock(),
                             self.disable_cgpu_isolation)
            if d.kind == "matched":
                p = pods[d.pod_index]
                try:
                    patch_pod_assigned(self.transport, p.namespace, p.name, d.patch)
                except Exception:                         # allocate.go:138-151
                    return api.allocate_error_response(self.table, self.node_name, request,
                                                       L.GS_ALLOC_ERR_PATCH)
            return d


# ---------------------------------------------------------------- node calls (start-up) --
def disable_cgpu_isolation_or_not(transport, node_name):
    """podmanager.go:62-75."""
    node = transport.get_node(node_name)
    labels = (node.get("metadata") or {}).get("labels") or {}
    return labels.get("cgpu.disable.isolation") == "true"


def patch_gpu_count(transport, node_name, gpu_count):
    """podmanager.go:160-185: set status.capacity/allocatable["aliyun.com/gpu-count"] unless
    capacity already says so. The reference builds a two-way strategic-merge patch of the
    whole node; the only difference between old and new is these two keys, so the patch is
    exactly this body."""
    node = transport.get_node(node_name)
    cap = ((node.get("status") or {}).get("capacity") or {})
    if "aliyun.com/gpu-count" in cap:
        try:
            if api.quantity_value(str(cap["aliyun.com/gpu-count"])) == gpu_count:
                return False
        except L.GpushareError:
            pass
    body = json.dumps({"status": {"allocatable": {"aliyun.com/gpu-count": str(gpu_count)},
                                  "capacity": {"aliyun.com/gpu-count": str(gpu_count)}}},
                      separators=(",", ":")).encode()
    transport.patch_node_status(node_name, body)
    return True

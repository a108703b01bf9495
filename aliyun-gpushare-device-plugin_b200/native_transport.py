"""The native gRPC transport (csrc/grpc_server.cpp, grpc_client.cpp) behind the names of
grpc_harness.py: Start / Stop / Register / Serve of pkg/gpu/nvidia/server.go:114-177,232-249.

With a plain api.NvidiaDevicePlugin the whole Allocate RPC -- HTTP/2 frames, request scan,
decision, response -- runs on the library's serving thread and never enters Python. A plugin
whose Allocate is Python (k8s_io.ClusterDevicePlugin: the pod source is the network; the
daemon's probe hook) is called back with the request bytes and returns the response bytes.
"""
import ctypes as C
import os

from . import _lib as L
from . import api
from .grpc_harness import KUBELET_SOCKET, RESOURCE_NAME, SERVER_SOCK, SERVICE


class NativeDevicePluginServer:
    """Drop-in for grpc_harness.DevicePluginServer, served by gs_server_*."""

    def __init__(self, plugin: api.NvidiaDevicePlugin, socket_path=SERVER_SOCK, on_allocate=None,
                 on_decision=None, max_request_bytes=0):
        self.lib = L.load()
        self.plugin = plugin
        self.socket = socket_path
        self.on_allocate = on_allocate      # (AllocDecision) -> None; forces the Python path
        self.on_decision = on_decision      # (kind, pod_index, gpu_index, pod_req_units) -> None
        self.max_request_bytes = max_request_bytes
        self.h = None
        self._keep = None
        self._stash = None
        python_allocate = ("Allocate" in vars(plugin)
                           or type(plugin).Allocate is not api.NvidiaDevicePlugin.Allocate
                           or on_allocate is not None)
        self.python_allocate = python_allocate

    # ---- callbacks (run on the library's serving thread) -------------------------------
    def _allocate_cb(self, _user, req, n, resp, cap, out_len):
        try:
            if self._stash is not None:          # the retry after GS_E_NOSPACE: same answer,
                b, self._stash = self._stash, None   # Allocate must not run twice
            else:
                d = self.plugin.Allocate(C.string_at(req, n))   # never an RPC error (H8)
                if self.on_allocate:
                    self.on_allocate(d)
                b = d.response
            out_len[0] = len(b)
            if len(b) > cap:
                self._stash = b
                return L.GS_E_NOSPACE
            C.memmove(resp, b, len(b))
            return L.GS_OK
        except L.GpushareError as e:
            return e.status if e.status in (L.GS_E_WIRE, L.GS_E_INVAL) else L.GS_E_STATE
        except Exception:                        # noqa: BLE001 -- must not unwind into C
            return L.GS_E_STATE

    def _decision_cb(self, _user, res):
        try:
            r = res.contents
            self.on_decision(api.ALLOC_KIND.get(r.kind, str(r.kind)), r.pod_index, r.gpu_index,
                             r.pod_req_units)
        except Exception:                        # noqa: BLE001
            pass

    # ---- lifecycle ------------------------------------------------------------------------
    def Start(self):
        """server.go:114-142."""
        cfg = L.ServerConfig()
        cfg.socket_path = self.socket.encode()
        cfg.max_request_bytes = self.max_request_bytes
        keep = [cfg.socket_path]
        if self.python_allocate:
            cb = L.ServerAllocateFn(self._allocate_cb)
            cfg.allocate = cb
            keep.append(cb)
        elif self.on_decision is not None:
            cb = L.ServerDecisionFn(self._decision_cb)
            cfg.on_decision = cb
            keep.append(cb)
        h = C.c_void_p()
        L.check(self.lib.gs_server_start(self.plugin.h, C.byref(cfg), C.byref(h)), "gs_server_start")
        self.h, self._keep = h, keep

    def Stop(self):
        """server.go:145-155."""
        if self.h is None:
            return
        h, self.h = self.h, None
        self.lib.gs_server_stop(h)               # also gs_plugin_stop: open streams end OK
        self._keep = None

    def Register(self, kubelet_endpoint=KUBELET_SOCKET, resource_name=RESOURCE_NAME):
        """server.go:158-177."""
        rc = self.lib.gs_register(kubelet_endpoint.encode(), os.path.basename(self.socket).encode(),
                                  resource_name.encode(), 5000)
        if rc != L.GS_OK:
            raise L.GpushareError(rc, "Register: " + self.lib.gs_grpc_last_error().decode("utf-8", "replace"))

    def Serve(self, kubelet_endpoint=KUBELET_SOCKET):
        """server.go:232-249."""
        self.Start()
        try:
            self.Register(kubelet_endpoint, RESOURCE_NAME)
        except Exception:
            self.Stop()
            raise

    def stats(self):
        st = L.ServerStats()
        L.check(self.lib.gs_server_get_stats(self.h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in st._fields_}


class RpcError(Exception):
    def __init__(self, code, message):
        super().__init__("grpc status %d: %s" % (code, message))
        self.code, self.message = code, message


class NativeClient:
    """The kubelet side over the native transport (tests, benchmarks): one caller at a time."""

    def __init__(self, socket_path, timeout_ms=5000, recv_cap=1 << 20):
        self.lib = L.load()
        h = C.c_void_p()
        rc = self.lib.gs_grpc_client_connect(socket_path.encode(), timeout_ms, C.byref(h))
        if rc != L.GS_OK:
            raise L.GpushareError(rc, "dial: " + self.lib.gs_grpc_last_error().decode("utf-8", "replace"))
        self.h = h
        self.buf = C.create_string_buffer(recv_cap)

    def close(self):
        if self.h:
            self.lib.gs_grpc_client_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def call(self, method, request=b"", service=SERVICE):
        """Unary call; returns the response bytes or raises RpcError."""
        path = ("/%s/%s" % (service, method)).encode()
        n, status, msg = C.c_size_t(), C.c_int32(), C.create_string_buffer(512)
        rc = self.lib.gs_grpc_client_unary(self.h, path, request, len(request), self.buf, len(self.buf),
                                           C.byref(n), C.byref(status), msg, len(msg))
        if rc == L.GS_E_NOSPACE:
            raise L.GpushareError(rc, "response of %d bytes exceeds recv_cap" % n.value)
        if rc != L.GS_OK:
            raise L.GpushareError(rc, self.lib.gs_grpc_last_error().decode("utf-8", "replace"))
        if status.value != 0:
            raise RpcError(status.value, msg.value.decode("utf-8", "replace"))
        return C.string_at(self.buf, n.value)

    def open_stream(self, method, request=b"", service=SERVICE):
        path = ("/%s/%s" % (service, method)).encode()
        sid = C.c_uint32()
        rc = self.lib.gs_grpc_client_stream_open(self.h, path, request, len(request), C.byref(sid))
        if rc != L.GS_OK:
            raise L.GpushareError(rc, self.lib.gs_grpc_last_error().decode("utf-8", "replace"))
        return sid.value

    def next_message(self, sid, timeout_ms=5000, copy=True):
        """Next message of a server stream: bytes (with copy=False an (address, length) pair
        borrowed from the client until the next call on this stream); None when the stream
        ended OK; RpcError otherwise."""
        n, status, ptr = C.c_size_t(), C.c_int32(), C.c_void_p()
        rc = self.lib.gs_grpc_client_stream_next_borrow(self.h, sid, C.byref(ptr), C.byref(n),
                                                        C.byref(status), timeout_ms)
        if rc == L.GS_OK:
            return C.string_at(ptr, n.value) if copy else (ptr.value, n.value)
        if rc == L.GS_E_TIMEOUT:
            raise TimeoutError("no message within %d ms" % timeout_ms)
        if rc == L.GS_E_STATE and status.value == 0:
            return None
        if rc == L.GS_E_STATE and status.value > 0:
            raise RpcError(status.value, "stream ended")
        raise L.GpushareError(rc, self.lib.gs_grpc_last_error().decode("utf-8", "replace"))

    def cancel(self, sid):
        self.lib.gs_grpc_client_stream_cancel(self.h, sid)

"""kubelet-facing gRPC face of the plugin (deviceplugin/v1beta1 over a unix socket).

Replaces the transport half of pkg/gpu/nvidia/server.go: Start (114-142), Stop (145-155),
Register (158-177), Serve (232-249) and the five RPC entry points. Message bodies are not
parsed or built here: every handler hands kubelet's bytes to the C ABI and returns the
bytes the C ABI wrote (identity (de)serialisers), so the wire format lives in one place.

The reference is a Go binary with grpc-go; neither image has a Go toolchain or C++ gRPC,
so this face is Python grpcio. INTEGRATION.md shows the cgo shim a Go maintainer would
use instead; it binds the same C ABI.

Note on sizes: one B200 at MiB granularity advertises 183 359 devices = 11.3 MB, eight of
them 90 MB, beyond gRPC's default 4 MiB receive limit (the reference has the
same defect). The server lifts its own limits; kubelet's client limit is kubelet's.
"""
import os
import threading
from concurrent import futures

import grpc

from . import api

VERSION = "v1beta1"                                    # pluginapi.Version
DEVICE_PLUGIN_PATH = "/var/lib/kubelet/device-plugins/"
KUBELET_SOCKET = DEVICE_PLUGIN_PATH + "kubelet.sock"
SERVER_SOCK = DEVICE_PLUGIN_PATH + "aliyungpushare.sock"   # const.go:13
RESOURCE_NAME = "aliyun.com/gpu-mem"                   # const.go:11
SERVICE = "v1beta1.DevicePlugin"
REGISTRATION = "v1beta1.Registration"
_UNLIMITED = [("grpc.max_send_message_length", -1), ("grpc.max_receive_message_length", -1)]


def _ident(b):
    return b


def _varint(n):
    out = bytearray()
    while n >= 0x80:
        out.append((n & 0x7F) | 0x80)
        n >>= 7
    out.append(n)
    return bytes(out)


def encode_register_request(version, endpoint, resource_name):
    """RegisterRequest{version=1, endpoint=2, resource_name=3, options=4 (unset)} as
    server.go:166-170 fills it."""
    out = b""
    for field, s in ((1, version), (2, endpoint), (3, resource_name)):
        b = s.encode()
        out += _varint((field << 3) | 2) + _varint(len(b)) + b
    return out


class DevicePluginServer:
    """Start/Stop/Register/Serve of server.go around an api.NvidiaDevicePlugin."""

    def __init__(self, plugin: api.NvidiaDevicePlugin, socket_path=SERVER_SOCK, on_allocate=None):
        self.plugin = plugin
        self.socket = socket_path
        self.server = None
        self.on_allocate = on_allocate   # hook: (AllocDecision) -> None, e.g. queue a probe

    # ---- RPC bodies: bytes in, bytes out -------------------------------------------
    def _options(self, request, context):
        return self.plugin.GetDevicePluginOptions(request)

    def _list_and_watch(self, request, context):
        for msg in self.plugin.ListAndWatch():
            if not context.is_active():
                return
            yield msg

    def _allocate(self, request, context):
        d = self.plugin.Allocate(request)          # never an RPC error (H8)
        if self.on_allocate:
            self.on_allocate(d)
        return d.response

    def _pre_start(self, request, context):
        return self.plugin.PreStartContainer(request)

    def _preferred(self, request, context):
        # server.go:37-40 panics; grpc-go turns that into a dead RPC. Never advertised.
        context.abort(grpc.StatusCode.UNIMPLEMENTED, "implement me")

    def _handler(self):
        u = grpc.unary_unary_rpc_method_handler
        return grpc.method_handlers_generic_handler(SERVICE, {
            "GetDevicePluginOptions": u(self._options, _ident, _ident),
            "ListAndWatch": grpc.unary_stream_rpc_method_handler(self._list_and_watch, _ident, _ident),
            "Allocate": u(self._allocate, _ident, _ident),
            "PreStartContainer": u(self._pre_start, _ident, _ident),
            "GetPreferredAllocation": u(self._preferred, _ident, _ident),
        })

    # ---- lifecycle -------------------------------------------------------------------
    def cleanup(self):
        """server.go:203-209."""
        try:
            os.remove(self.socket)
        except FileNotFoundError:
            pass

    def Start(self):
        """server.go:114-142: remove a stale socket, listen, serve, self-dial within 5 s."""
        self.cleanup()
        self.server = grpc.server(futures.ThreadPoolExecutor(max_workers=8), options=_UNLIMITED)
        self.server.add_generic_rpc_handlers((self._handler(),))
        self.server.add_insecure_port("unix://" + self.socket)
        self.server.start()
        ch = grpc.insecure_channel("unix://" + self.socket)
        try:
            grpc.channel_ready_future(ch).result(timeout=5)
        finally:
            ch.close()

    def Stop(self):
        """server.go:145-155."""
        if self.server is None:
            return
        self.plugin.Stop()
        self.server.stop(grace=0.5).wait()
        self.server = None
        self.cleanup()

    def Register(self, kubelet_endpoint=KUBELET_SOCKET, resource_name=RESOURCE_NAME):
        """server.go:158-177."""
        ch = grpc.insecure_channel("unix://" + kubelet_endpoint)
        try:
            grpc.channel_ready_future(ch).result(timeout=5)
            call = ch.unary_unary("/%s/Register" % REGISTRATION, request_serializer=_ident,
                                  response_deserializer=_ident)
            call(encode_register_request(VERSION, os.path.basename(self.socket), resource_name),
                 timeout=5)
        finally:
            ch.close()

    def Serve(self, kubelet_endpoint=KUBELET_SOCKET):
        """server.go:232-249."""
        self.Start()
        try:
            self.Register(kubelet_endpoint, RESOURCE_NAME)
        except Exception:
            self.Stop()
            raise


class KubeletStub:
    """The kubelet side, for tests and benchmarks: a Registration server that records what
    the plugin sent, and a DevicePlugin client with its receive limit lifted (H12)."""

    def __init__(self, kubelet_socket):
        self.socket = kubelet_socket
        self.registered = []
        self._event = threading.Event()
        self.server = None

    def _register(self, request, context):
        self.registered.append(request)
        self._event.set()
        return b""

    def start(self):
        try:
            os.remove(self.socket)
        except FileNotFoundError:
            pass
        self.server = grpc.server(futures.ThreadPoolExecutor(max_workers=2))
        self.server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(REGISTRATION, {
            "Register": grpc.unary_unary_rpc_method_handler(self._register, _ident, _ident)}),))
        self.server.add_insecure_port("unix://" + self.socket)
        self.server.start()

    def wait_registered(self, timeout=5):
        return self._event.wait(timeout)

    def stop(self):
        if self.server:
            self.server.stop(grace=0).wait()
            self.server = None
        try:
            os.remove(self.socket)
        except FileNotFoundError:
            pass

    @staticmethod
    def connect(plugin_socket, unlimited=True):
        return grpc.insecure_channel("unix://" + plugin_socket, options=_UNLIMITED if unlimited else None)

    @staticmethod
    def method(channel, name, stream=False):
        path = "/%s/%s" % (SERVICE, name)
        if stream:
            return channel.unary_stream(path, request_serializer=_ident, response_deserializer=_ident)
        return channel.unary_unary(path, request_serializer=_ident, response_deserializer=_ident)

"""The node daemon's host logic in Python -- a TEST HARNESS, not the product.

The daemon that ships is the native binary (daemon_cpp/main.cpp -> bin/gpushare-device-plugin-v2,
what deploy/ runs). This twin exists because the retry-ladder and supervisor tests inject their
sleeps, clocks and fake sources here, and because it reads as a specification of
cmd/nvidia/main.go and gpumanager.go. It serves the same library over the same transport, but it
does NOT carry the slice probe: the probe worker (batching, concurrent health sweeps, the
no-allocation read-back) lives in the binary only, so there is one implementation to keep right.

Keeps the reference's command line (flag names, defaults, single- or double-dash forms as
Go's `flag` accepts) and environment (NODE_NAME mandatory, podmanager.go:55-58), and wires
the hot path: NVML enumerate -> device table -> plugin -> gRPC server -> Register ->
XID health watch.

Pod source: by default the live one (k8s_io: kubelet /pods or apiserver LIST with the
reference's retry ladders, PATCH with its one retry; KUBECONFIG or the in-cluster service
account). With --pod-source-file the pods come from a JSON file in the Kubernetes shape,
re-read on every Allocate, which is what the test harness drives.

Supervisor (gpumanager.go:33-111, watchers.go): the plugin instance is torn down, devices
re-enumerated and registration repeated when kubelet.sock is re-created in the
device-plugins directory (kubelet restart) or on SIGHUP; SIGQUIT dumps every thread's
stack to /etc/kubernetes/go_<timestamp>.txt; SIGINT/SIGTERM stop the plugin and exit 0.

    python -m gpushare_b200.daemon --memory-unit=MiB --health-check --pod-source-file pods.json
"""
import argparse
import ctypes
import faulthandler
import json
import os
import select
import signal
import struct
import sys
import threading
import time

from . import _lib as L
from . import api
from .grpc_harness import KUBELET_SOCKET, SERVER_SOCK, DevicePluginServer


_GPU_MEMORY = 0          # nvidia.go:13


def parse_args(argv=None):
    """The ten flags of cmd/nvidia/main.go:15-26 with the same defaults."""
    ap = argparse.ArgumentParser(prog="gpushare-device-plugin-v2", allow_abbrev=False)

    def flag(name, **kw):
        ap.add_argument("-" + name, "--" + name, **kw)

    def boolflag(name, help):
        # Go: -x, -x=true, -x=false
        flag(name, nargs="?", const="true", default="false", help=help)

    boolflag("mps", "Enable or Disable MPS")                      # stored, never read (main.go:16)
    boolflag("health-check", "Enable or disable Health check")
    flag("memory-unit", default="GiB", help="Set memoryUnit of the GPU Memroy, support 'GiB' and 'MiB'")
    boolflag("query-kubelet", "Query pending pods from kubelet instead of kube-apiserver")
    flag("kubelet-address", default="0.0.0.0", help="Kubelet IP Address")
    flag("kubelet-port", type=int, default=10250, help="Kubelet listened Port")
    flag("client-cert", default="", help="Kubelet TLS client certificate")
    flag("client-key", default="", help="Kubelet TLS client key")
    flag("token", default="", help="Kubelet client bearer token")
    flag("timeout", type=int, default=10, help="Kubelet client http timeout duration")
    # glog flags the DaemonSet passes (device-plugin-ds.yaml:29-33); accepted, verbosity only
    boolflag("logtostderr", "log to stderr")
    flag("v", type=int, default=0, help="log verbosity")
    # additions (not in the reference)
    flag("pod-source-file", default="", help="JSON file with the node's pods (stands in for the lister)")
    flag("nvml-lib", default="", help="path of libnvidia-ml (tests: the fake NVML)")
    flag("socket-dir", default=os.path.dirname(SERVER_SOCK), help="device-plugins directory")
    boolflag("probe", "(native binary only) sweep each granted slice on the GPU after Allocate")
    flag("health-probe-interval", type=int, default=0, help="(native binary only)")
    boolflag("no-register", "serve without registering with kubelet (harness)")
    flag("transport", default="native", choices=("native", "grpcio"),
         help="gRPC transport: the library's own HTTP/2 server (native) or Python grpcio")
    a = ap.parse_args(argv)
    for k in ("mps", "health_check", "query_kubelet", "logtostderr", "probe", "no_register"):
        v = str(getattr(a, k)).lower()
        if v not in ("true", "false", "1", "0", "t", "f"):
            ap.error("invalid boolean value %r for -%s" % (getattr(a, k), k.replace("_", "-")))
        setattr(a, k, v in ("true", "1", "t"))
    if a.probe or a.health_probe_interval:
        ap.error("the slice probe runs in the native daemon (bin/gpushare-device-plugin-v2) only; "
                 "this Python twin is the test harness of the host logic")
    return a


class PodSourceFile:
    """Stand-in lister: the pods of this node, re-read per Allocate like a LIST."""

    def __init__(self, path):
        self.path = path

    def list(self):
        if not self.path:
            return []
        with open(self.path) as f:
            doc = json.load(f)
        items = doc.get("items", doc) if isinstance(doc, dict) else doc
        return [api.Pod.from_k8s(p) for p in items]


def log(v, level, msg):
    if v >= level:
        print(msg, file=sys.stderr, flush=True)


class FSWatcher:
    """newFSWatcher (watchers.go:10-25): inotify on the device-plugins directory."""
    IN_CREATE = 0x100

    def __init__(self, path):
        self.libc = ctypes.CDLL(None, use_errno=True)
        self.fd = self.libc.inotify_init1(os.O_NONBLOCK | os.O_CLOEXEC)
        if self.fd < 0:
            raise OSError(ctypes.get_errno(), "inotify_init1")
        if self.libc.inotify_add_watch(self.fd, path.encode(), self.IN_CREATE) < 0:
            err = ctypes.get_errno()
            os.close(self.fd)
            raise OSError(err, "inotify_add_watch(%s)" % path)

    def events(self):
        """-> [(file name, mask)] that arrived since the last call."""
        out = []
        try:
            buf = os.read(self.fd, 65536)
        except BlockingIOError:
            return out
        off = 0
        while off + 16 <= len(buf):
            _wd, mask, _cookie, ln = struct.unpack_from("iIII", buf, off)
            name = buf[off + 16:off + 16 + ln].split(b"\0", 1)[0].decode()
            out.append((name, mask))
            off += 16 + ln
        return out

    def close(self):
        os.close(self.fd)


class PluginInstance:
    """One NewNvidiaDevicePlugin + Serve (gpumanager.go:70-79): everything that is rebuilt on
    a restart."""

    def __init__(self, a, nv, unit, node):
        self.a, self.nv = a, nv
        infos = nv.devices()
        self.infos = infos
        # nvidia.go:12-15,67-69: gpuMemory is a process global, set once from the first GPU and kept
        # across plugin restarts; a restart re-reads UUIDs, never the unit count
        global _GPU_MEMORY
        if _GPU_MEMORY == 0:
            self.table = api.DeviceTable.from_gpus(infos, unit)     # getDevices
            _GPU_MEMORY = self.table.units_per_gpu
        else:
            self.table = api.DeviceTable([i.uuid.decode() for i in infos], _GPU_MEMORY, unit)
        self.source = None
        if a.pod_source_file:
            self.plugin = api.NvidiaDevicePlugin(self.table, node)
            self.source = PodSourceFile(a.pod_source_file)
        else:
            from . import k8s_io
            token = a.token
            if not (a.client_cert or a.client_key or token):        # main.go:28-36
                try:
                    token = open(k8s_io.SA_DIR + "/token").read()
                except OSError as e:
                    sys.exit("in cluster mode, find token failed, error: %s" % e)
            transport = k8s_io.HttpTransport(
                k8s_io.ApiConfig.load(), "https://%s:%d" % (a.kubelet_address, a.kubelet_port), token,
                a.client_cert or None, a.client_key or None, timeout=float(a.timeout))
            try:                                                    # server.go:57-64
                k8s_io.patch_gpu_count(transport, node, len(infos))
                cgpu = k8s_io.disable_cgpu_isolation_or_not(transport, node)
            except Exception as e:
                log(a.v, 0, "Failed to get device plugin due to %s" % e)
                sys.exit(1)                                         # gpumanager.go:71-73
            self.plugin = k8s_io.ClusterDevicePlugin(
                self.table, node, k8s_io.PodLister(transport, node, a.query_kubelet), transport, cgpu)

        if self.source is not None:
            self._allocate = self.plugin.Allocate
            self.plugin.Allocate = self.allocate_from_file
        if a.transport == "native":
            from .native_transport import NativeDevicePluginServer as server_cls
        else:
            server_cls = DevicePluginServer
        self.server = server_cls(self.plugin, os.path.join(a.socket_dir, os.path.basename(SERVER_SOCK)))
        self.stop = threading.Event()

    def allocate_from_file(self, request):
        self.plugin.set_pods(self.source.list())     # the file stands in for the lister: re-read per Allocate
        return self._allocate(request)

    def Serve(self):
        a = self.a
        if a.no_register:
            self.server.Start()
        else:
            self.server.Serve(os.path.join(a.socket_dir, os.path.basename(KUBELET_SOCKET)))
        log(a.v, 1, "Starting to serve on %s" % self.server.socket)
        if a.health_check:                                          # server.go:211-229
            bad_mask = self.nv.health_open()
            for g in range(len(self.infos)):
                if bad_mask >> g & 1:
                    self.plugin.unhealthy(g)                        # nvidia.go:107-111

            def pump():
                while not self.stop.is_set():
                    ev = self.nv.health_wait(1000)
                    if ev and not ev[2]:
                        self.plugin.unhealthy(ev[0])
            self.pump_thread = threading.Thread(target=pump, daemon=True)
            self.pump_thread.start()

    def Stop(self):
        self.stop.set()
        self.server.Stop()
        if self.a.health_check:
            self.pump_thread.join(3)
            self.nv.health_close()


def coredump(path):
    """coredump.go:10-30, for threads instead of goroutines."""
    try:
        with open(path, "w") as f:
            faulthandler.dump_traceback(file=f, all_threads=True)
    except OSError as e:
        print("failed to write %s: %s" % (path, e), file=sys.stderr)


def main(argv=None):
    a = parse_args(argv)
    unit, warned = api.translate_memory_units(a.memory_unit)       # main.go:67-78
    if warned:
        log(a.v, 0, "Unsupported memory unit: %s, use memoryUnit Gi as default" % a.memory_unit)
    node = os.environ.get("NODE_NAME", "")
    if not node:
        sys.exit("Please set env NODE_NAME")                        # podmanager.go:55-58
    log(a.v, 1, "Start gpushare device plugin")
    # ---- sharedGPUManager.Run (gpumanager.go:33-111) ---------------------------------
    log(a.v, 1, "Loading NVML")
    try:
        nv = api.Nvml(a.nvml_lib or None)                           # :36
    except L.GpushareError as e:
        log(a.v, 1, "Failed to initialize NVML: %s." % e)
        log(a.v, 1, "If this is a GPU node, did you set the docker default runtime to `nvidia`?")
        threading.Event().wait()                                    # select {} (:39)
    log(a.v, 1, "Fetching devices.")
    if nv.device_count() == 0:
        log(a.v, 1, "No devices found. Waiting indefinitely.")
        threading.Event().wait()                                    # :44-47
    log(a.v, 1, "Starting FS watcher.")
    watcher = FSWatcher(a.socket_dir)                               # :50
    log(a.v, 1, "Starting OS watcher.")
    rd, wr = os.pipe()                                              # newOSWatcher (watchers.go:27-32)
    os.set_blocking(wr, False)
    for s in (signal.SIGHUP, signal.SIGINT, signal.SIGTERM, signal.SIGQUIT):
        signal.signal(s, lambda signum, _f: os.write(wr, bytes([signum])))
    kubelet_sock = os.path.basename(KUBELET_SOCKET)
    restart, inst = True, None
    while True:
        if restart:                                                 # :66-79
            if inst is not None:
                inst.Stop()
            inst = PluginInstance(a, nv, unit, node)
            try:
                inst.Serve()
            except Exception as e:
                log(a.v, 0, "Failed to start device plugin due to %s" % e)
                sys.exit(2)                                         # :74-76
            restart = False
        ready, _, _ = select.select([watcher.fd, rd], [], [], 1.0)
        if watcher.fd in ready:
            for name, mask in watcher.events():                     # :82-87
                if name == kubelet_sock and mask & FSWatcher.IN_CREATE:
                    log(a.v, 1, "inotify: %s created, restarting." % KUBELET_SOCKET)
                    restart = True
        if rd in ready:
            for signum in os.read(rd, 64):
                if signum == signal.SIGHUP:                         # :94-96
                    log(a.v, 1, "Received SIGHUP, restarting.")
                    restart = True
                elif signum == signal.SIGQUIT:                      # :97-101
                    log(a.v, 0, "generate core dump")
                    coredump("/etc/kubernetes/go_" + time.strftime("%Y%m%d%H%M%S") + ".txt")
                else:                                               # :102-106
                    log(a.v, 1, 'Received signal "%d", shutting down.' % signum)
                    inst.Stop()
                    watcher.close()
                    log(a.v, 1, "Shutdown of NVML returned: %s" % nv.close())
                    return 0


if __name__ == "__main__":
    sys.exit(main())

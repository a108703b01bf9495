"""gpushare-device-plugin-v2 -- the daemon entry point (cmd/nvidia/main.go, gpumanager.go).

Keeps the reference's command line (flag names, defaults, single- or double-dash forms as
Go's `flag` accepts) and environment (NODE_NAME mandatory, podmanager.go:55-58), and wires
the hot path: NVML enumerate -> device table -> plugin -> gRPC server -> Register ->
XID health watch -> probe at Allocate time.

Pod source: by default the live one (k8s_io: kubelet /pods or apiserver LIST with the
reference's retry ladders, PATCH with its one retry; KUBECONFIG or the in-cluster service
account). With --pod-source-file the pods come from a JSON file in the Kubernetes shape,
re-read on every Allocate, which is what the test harness drives. The fsnotify/SIGHUP
supervisor loop (gpumanager.go:63-108) is not built.

    python -m gpushare_b200.daemon --memory-unit=MiB --health-check --pod-source-file pods.json
"""
import argparse
import json
import os
import signal
import sys
import threading

from . import _lib as L
from . import api
from .grpc_harness import KUBELET_SOCKET, SERVER_SOCK, DevicePluginServer


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
    boolflag("probe", "sweep each granted slice on the GPU after Allocate")
    boolflag("no-register", "serve without registering with kubelet (harness)")
    a = ap.parse_args(argv)
    for k in ("mps", "health_check", "query_kubelet", "logtostderr", "probe", "no_register"):
        v = str(getattr(a, k)).lower()
        if v not in ("true", "false", "1", "0", "t", "f"):
            ap.error("invalid boolean value %r for -%s" % (getattr(a, k), k.replace("_", "-")))
        setattr(a, k, v in ("true", "1", "t"))
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


def main(argv=None):
    a = parse_args(argv)
    unit, warned = api.translate_memory_units(a.memory_unit)       # main.go:67-78
    if warned:
        log(a.v, 0, "Unsupported memory unit: %s, use memoryUnit Gi as default" % a.memory_unit)
    node = os.environ.get("NODE_NAME", "")
    if not node:
        sys.exit("Please set env NODE_NAME")                        # podmanager.go:55-58
    log(a.v, 1, "Start gpushare device plugin")
    log(a.v, 1, "Loading NVML")
    try:
        nv = api.Nvml(a.nvml_lib or None)                           # gpumanager.go:36
    except L.GpushareError as e:
        log(a.v, 1, "Failed to initialize NVML: %s." % e)
        log(a.v, 1, "If this is a GPU node, did you set the docker default runtime to `nvidia`?")
        threading.Event().wait()                                    # select {} (gpumanager.go:39)
    log(a.v, 1, "Fetching devices.")
    infos = nv.devices()
    if len(infos) == 0:
        log(a.v, 1, "No devices found. Waiting indefinitely.")
        threading.Event().wait()                                    # gpumanager.go:44-47
    table = api.DeviceTable.from_gpus(infos, unit)                  # getDevices
    source = None
    if a.pod_source_file:
        plugin = api.NvidiaDevicePlugin(table, node)
        source = PodSourceFile(a.pod_source_file)
    else:
        from . import k8s_io
        token = a.token
        if not (a.client_cert or a.client_key or token):           # main.go:28-36
            try:
                token = open(k8s_io.SA_DIR + "/token").read()
            except OSError as e:
                sys.exit("in cluster mode, find token failed, error: %s" % e)
        transport = k8s_io.HttpTransport(
            k8s_io.ApiConfig.load(), "https://%s:%d" % (a.kubelet_address, a.kubelet_port), token,
            a.client_cert or None, a.client_key or None, timeout=float(a.timeout))
        try:                                                       # server.go:57-64
            k8s_io.patch_gpu_count(transport, node, len(infos))
            cgpu = k8s_io.disable_cgpu_isolation_or_not(transport, node)
        except Exception as e:
            log(a.v, 0, "Failed to get device plugin due to %s" % e)
            sys.exit(1)                                            # gpumanager.go:71-73
        plugin = k8s_io.ClusterDevicePlugin(
            table, node, k8s_io.PodLister(transport, node, a.query_kubelet), transport, cgpu)

    probes = {}
    if a.probe:
        from .probe import Probe, ProbeScheduler
        for info, o in zip(infos, nv.map_cuda_ordinals(infos)):
            if o >= 0:
                probes[info.nvml_index] = ProbeScheduler(
                    Probe(o), on_bad=lambda tag, v, g=info.nvml_index: plugin.unhealthy(g))

    unit_bytes = (1 << 20) if unit == api.MiB else (1 << 30)
    arenas = {}

    def refresh_pods_then_allocate(request, _orig=plugin.Allocate):
        if source is not None:
            plugin.set_pods(source.list())
        d = _orig(request)
        sched = probes.get(d.gpu_index)
        if sched is not None and d.kind == "matched":
            # exactly the granted range, capped by what is free right now; never lets a
            # failed allocation change the answer (parity)
            nbytes = d.pod_req_units * unit_bytes
            try:
                free, _ = sched.probe.mem_info()
                nbytes = min(nbytes, max(0, free - (1 << 30))) & ~0xFFFFF
                if nbytes:
                    old = arenas.pop(d.gpu_index, None)
                    sched.drain()
                    if old:
                        sched.probe.free(old)
                    arenas[d.gpu_index] = sched.probe.alloc(nbytes)
                    sched.grant(arenas[d.gpu_index], nbytes, d.pod_index & 0xFFFFFFFF, d.pod_index)
                    sched.pump()
            except L.GpushareError as e:
                log(a.v, 1, "probe skipped: %s" % e)
        return d

    plugin.Allocate = refresh_pods_then_allocate
    server = DevicePluginServer(plugin, os.path.join(a.socket_dir, os.path.basename(SERVER_SOCK)))
    if a.no_register:
        server.Start()
    else:
        try:
            server.Serve(os.path.join(a.socket_dir, os.path.basename(KUBELET_SOCKET)))
        except Exception as e:
            log(a.v, 0, "Failed to start device plugin due to %s" % e)
            sys.exit(2)                                             # gpumanager.go:74-76
    log(a.v, 1, "Starting to serve on %s" % server.socket)

    stop = threading.Event()
    if a.health_check:                                              # server.go:211-229
        bad_mask = nv.health_open()
        for g in range(len(infos)):
            if bad_mask >> g & 1:
                plugin.unhealthy(g)                                 # nvidia.go:107-111

        def pump():
            while not stop.is_set():
                ev = nv.health_wait(5000)                           # nvidia.go:125
                if ev and not ev[2]:
                    plugin.unhealthy(ev[0])
                for s in probes.values():
                    s.pump()
        threading.Thread(target=pump, daemon=True).start()

    def on_signal(signum, _frame):                                  # gpumanager.go:102-106
        log(a.v, 1, 'Received signal "%d", shutting down.' % signum)
        stop.set()
    for s in (signal.SIGINT, signal.SIGTERM):
        signal.signal(s, on_signal)
    stop.wait()
    server.Stop()
    nv.close()


if __name__ == "__main__":
    main()

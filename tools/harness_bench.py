"""Host-side metrics of SURVEY 8(d) that bench.py's single line does not carry:

  metric 1(ii)  Allocate decisions/s end to end over the unix-socket gRPC harness
  metric 2      ListAndWatch latency: stream open -> last byte; health event -> last byte
  config 5      health churn: XID events (79 on GPU 3, 31 ignored, one without a device)
                injected through the fake NVML while an Allocate loop runs

    python tools/harness_bench.py [--gpus 8] [--transport native|grpcio] > profiles/r01_harness_8gpu.json

--transport picks BOTH ends: "native" is the library's own HTTP/2 + gRPC (gs_server_* and
gs_grpc_client_*: nothing of the RPC runs in Python), "grpcio" is grpc-core driven from
Python on both sides (grpc_harness.py), the only transport the first measurements had.

No GPU needed (the probe is not involved); run on the B200 box for a like-for-like host.
The reference-side figures come from oracle/ref_model.cpp (cost-faithful restatement).
"""
import argparse
import ctypes as C
import json
import os
import statistics
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gpushare_b200 import api  # noqa: E402
from gpushare_b200.grpc_harness import DevicePluginServer, KubeletStub  # noqa: E402
from gpushare_b200.native_transport import NativeClient, NativeDevicePluginServer  # noqa: E402
from oracle import ref_cpp  # noqa: E402
from oracle import ref_model as R  # noqa: E402

FAKE = os.path.join(ROOT, "tests", "fake_nvml", "libfake_nvml.so")
UNITS = 183359


def pct(xs, p):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(round(p / 100.0 * (len(xs) - 1))))]


class GrpcioEnds:
    """grpc-core on both ends, driven from Python."""

    def __init__(self, plug, sock):
        self.srv = DevicePluginServer(plug, sock)
        self.srv.Start()
        self.ch = KubeletStub.connect(sock)
        self.stream = None

    def open_law(self):
        self.stream = KubeletStub.method(self.ch, "ListAndWatch", stream=True)(b"")

    def next_law(self):
        """-> (length, fetch) where fetch() returns the bytes (outside any timed interval)."""
        msg = next(self.stream)
        return len(msg), lambda: msg

    def cancel_law(self):
        self.stream.cancel()

    def allocate_caller(self):
        ch = KubeletStub.connect(self.srv.socket)
        return KubeletStub.method(ch, "Allocate"), ch.close

    def close(self):
        self.ch.close()
        self.srv.Stop()


class NativeEnds:
    """gs_server_* and gs_grpc_client_*: the RPC never enters Python on either end."""

    def __init__(self, plug, sock):
        self.srv = NativeDevicePluginServer(plug, sock)
        self.srv.Start()
        self.cl = NativeClient(sock)
        self.sid = None

    def open_law(self):
        self.sid = self.cl.open_stream("ListAndWatch")

    def next_law(self):
        ptr, n = self.cl.next_message(self.sid, timeout_ms=60000, copy=False)
        return n, lambda: C.string_at(ptr, n)

    def cancel_law(self):
        self.cl.cancel(self.sid)

    def allocate_caller(self):
        cl = NativeClient(self.srv.socket)
        return (lambda w: cl.call("Allocate", w)), cl.close

    def close(self):
        self.cl.close()
        self.srv.Stop()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--transport", choices=("native", "grpcio"), default="native")
    args = ap.parse_args()
    os.environ["FAKE_NVML_GPUS"] = str(args.gpus)
    os.environ["FAKE_NVML_MEM_MIB"] = str(UNITS)
    out = {"gpus": args.gpus, "unit": "MiB", "devices": args.gpus * UNITS,
           "transport": args.transport, "host": {"nproc": os.cpu_count()}}
    fake = C.CDLL(FAKE)
    fake.fake_nvml_push_event.argtypes = [C.c_int, C.c_ulonglong]
    nv = api.Nvml(FAKE)
    infos = nv.devices()
    t0 = time.perf_counter()
    table = api.DeviceTable.from_gpus(infos, api.MiB)
    out["table_build_ms"] = (time.perf_counter() - t0) * 1e3
    out["payload_bytes"] = table.payload_size()
    plug = api.NvidiaDevicePlugin(table, "node-a")
    d = tempfile.mkdtemp(prefix="gsh")
    ends = (NativeEnds if args.transport == "native" else GrpcioEnds)(
        plug, os.path.join(d, "aliyungpushare.sock"))

    # ---- metric 2: stream open -> full list ------------------------------------------
    lat = []
    for _ in range(5):
        t0 = time.perf_counter()
        ends.open_law()
        n, _fetch = ends.next_law()
        lat.append((time.perf_counter() - t0) * 1e3)
        assert n == out["payload_bytes"]
        ends.cancel_law()
    out["law_open_to_last_byte_ms"] = {"median": statistics.median(lat), "min": min(lat), "n": len(lat)}

    # ---- config 5: health churn under an Allocate loop --------------------------------
    uuids = [i.uuid.decode() for i in infos]
    pods = [api.Pod("u%d" % k, "default", "p%d" % k, "node-a", 1024, str(k % args.gpus),
                    str(1000 + k), "false") for k in range(64)]
    pod_list = api.PodList(pods)
    req = R.encode_allocate_request([["%s-_-%d" % (uuids[0], j) for j in range(1024)]])
    stop = threading.Event()
    alloc_count = [0]

    def alloc_loop():
        call, close = ends.allocate_caller()
        while not stop.is_set():
            plug.set_pods(pod_list)
            for _ in range(64):
                call(req)
                alloc_count[0] += 1
        close()

    nv.health_open()

    def pump():
        while not stop.is_set():
            ev = nv.health_wait(50)
            if ev and not ev[2]:
                plug.unhealthy(ev[0])
    threading.Thread(target=pump, daemon=True).start()
    loop = threading.Thread(target=alloc_loop, daemon=True)
    loop.start()
    ends.open_law()
    ends.next_law()
    readv = []
    script = [(3 % args.gpus, 79), (1 % args.gpus, 31), (-1, 48)] * 7
    t_start, a_start = time.perf_counter(), alloc_count[0]
    for gpu, xid in script:
        table.reset_health()
        plug.notify_health()
        ends.next_law()                                # the reset is a change too
        t0 = time.perf_counter()
        fake.fake_nvml_push_event(gpu, xid)
        if xid in (31, 43, 45):
            time.sleep(0.05)                           # must NOT re-advertise
            continue
        _n, fetch = ends.next_law()
        readv.append((time.perf_counter() - t0) * 1e3)
        want_bad = UNITS * (args.gpus if gpu < 0 else 1)
        got_bad = fetch().count(b"Unhealthy")
        assert got_bad == want_bad, (gpu, xid, got_bad)
    churn_s = time.perf_counter() - t_start
    stop.set()
    loop.join(10)
    ends.cancel_law()
    out["health_readvertise_ms"] = {"p50": pct(readv, 50), "p99": pct(readv, 99), "n": len(readv),
                                    "what": "fake-NVML XID push -> last byte of the full list at the "
                                            "client, Allocate loop running; one re-send per event"}
    out["allocate_under_churn_per_s"] = (alloc_count[0] - a_start) / churn_s
    table.reset_health()

    # ---- metric 1(ii): Allocate over gRPC ---------------------------------------------
    call, close_caller = ends.allocate_caller()
    grpc_rates = {}
    for n in (2, 1024, 4096, 16384, 40960):
        ids = ["%s-_-%d" % (uuids[j % args.gpus], j) for j in range(n)]
        w = R.encode_allocate_request([ids])
        pods_n = [api.Pod("u%d" % k, "default", "p%d" % k, "node-a", n, str(k % args.gpus),
                          str(1000 + k), "false") for k in range(64)]
        pl = api.PodList(pods_n)
        reps = 4 if args.transport == "grpcio" else (64 if n <= 4096 else 16)
        plug.set_pods(pl)
        for _ in range(32):
            call(w)                                    # warm the connection's buffers
        t0 = time.perf_counter()
        for _ in range(reps):
            plug.set_pods(pl)
            for _ in range(64):
                call(w)
        grpc_rates[str(n)] = {"decisions_per_s": reps * 64 / (time.perf_counter() - t0),
                              "request_bytes": len(w)}
    out["allocate_over_grpc"] = grpc_rates
    close_caller()
    ends.close()
    nv.health_close()
    nv.close()

    # ---- the reference model on the same box -------------------------------------------
    ref = ref_cpp.RefPlugin(uuids, UNITS, "MiB", "node-a")
    t0 = time.perf_counter()
    ref.get_devices()
    gd = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    nbytes = ref.l.ref_list_and_watch(ref.h, None, 0)
    ms = (time.perf_counter() - t0) * 1e3
    out["reference_model"] = {
        "get_devices_ms": gd, "marshal_per_send_ms": ms, "payload_bytes": nbytes,
        "resends_per_gpu_event_by_design": UNITS,
        "readvertise_one_gpu_event_s_extrapolated": UNITS * ms / 1e3,
        "note": "server.go:187-190 re-sends the full list once per fake device of the sick GPU; "
                "the per-send cost is measured, the product with 183 359 is arithmetic, not a run"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()

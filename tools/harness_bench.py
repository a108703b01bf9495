"""Host-side metrics of SURVEY 8(d) that bench.py's single line does not carry:

  metric 1(ii)  Allocate decisions/s end to end over the unix-socket gRPC harness
  metric 2      ListAndWatch latency: stream open -> last byte; health event -> last byte
  config 5      health churn: XID events (79 on GPU 3, 31 ignored, one without a device)
                injected through the fake NVML while an Allocate loop runs

    python tools/harness_bench.py [--gpus 8] > profiles/r01_harness_8gpu.json

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
from oracle import ref_cpp  # noqa: E402
from oracle import ref_model as R  # noqa: E402

FAKE = os.path.join(ROOT, "tests", "fake_nvml", "libfake_nvml.so")
UNITS = 183359


def pct(xs, p):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(round(p / 100.0 * (len(xs) - 1))))]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=8)
    args = ap.parse_args()
    os.environ["FAKE_NVML_GPUS"] = str(args.gpus)
    os.environ["FAKE_NVML_MEM_MIB"] = str(UNITS)
    out = {"gpus": args.gpus, "unit": "MiB", "devices": args.gpus * UNITS,
           "host": {"nproc": os.cpu_count()}}
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
    srv = DevicePluginServer(plug, os.path.join(d, "aliyungpushare.sock"))
    srv.Start()
    ch = KubeletStub.connect(srv.socket)

    # ---- metric 2: stream open -> full list ------------------------------------------
    lat = []
    for _ in range(5):
        t0 = time.perf_counter()
        stream = KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")
        msg = next(stream)
        lat.append((time.perf_counter() - t0) * 1e3)
        assert len(msg) == out["payload_bytes"]
        stream.cancel()
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
        ch2 = KubeletStub.connect(srv.socket)
        call = KubeletStub.method(ch2, "Allocate")
        while not stop.is_set():
            plug.set_pods(pod_list)
            for _ in range(64):
                call(req)
                alloc_count[0] += 1
        ch2.close()

    nv.health_open()

    def pump():
        while not stop.is_set():
            ev = nv.health_wait(50)
            if ev and not ev[2]:
                plug.unhealthy(ev[0])
    threading.Thread(target=pump, daemon=True).start()
    loop = threading.Thread(target=alloc_loop, daemon=True)
    loop.start()
    stream = KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")
    next(stream)
    readv = []
    script = [(3 % args.gpus, 79), (1 % args.gpus, 31), (-1, 48)] * 7
    t_start, a_start = time.perf_counter(), alloc_count[0]
    for gpu, xid in script:
        table.reset_health()
        next(stream)                                   # the reset is a change too
        t0 = time.perf_counter()
        fake.fake_nvml_push_event(gpu, xid)
        if xid in (31, 43, 45):
            time.sleep(0.05)                           # must NOT re-advertise
            continue
        msg = next(stream)
        readv.append((time.perf_counter() - t0) * 1e3)
        want_bad = UNITS * (args.gpus if gpu < 0 else 1)
        assert msg.count(b"Unhealthy") == want_bad, (gpu, xid, msg.count(b"Unhealthy"))
    churn_s = time.perf_counter() - t_start
    stop.set()
    loop.join(10)
    stream.cancel()
    out["health_readvertise_ms"] = {"p50": pct(readv, 50), "p99": pct(readv, 99), "n": len(readv),
                                    "what": "fake-NVML XID push -> last byte of the full list at the "
                                            "client, Allocate loop running; one re-send per event"}
    out["allocate_under_churn_per_s"] = (alloc_count[0] - a_start) / churn_s
    table.reset_health()

    # ---- metric 1(ii): Allocate over gRPC ---------------------------------------------
    call = KubeletStub.method(ch, "Allocate")
    grpc_rates = {}
    for n in (2, 1024, 4096, 16384, 40960):
        ids = ["%s-_-%d" % (uuids[j % args.gpus], j) for j in range(n)]
        w = R.encode_allocate_request([ids])
        pods_n = [api.Pod("u%d" % k, "default", "p%d" % k, "node-a", n, str(k % args.gpus),
                          str(1000 + k), "false") for k in range(64)]
        pl = api.PodList(pods_n)
        reps = max(1, 256 // 64)
        t0 = time.perf_counter()
        for _ in range(reps):
            plug.set_pods(pl)
            for _ in range(64):
                call(w)
        grpc_rates[str(n)] = {"decisions_per_s": reps * 64 / (time.perf_counter() - t0),
                              "request_bytes": len(w)}
    out["allocate_over_grpc"] = grpc_rates
    ch.close()
    srv.Stop()
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

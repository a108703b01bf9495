"""Diagnostic: Allocate decisions/s at the C ABI with and without the probe running."""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200 import api  # noqa: E402

uuid = "GPU-%08x-0000-0000-0000-%012x" % (0, 0)
pods_k8s, wires, sizes = bench.build_workload(0, uuid, 183359)
pods = [api.Pod.from_k8s(p) for p in pods_k8s]
table = api.DeviceTable([uuid], 183359, "MiB")
plug = api.NvidiaDevicePlugin(table, "node-a")
views, keep = api.pod_views(pods)
lib = L.load()
resp = C.create_string_buffer(4096)
res = L.AllocResult()


def wave():
    lib.gs_plugin_set_pods(plug.h, views, len(pods))
    for k, w in enumerate(wires):
        rc = lib.gs_plugin_allocate(plug.h, w, len(w), resp, 4096, C.byref(res))
        assert rc == 0 and res.kind == 0 and res.pod_index == k


def rate(n=200):
    for _ in range(10):
        wave()
    t0 = time.perf_counter()
    for _ in range(n):
        wave()
    return n * 8 / (time.perf_counter() - t0)


print("no GPU activity: %.0f decisions/s" % rate())
cnt = (C.c_uint32 * 4)()
nc, tot = C.c_uint32(), C.c_uint64()
for w in (wires[4], wires[5]):
    t0 = time.perf_counter()
    for _ in range(500):
        lib.gs_request_scan(w, len(w), cnt, 4, C.byref(nc), C.byref(tot))
    dt = (time.perf_counter() - t0) / 500
    print("scan %d ids: %.1f us, %.1f GB/s" % (tot.value, dt * 1e6, len(w) / dt / 1e9))
big = bytearray(64 << 20)
t0 = time.perf_counter()
for _ in range(10):
    b2 = bytes(big)
print("host memcpy %.1f GB/s" % (10 * len(big) / (time.perf_counter() - t0) / 1e9))
if len(sys.argv) > 1:
    from gpushare_b200.probe import Probe, ProbeScheduler
    p = Probe(0)
    print("probe open: %.0f decisions/s" % rate())
    arena = p.alloc(64 << 30)
    sl = [(arena, 64 << 30, 1)]
    p.submit(sl)
    r = rate(40)
    p.wait()
    print("probe running (no polling): %.0f decisions/s" % r)
    s = ProbeScheduler(p)
    t0 = time.perf_counter()
    for _ in range(100):
        wave()
        s.grant(arena, 64 << 30, 1)
        s.pump()
    print("with scheduler: %.0f decisions/s" % (800 / (time.perf_counter() - t0)))
    s.drain()

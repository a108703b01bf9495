"""BASELINE config 3 sweep at the C ABI: 8 x B200 at MiB granularity (1 466 872 ids),
requests of n ids drawn without replacement from the free pool, pod sets of 1 / 8 / 64.
Prints decisions/s for this library and for oracle/ref_model.cpp (the reference algorithm
restated with its per-unit costs) on the same host, single thread both.

    python tools/allocate_sweep.py > profiles/r01_allocate_sweep.json
"""
import ctypes as C
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200 import api  # noqa: E402
from oracle import ref_cpp  # noqa: E402
from oracle import ref_model as R  # noqa: E402

UNITS, GPUS, NODE = 183359, 8, "node-a"
uuids = ["GPU-%08x-0000-0000-0000-%012x" % (i, i) for i in range(GPUS)]
rng = random.Random(20261106)
table = api.DeviceTable(uuids, UNITS, api.MiB)
lib = L.load()
drv = C.CDLL(os.path.join(ROOT, "oracle", "libgs_oracle.so"))   # C timing loops (oracle/driver.c)
rows = []
for n in (2, 1024, 4096, 16384, 40960):      # 2 = the GiB-unit binpack-1 request of configs[0]/[1]
    pool = rng.sample(range(GPUS * UNITS), n)
    wire = R.encode_allocate_request([["%s-_-%d" % (uuids[x // UNITS], x % UNITS) for x in pool]])
    for npods in (1, 8, 64):
        # every pod is a candidate of the requested size with a distinct assume time: each
        # decision claims the oldest; the list is re-installed when all are claimed
        k8s = [{"metadata": {"uid": "u%d" % k, "name": "p%d" % k, "namespace": "default",
                             "annotations": {"ALIYUN_COM_GPU_MEM_IDX": str(k % GPUS),
                                             "ALIYUN_COM_GPU_MEM_ASSUME_TIME": str(10**18 + 31 * ((k * 37) % npods)),
                                             "ALIYUN_COM_GPU_MEM_ASSIGNED": "false"}},
                "spec": {"nodeName": NODE, "containers": [
                    {"name": "c", "resources": {"limits": {"aliyun.com/gpu-mem": str(n)}}}]},
                "status": {"phase": "Pending"}} for k in range(npods)]
        plist = api.PodList([api.Pod.from_k8s(p) for p in k8s])
        plug = api.NvidiaDevicePlugin(table, NODE)
        resp = C.create_string_buffer(4096)
        res = L.AllocResult()
        ref = ref_cpp.RefPlugin(uuids, UNITS, "MiB", NODE)
        rbuf = C.create_string_buffer(1 << 16)
        rn = C.c_size_t()

        failed = C.c_uint64()
        drv.gs_drive_rounds.restype = C.c_double
        o = drv.gs_drive_rounds(C.cast(lib.gs_plugin_set_pods, C.c_void_p), C.cast(lib.gs_plugin_allocate, C.c_void_p),
                                plug.h, plist.views, C.c_size_t(plist.n), wire, C.c_size_t(len(wire)), npods,
                                C.c_double(0.6), C.byref(failed))
        assert failed.value == 0, failed.value
        ref.set_pods_k8s(k8s)
        ref.save_pods()
        drv.ref_drive_rounds.restype = C.c_double
        r = drv.ref_drive_rounds(C.c_void_p(ref.h), wire, C.c_size_t(len(wire)), npods, C.c_double(0.8),
                                 C.byref(failed))
        assert failed.value == 0, failed.value
        # same answers on both sides for this workload
        lib.gs_plugin_set_pods(plug.h, plist.views, plist.n)
        ref.restore_pods()
        lib.gs_plugin_allocate(plug.h, wire, len(wire), resp, 4096, C.byref(res))
        kind, rresp, rp = ref.allocate(wire, 1)
        assert R.decode_allocate_response(C.string_at(resp, res.resp_len)) == R.decode_allocate_response(rresp)
        assert res.pod_index == rp
        rows.append({"ids": n, "request_bytes": len(wire), "pods": npods, "ours_per_s": o, "reference_model_per_s": r,
                     "ratio": o / r, "ours_request_gbps": o * len(wire) / 1e9})
        print("ids %6d pods %2d: ours %9.0f/s  ref %8.0f/s  x%.1f  (scan %.1f GB/s)" % (
            n, npods, o, r, o / r, o * len(wire) / 1e9), file=sys.stderr, flush=True)
print(json.dumps({"config": "BASELINE configs[2]: 8 x B200, MiB units, %d fake ids" % (GPUS * UNITS),
                  "host": {"nproc": os.cpu_count()}, "threads": 1,
                  "timing": "C loops on both sides (oracle/driver.c, ref_drive_rounds): no interpreter in the timed region",
                  "rows": rows}, indent=1))

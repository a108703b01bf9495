#!/usr/bin/env python
"""bench.py -- the gpushare hot path on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Metric (BASELINE.json): Allocate decisions/s at MiB granularity, with the slice probe's
HBM GB/s as the kernel roofline.

Two shapes are measured, and both are printed:

  per shard (`value`, `e2e`, `over_socket`): one rank = one GPU's shard of the node -- its
    183 359 fake-device ids, the tenant pods the scheduler extender bound to it, its probe
    engine. Nothing is shared between shards (no collective on the data path): weak scaling.
  single queue (`single_queue`, rank 0, every N): the reference's own deployment shape -- ONE
    process, ONE plugin and ONE lock (allocate.go:60-61) over an N-GPU table (N x 183 359 ids),
    one request stream carrying the tenants of all N GPUs, N probe engines fed from it.
    `--impl reference` prints the reference model on the very same N-GPU stream beside it.

A step = WAVES_PER_STEP admission waves. One wave = the lister's pod list is refreshed
(64 pods per GPU: 8 pending tenants, sizes 1-40 GiB as BASELINE config 4 draws them, plus 56
already-assigned pods) and kubelet's 8 AllocateRequests arrive, each carrying n x 1024 device
ids (n GiB at MiB granularity, 51 KB - 2 MB of wire bytes). Every decision = decode request +
candidate filter/sort/match + response and patch encode + patch applied (SURVEY 8(d) metric
1). Every grant is queued for the slice probe, which sweeps exactly the granted range on the
GPU beside the decisions, with a seed of its own per sweep, and never gates them.

  value : decisions/s at the bare C ABI (gs_plugin_allocate, caller-owned buffers), probe
          engine running on HBM-resident ranges. The timed region runs from the first
          decision to the end of the last probe batch's drain (`decision_clock_s` and
          `drain_ms` say how it splits); at the default step size 20 steps are > 2 s.
  e2e   : the same through the package's public API (NvidiaDevicePlugin.Allocate: bytes
          in, bytes out) with the probe scheduler, result read-back included.
  over_socket : the same with kubelet's bytes arriving over the unix socket (the drop-in
          boundary, server.go:114-142), with the reference model served through the same
          transport beside it.
  roofline : probe kernels, algorithmic bytes (1 B stored + 1 B loaded per slice byte)
          over CUDA-event time on the probe's stream, against MEASURED_PEAKS.json.
  cpu_baseline / --impl reference : oracle/ref_model.cpp (the reference algorithm restated
          with the reference's per-unit costs; the Go binary cannot be built here), one
          thread because the reference serialises Allocate under one lock.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NODE = "node-a"
UNITS_B200_MIB = 183359          # nvidia-smi / NVML total of a B200 in MiB
TENANTS = 8                      # config 4: 64 tenants over 8 GPUs
PODS_IN_LIST = 64                # per GPU
SMALL = bool(os.environ.get("GS_BENCH_SMALL"))        # CPU tests: MiB-sized tenants, short steps
WAVES_PER_STEP = 8 if SMALL else 640                  # 5120 decisions per step: 20 steps > 2 s
SEED = 20261106


# ---------------------------------------------------------------- workload ------------
def tenant_sizes_gib(rank):
    """BASELINE config 4 per-GPU share: 8 sizes uniform in 1..40 GiB, redrawn until they fit
    one B200 (sum <= 160 GiB so the probed ranges fit beside the CUDA context). Every GPU gets
    the SAME eight sizes (weak scaling: the work per GPU is fixed; only the ids, the UUID and the
    pod names differ from shard to shard), so `rank` does not enter the draw."""
    import random
    del rank
    rng = random.Random(SEED)
    if SMALL:
        return rng.sample(range(1, 40), TENANTS)
    while True:
        s = [rng.randint(1, 40) for _ in range(TENANTS)]
        if len(set(s)) == TENANTS and sum(s) <= 160:   # distinct sizes: the size picks the pod
            return s


def _varint(n):
    out = bytearray()
    while n >= 0x80:
        out.append((n & 0x7F) | 0x80)
        n >>= 7
    out.append(n)
    return bytes(out)


def encode_allocate_request(containers):
    """What kubelet sends: AllocateRequest{container_requests=1{devices_ids=1}} (synthetic input;
    deliberately not taken from oracle/, which only the baseline legs may touch)."""
    out = bytearray()
    for ids in containers:
        inner = bytearray()
        for i in ids:
            b = i.encode()
            inner += b"\x0a" + _varint(len(b)) + b
        out += b"\x0a" + _varint(len(inner)) + inner
    return bytes(out)


def build_workload(rank, uuid, units, gpu_index=0, extra_units=0):
    """One GPU's shard -> (pods as k8s dicts, request wires in arrival order, sizes GiB,
    units per tenant). `gpu_index` is the ALIYUN_COM_GPU_MEM_IDX the extender wrote;
    `extra_units` (the single-queue node: + gpu index) keeps tenant sizes distinct across the
    GPUs of a node, because the request's size is what picks the pod (allocate.go:79-84)."""
    import numpy as np
    sizes = tenant_sizes_gib(rank)
    per = 16 if SMALL else 1024                        # units per size step
    rng = np.random.default_rng(SEED + 1000 * rank)
    free = rng.permutation(units)                      # ids drawn without replacement
    wires, off, tenant_units = [], 0, []
    for gib in sizes:
        n = gib * per + extra_units
        ids = ["%s-_-%d" % (uuid, j) for j in free[off:off + n]]
        off += n
        wires.append(encode_allocate_request([ids]))
        tenant_units.append(n)
    pods = []
    for k, n in enumerate(tenant_units):
        pods.append({"metadata": {"uid": "tenant-%d-%d" % (rank, k), "name": "tenant-%d-%d" % (rank, k),
                                  "namespace": "default",
                                  "annotations": {"ALIYUN_COM_GPU_MEM_IDX": str(gpu_index),
                                                  "ALIYUN_COM_GPU_MEM_ASSUME_TIME": str(10**18 + 977 * k + rank),
                                                  "ALIYUN_COM_GPU_MEM_ASSIGNED": "false"}},
                     "spec": {"nodeName": NODE, "containers": [
                         {"name": "main", "resources": {"limits": {"aliyun.com/gpu-mem": str(n)}}}]},
                     "status": {"phase": "Pending"}})
    for k in range(PODS_IN_LIST - TENANTS):
        pods.append({"metadata": {"uid": "bg-%d-%d" % (rank, k), "name": "running-%d-%d" % (rank, k),
                                  "namespace": "default",
                                  "annotations": {"ALIYUN_COM_GPU_MEM_IDX": str(gpu_index),
                                                  "ALIYUN_COM_GPU_MEM_ASSUME_TIME": str(10**17 + k),
                                                  "ALIYUN_COM_GPU_MEM_ASSIGNED": "true"}},
                     "spec": {"nodeName": NODE, "containers": [
                         {"name": "main", "resources": {"limits": {"aliyun.com/gpu-mem": str(1024 * (1 + k % 7))}}}]},
                     "status": {"phase": "Pending"}})
    return pods, wires, sizes, tenant_units


def build_node_workload(n_gpus, uuids, units):
    """The single-queue node: the shards of all N GPUs behind one lister and one request stream.
    -> (pods, [(gpu, tenant k, wire)] in arrival order: GPUs interleaved, sizes GiB per GPU)."""
    pods, stream, sizes = [], [], []
    per_gpu = []
    for g in range(n_gpus):
        p, wires, sz, _tu = build_workload(g, uuids[g], units, gpu_index=g, extra_units=g)
        base = len(pods)
        pods += p
        sizes.append(sz)
        per_gpu.append((base, wires))
    for k in range(TENANTS):
        for g in range(n_gpus):
            base, wires = per_gpu[g]
            stream.append((g, k, base + k, wires[k]))
    return pods, stream, sizes


def workload_config(sizes, waves, req_avg):
    return {"workload": "BASELINE configs[2]/[3] per-GPU shard: 1xB200 at MiB granularity "
                        "(183359 fake ids), 8 tenant pods %s GiB + 56 assigned pods in the lister's "
                        "list, requests of n*1024 ids; probe of each granted range" % sizes,
            "waves_per_step": waves, "decisions_per_step": waves * TENANTS,
            "request_bytes_avg": req_avg, "l2": "probed ranges (>= 100 GiB per batch) exceed the 126 MB L2",
            "parallelism": "value/e2e/over_socket: one process per GPU, no collective, each rank decides for its "
                           "own GPU's shard; single_queue: one process, one plugin and lock over the N-GPU node "
                           "(the reference's shape, allocate.go:60-61)"}


# ---------------------------------------------------------------- host placement ------
def physical_cores():
    """-> list of cpu-id tuples, one tuple of hyper-thread siblings per physical core, restricted
    to the cpus this process may run on."""
    allowed = sorted(os.sched_getaffinity(0))
    seen, cores = set(), []
    for c in allowed:
        if c in seen:
            continue
        sib = (c,)
        try:
            txt = open("/sys/devices/system/cpu/cpu%d/topology/thread_siblings_list" % c).read().strip()
            ids = set()
            for part in txt.split(","):
                a, _, b = part.partition("-")
                ids.update(range(int(a), int(b or a) + 1))
            sib = tuple(sorted(i for i in ids if i in allowed)) or (c,)
        except (OSError, ValueError):
            pass
        seen.update(sib)
        cores.append(sib)
    return cores


def pin_rank(local_rank, world):
    """Each rank gets physical cores of its own (all their hyper-threads, never shared with
    another rank): the Python loop, the library's serving thread of the over-socket leg and
    NCCL's helper threads of one rank then contend with nobody else's. The last core is left to
    the nvidia-smi sampler and whatever else the box runs. Returns a description for the line."""
    global _BOX_CORES
    cores = physical_cores()
    _BOX_CORES = len(cores)               # counted before the affinity mask shrinks the view
    usable = cores[:-1] if len(cores) > world else cores
    per = max(1, min(2, len(usable) // max(world, 1)))
    # spread the ranks over the whole socket instead of packing them into its first cores: the
    # request scan is memory traffic, and neighbours share mesh stops and L3 slices
    stride = max(per, len(usable) // max(world, 1))
    mine = usable[local_rank * stride:local_rank * stride + per] or [usable[local_rank % len(usable)]]
    cpus = sorted(c for core in mine for c in core)
    try:
        os.sched_setaffinity(0, cpus)
    except OSError:
        return {"pinned": False}
    return {"pinned": True, "cpus": cpus, "physical_cores_on_box": len(cores), "cores_per_rank": per,
            "stride": stride,
            "spare_cpus": list(cores[-1]) if len(cores) > world else []}


# ---------------------------------------------------------------- clocks --------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, cpus=None):
        def place():
            if cpus:
                try:
                    os.sched_setaffinity(0, cpus)
                except OSError:
                    pass
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
                                      preexec_fn=place)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        out = self.p.communicate()[0]
        sm, mx, reasons, watts = [], 0, set(), []
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
                watts.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [x for x in sm if x > 0.5 * mx] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm),
                "power_w_median": statistics.median(watts) if watts else None}


def host_memcpy_gbps(n=256 << 20, reps=3):
    """Single-thread memcpy of a buffer larger than the L3 (bytes copied / s, reads + writes not
    double counted): the host-side yardstick for the request scan and the payload copy."""
    import ctypes
    a, b = ctypes.create_string_buffer(n), ctypes.create_string_buffer(n)
    ctypes.memmove(b, a, n)
    best = 0.0
    for _ in range(reps):
        t0 = time.perf_counter()
        ctypes.memmove(b, a, n)
        best = max(best, n / (time.perf_counter() - t0) / 1e9)
    return best


_BOX_CORES = None


def host_info():
    model = ""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return {"nproc": os.cpu_count(), "cpu": model,
            "physical_cores": _BOX_CORES if _BOX_CORES is not None else len(physical_cores())}


# ---------------------------------------------------------------- reference model -----
class RefNode:
    """oracle/ref_model.cpp set up on a workload: in-process calls, or served through a socket.
    Only bench.py's baseline legs construct this (the one place outside tests/ that may touch
    oracle/)."""

    def __init__(self, uuids, units, pods):
        import ctypes as C

        from oracle import ref_cpp
        self.C = C
        self.ref = ref_cpp.RefPlugin(uuids, units, "MiB", NODE)
        self.lib, self.h = self.ref.l, self.ref.h
        self.ref.set_pods_k8s(pods)
        self.ref.save_pods()
        self.buf = C.create_string_buffer(1 << 16)
        self.n = C.c_size_t()

    def wave(self, wires):
        self.ref.restore_pods()          # the lister's answer, re-installed in C++ (no ctypes cost)
        for w in wires:
            kind = self.lib.ref_allocate(self.h, w, len(w), 1, self.buf, len(self.buf), self.C.byref(self.n))
            if kind != 0:
                raise RuntimeError("reference model: unexpected decision kind %d" % kind)

    def run_for(self, wires, budget_s, min_waves=1):
        waves, t0 = 0, time.perf_counter()
        while waves < min_waves or time.perf_counter() - t0 < budget_s:
            self.wave(wires)
            waves += 1
        return waves, time.perf_counter() - t0

    def serve_native(self):
        """The reference model behind the product's own transport: gs_server_config.allocate
        points at oracle's ref_server_allocate (C++ all the way, no Python on the serving side).
        -> (server handle, socket path, a callable that stops it)."""
        C = self.C
        from gpushare_b200 import _lib as L
        from gpushare_b200 import api
        lib = L.load()
        self._table = api.DeviceTable(["GPU-00000000-0000-0000-0000-000000000000"], 16, api.MiB)
        self._plug = api.NvidiaDevicePlugin(self._table, NODE)   # the server wants a plugin; its Allocate is replaced
        d = tempfile.mkdtemp(prefix="gsr")
        sock = os.path.join(d, "aliyungpushare.sock")
        cfg = L.ServerConfig()
        cfg.socket_path = sock.encode()
        fn = C.cast(self.lib.ref_server_allocate, L.ServerAllocateFn)
        cfg.allocate = fn
        cfg.user = self.h
        h = C.c_void_p()
        L.check(lib.gs_server_start(self._plug.h, C.byref(cfg), C.byref(h)), "gs_server_start")

        def stop():
            lib.gs_server_stop(h)
            try:
                os.rmdir(d)
            except OSError:
                pass
        return sock, stop, (cfg, fn)

    def over_socket(self, wires, budget_s, min_waves=1):
        """Requests over the unix socket, the library's client standing in for kubelet, the
        reference model answering: the matched leg for our `over_socket`."""
        C = self.C
        from gpushare_b200 import _lib as L
        from gpushare_b200.native_transport import NativeClient
        lib = L.load()
        sock, stop, keep = self.serve_native()
        cl = NativeClient(sock)
        resp = C.create_string_buffer(1 << 16)
        rn, rstatus, rmsg = C.c_size_t(), C.c_int32(), C.create_string_buffer(128)
        path = b"/v1beta1.DevicePlugin/Allocate"
        waves, t0 = 0, time.perf_counter()
        try:
            while waves < min_waves or time.perf_counter() - t0 < budget_s:
                self.ref.restore_pods()
                for w in wires:
                    rc = lib.gs_grpc_client_unary(cl.h, path, w, len(w), resp, len(resp), C.byref(rn),
                                                  C.byref(rstatus), rmsg, 128)
                    if rc != 0 or rstatus.value != 0 or rn.value < 100:
                        raise RuntimeError("reference over the socket: rc %d status %d" % (rc, rstatus.value))
                waves += 1
            dt = time.perf_counter() - t0
        finally:
            cl.close()
            stop()
        del keep
        return waves, dt

    def over_grpcio(self, wires, budget_s, min_waves=1):
        """The same behind Python grpcio on both ends (the only other gRPC stack on this image)."""
        from concurrent import futures

        import grpc
        ref = self

        def allocate(request, _ctx):
            kind = ref.lib.ref_allocate(ref.h, request, len(request), 1, ref.buf, len(ref.buf),
                                        ref.C.byref(ref.n))
            if kind != 0:
                raise RuntimeError("kind %d" % kind)
            return ref.buf.raw[:ref.n.value]
        ident = lambda b: b                                                          # noqa: E731
        d = tempfile.mkdtemp(prefix="gsg")
        sock = os.path.join(d, "aliyungpushare.sock")
        opts = [("grpc.max_send_message_length", -1), ("grpc.max_receive_message_length", -1)]
        server = grpc.server(futures.ThreadPoolExecutor(max_workers=1), options=opts)
        server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler("v1beta1.DevicePlugin", {
            "Allocate": grpc.unary_unary_rpc_method_handler(allocate, ident, ident)}),))
        server.add_insecure_port("unix://" + sock)
        server.start()
        ch = grpc.insecure_channel("unix://" + sock, options=opts)
        call = ch.unary_unary("/v1beta1.DevicePlugin/Allocate", ident, ident)
        waves, t0 = 0, time.perf_counter()
        try:
            while waves < min_waves or time.perf_counter() - t0 < budget_s:
                self.ref.restore_pods()
                for w in wires:
                    if len(call(w, timeout=30)) < 100:
                        raise RuntimeError("short response")
                waves += 1
            dt = time.perf_counter() - t0
        finally:
            ch.close()
            server.stop(0)
            try:
                os.unlink(sock)
            except OSError:
                pass
            os.rmdir(d)
        return waves, dt


def synthetic_uuid(i):
    return "GPU-%08x-0000-0000-0000-%012x" % (i, i)


# ---------------------------------------------------------------- reference arm -------
def run_reference(args, rank, world):
    """The reference algorithm (oracle/ref_model.cpp) on the host cores: the same per-shard
    workload and step as our arm (rank 0 only), then the same N-GPU single-queue stream."""
    if rank != 0:
        return
    uuid = synthetic_uuid(0)
    pods, wires, sizes, _tu = build_workload(0, uuid, UNITS_B200_MIB)
    req_avg = sum(len(w) for w in wires) / len(wires)
    node = RefNode([uuid], UNITS_B200_MIB, pods)

    def step():
        for _ in range(WAVES_PER_STEP):
            node.wave(wires)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    decisions = args.steps * WAVES_PER_STEP * TENANTS
    v = decisions / dt

    budget = 1.0 if SMALL else 8.0
    sw, sdt = node.over_socket(wires, budget)
    gw, gdt = node.over_grpcio(wires, budget)

    t1 = time.perf_counter()
    for _ in range(3):
        node.ref.get_devices()
    law_build = (time.perf_counter() - t1) / 3
    t1 = time.perf_counter()
    for _ in range(3):
        nbytes = node.lib.ref_list_and_watch(node.h, None, 0)
    law_send = (time.perf_counter() - t1) / 3

    # the N-GPU node behind one lock: what the reference actually is at N GPUs
    n = max(1, args.gpus)
    uuids = [synthetic_uuid(g) for g in range(n)]
    npods, stream, _nsizes = build_node_workload(n, uuids, UNITS_B200_MIB)
    nnode = RefNode(uuids, UNITS_B200_MIB, npods)
    nwires = [w for _g, _k, _pi, w in stream]
    qw, qdt = nnode.run_for(nwires, budget)
    single_queue = {"value": qw * len(nwires) / qdt, "unit": "decisions/s", "n_gpus": n,
                    "fake_ids": n * UNITS_B200_MIB, "pods_in_list": len(npods),
                    "sample": "%d waves of %d requests in %.1f s" % (qw, len(nwires), qdt)}

    print(json.dumps({
        "impl": "reference", "metric": "allocate_decisions_per_s", "value": v, "unit": "decisions/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "timed_region_s": dt, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(sizes, WAVES_PER_STEP, req_avg),
        "cpu_baseline": {"value": v, "unit": "decisions/s", "cores": 1, "kind": "port",
                         "sample": "%d decisions (%d waves x 8 tenant requests per step); one thread: "
                                   "the reference holds one lock across Allocate (allocate.go:60)" %
                                   (decisions, WAVES_PER_STEP)},
        "e2e": {"value": v, "unit": "decisions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "over_socket": {"value": sw * TENANTS / sdt, "unit": "decisions/s",
                        "transport": "the reference model behind this repository's transport (gs_server_* with "
                                     "its allocate callback pointing at oracle's ref_server_allocate; "
                                     "gs_grpc_client_* as kubelet)",
                        "sample": "%d waves in %.1f s" % (sw, sdt),
                        "grpcio": {"value": gw * TENANTS / gdt, "unit": "decisions/s",
                                   "transport": "Python grpcio on both ends", "sample": "%d waves in %.1f s" % (gw, gdt)}},
        "single_queue": single_queue,
        "list_and_watch": {"devices": UNITS_B200_MIB, "payload_bytes": nbytes,
                           "get_devices_ms": law_build * 1e3, "marshal_ms": law_send * 1e3},
        "host": host_info(),
        "note": "reference algorithm restated in C++ (Go binary unavailable on this box); C++ is "
                "likely faster than the Go original, so this baseline is conservative",
    }))


# ---------------------------------------------------------------- our arm -------------
def enumerate_gpus(lib, api, L):
    """NVML -> {cuda ordinal: (uuid, units at MiB)}. A box whose NVML cannot be read is not
    silently benchmarked on invented UUIDs: the failure is returned and printed in the line."""
    found, err = {}, None
    try:
        with api.Nvml() as nv:
            infos = nv.devices()
            ords = nv.map_cuda_ordinals(infos)
            for info, o in zip(infos, ords):
                if o >= 0:
                    found[o] = (info.uuid.decode(), lib.gs_units_from_raw(info.total_bytes >> 20, L.GS_UNIT_MIB))
    except L.GpushareError as e:
        err = str(e)
    return found, err


def make_arena(probe, sizes, reserve=6 << 30):
    """One allocation holding the tenants' granted ranges back to back, scaled down together
    if the GPU has less free memory than the tenants add up to."""
    free, _total = probe.mem_info()
    want = sum(sizes) << 30
    arena_bytes = min(want, free - reserve)
    if arena_bytes <= 0:
        raise RuntimeError("no free HBM for the probe arena")
    scale = arena_bytes / want
    arena = probe.alloc(arena_bytes)
    slices, off = [], 0
    for k, gib in enumerate(sizes):
        nb = (int((gib << 30) * scale) >> 20) << 20     # exactly the granted range (MiB units)
        slices.append((arena + off, nb, k))
        off += nb
    return arena, arena_bytes, slices


def run_ours(args, rank, world, local_rank):
    import ctypes as C

    import torch

    from gpushare_b200 import _lib as L
    from gpushare_b200 import api
    from gpushare_b200.native_transport import NativeClient, NativeDevicePluginServer
    from gpushare_b200.probe import Probe, ProbeScheduler

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the gpushare path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    placement = pin_rank(local_rank, world)
    torch.cuda.set_device(local_rank)
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner)
    # are sent to stderr until the line is written
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    dist = cpu_group = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        cpu_group = dist.new_group(backend="gloo")      # waits that must not occupy a GPU

    lib = L.load()
    gpus, nvml_error = enumerate_gpus(lib, api, L)
    if local_rank in gpus:
        uuid, units = gpus[local_rank]
        uuid_source = "nvml"
    else:
        print("bench.py: NVML enumeration failed (%s); UUIDs are synthetic" % nvml_error, file=sys.stderr)
        uuid, units, uuid_source = synthetic_uuid(rank), UNITS_B200_MIB, "synthetic (NVML: %s)" % nvml_error
    ordinal = local_rank

    pods_k8s, wires, sizes, _tu = build_workload(rank, uuid, units)
    pods = [api.Pod.from_k8s(p) for p in pods_k8s]
    table = api.DeviceTable([uuid], units, api.MiB)
    plug = api.NvidiaDevicePlugin(table, NODE)
    pod_list = api.PodList(pods)         # the lister's output, laid out once
    views = pod_list.views

    probe = Probe(ordinal)
    arena, arena_bytes, slices = make_arena(probe, sizes)

    resp = C.create_string_buffer(4096)
    res = L.AllocResult()

    def step_abi(sched):
        """value leg: raw C ABI."""
        for _ in range(WAVES_PER_STEP):
            L.check(lib.gs_plugin_set_pods(plug.h, views, len(pods)))
            for k, w in enumerate(wires):
                rc = lib.gs_plugin_allocate(plug.h, w, len(w), resp, 4096, C.byref(res))
                if rc != 0 or res.kind != L.GS_ALLOC_MATCHED or res.pod_index != k:
                    raise RuntimeError("unexpected decision %d kind %d pod %d" % (rc, res.kind, res.pod_index))
                sched.grant(slices[k][0], slices[k][1], None, k)
            sched.pump()

    def step_api(sched):
        """e2e leg: public Python API, bytes in / bytes out."""
        for _ in range(WAVES_PER_STEP):
            plug.set_pods(pod_list)
            for k, w in enumerate(wires):
                d = plug.Allocate(w)
                if d.kind != "matched" or d.pod_index != k:
                    raise RuntimeError("unexpected decision " + repr(d))
                sched.grant(slices[k][0], slices[k][1], None, k)
            sched.pump()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(step_fn, steps=None, warmup=None):
        steps = args.steps if steps is None else steps
        sched = ProbeScheduler(probe)
        for _ in range(args.warmup if warmup is None else warmup):
            step_fn(sched)
        sched.drain()
        sched = ProbeScheduler(probe)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            step_fn(sched)
        t_dec = time.perf_counter() - t0   # the last decision has been answered
        sched.drain()                      # the in-flight probe batch belongs to the job
        barrier()
        dt = time.perf_counter() - t0
        mine = dt
        if dist is not None:
            t = torch.tensor([dt], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return {"dt": dt, "mine": mine, "decision_clock": t_dec, "sched": sched, "steps": steps}

    sampler = ClockSampler(local_rank, placement.get("spare_cpus")) if rank == 0 else None
    r_abi = timed(step_abi)
    r_api = timed(step_api)

    # third leg (SURVEY 8(d) metric 1(ii)): the same job with kubelet's bytes arriving over the
    # unix socket -- the library's HTTP/2 + gRPC transport on both ends, the client standing in
    # for kubelet. The serving thread makes the decision; this thread only queues the sweep.
    sockdir = tempfile.mkdtemp(prefix="gsb")
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "aliyungpushare.sock"))
    srv.Start()
    cl = NativeClient(srv.socket)
    rpc_path = b"/v1beta1.DevicePlugin/Allocate"
    rn, rstatus, rmsg = C.c_size_t(), C.c_int32(), C.create_string_buffer(128)

    def step_socket(sched):
        for _ in range(WAVES_PER_STEP):
            L.check(lib.gs_plugin_set_pods(plug.h, views, len(pods)))
            for k, w in enumerate(wires):
                rc = lib.gs_grpc_client_unary(cl.h, rpc_path, w, len(w), resp, 4096, C.byref(rn),
                                              C.byref(rstatus), rmsg, 128)
                if rc != 0 or rstatus.value != 0 or rn.value < 100:
                    raise RuntimeError("Allocate over the socket: rc %d status %d %r" % (
                        rc, rstatus.value, rmsg.value))
                sched.grant(slices[k][0], slices[k][1], None, k)
            sched.pump()

    # the socket leg moves ~4x fewer decisions per second: a quarter of the steps keeps the run short
    sock_steps = max(2, args.steps // 4)
    r_sock = timed(step_socket, steps=sock_steps, warmup=max(1, args.warmup // 3))
    sock_stats = srv.stats()
    cl.close()
    srv.Stop()
    try:
        os.rmdir(sockdir)
    except OSError:
        pass
    clocks = sampler.stop() if sampler else None

    # probe kernels alone (device-resident ranges, CUDA events on the probe stream), both patterns
    def alone(pattern):
        runs = []
        for _ in range(3):
            _v, tm = probe.run([(b, n, probe.next_seed(), k, pattern) for b, n, k in slices])
            runs.append(tm)
        best = min(runs, key=lambda t: t.write_ms + t.verify_ms)
        return {"write_gbps": best.bytes / best.write_ms / 1e6, "verify_gbps": best.bytes / best.verify_ms / 1e6,
                "combined_gbps": best.gbps()}
    alone_low = alone(L.GS_PROBE_PATTERN_LOWTOGGLE)
    alone_hash = alone(L.GS_PROBE_PATTERN_HASH)

    size_curve = probe_size_curve(probe, slices[0][0], arena_bytes, L) if rank == 0 and world == 1 else None

    # ListAndWatch payload for this shard
    t1 = time.perf_counter()
    for _ in range(5):
        payload = table.payload_size()
        table.payload_segments()
    law_borrow = (time.perf_counter() - t1) / 5
    buf = C.create_string_buffer(payload)
    outn = C.c_size_t()
    t1 = time.perf_counter()
    for _ in range(5):
        L.check(lib.gs_law_encode(table.h, buf, payload, C.byref(outn)))
    law_copy = (time.perf_counter() - t1) / 5
    t1 = time.perf_counter()
    t2 = api.DeviceTable([uuid], units, api.MiB)
    law_build = time.perf_counter() - t1
    t2.close()

    # all ranks copy at once: what one rank's memcpy gets while the others run theirs
    barrier()
    memcpy_concurrent = host_memcpy_gbps(reps=4)
    barrier()

    decisions = args.steps * WAVES_PER_STEP * TENANTS * world
    totals = {}
    for name, r in (("abi", r_abi), ("api", r_api)):
        s = r["sched"]
        vals = [s.launches, s.bytes_probed, s.bytes_granted, s.dropped, s.write_ms, s.verify_ms, s.bad,
                s.batches]
        if dist is not None:
            t = torch.tensor(vals, dtype=torch.float64, device="cuda")
            dist.all_reduce(t)
            vals = t.tolist()
        totals[name] = vals
    per_rank = [r_abi["mine"] / args.steps * 1e3, r_abi["decision_clock"] / args.steps * 1e3,
                r_api["mine"] / args.steps * 1e3, r_sock["mine"] / sock_steps * 1e3, memcpy_concurrent]
    if dist is not None:
        t = torch.tensor(per_rank, dtype=torch.float64, device="cuda")
        gathered = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
        per_rank_all = [g.tolist() for g in gathered]
    else:
        per_rank_all = [per_rank]

    # the shard legs are over: release this GPU for the single-queue leg, which rank 0 drives
    # over all N GPUs while the other ranks wait on the CPU
    probe.free(arena)
    probe.close()
    torch.cuda.empty_cache()
    barrier()
    single_queue = None
    if rank == 0:
        single_queue = run_single_queue(args, world, gpus, lib, api, L, Probe, ProbeScheduler, C)
    if cpu_group is not None:
        dist.barrier(group=cpu_group)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak, which = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "probe_traffic.json")))
            traffic_ratio = traffic["traffic_over_algorithmic"]
        except (OSError, KeyError, ValueError):
            traffic, traffic_ratio = {}, None
        la = totals["abi"]
        live_ms = la[4] + la[5]
        live_gbps = 2 * la[1] / (live_ms * 1e-3) / 1e9 if live_ms else 0.0
        req_avg = sum(len(w) for w in wires) / len(wires)
        cfg = workload_config(sizes, WAVES_PER_STEP, req_avg)      # identical in both arms
        sock_decisions = sock_steps * WAVES_PER_STEP * TENANTS * world
        line = {
            "metric": "allocate_decisions_per_s", "value": decisions / r_abi["dt"], "unit": "decisions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": r_abi["dt"] / args.steps * 1e3, "timed_region_s": r_abi["dt"],
            "decision_clock_s": r_abi["decision_clock"], "drain_ms": (r_abi["mine"] - r_abi["decision_clock"]) * 1e3,
            "decisions_per_s_on_the_decision_clock_rank0": decisions / world / r_abi["decision_clock"],
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
            "e2e": {"value": decisions / r_api["dt"], "unit": "decisions/s", "timed_region_s": r_api["dt"],
                    "h2d_bytes_per_step": int(totals["api"][0] / world / args.steps * 2584),
                    "d2h_bytes_per_step": int(totals["api"][7] / world / args.steps * 32 * TENANTS),
                    "note": "request and response bytes live on the host (the kubelet side); the GPU "
                            "leg's host->device traffic is the slice descriptors passed as kernel "
                            "parameters, device->host the 32-byte verdict per slice"},
            "over_socket": {"value": sock_decisions / r_sock["dt"], "unit": "decisions/s", "steps": sock_steps,
                            "timed_region_s": r_sock["dt"],
                            "transport": "gs_server_* <- unix socket <- gs_grpc_client_* (HTTP/2 + gRPC "
                                         "in the library on both ends; the client stands in for kubelet)",
                            "allocate_calls_served_rank0": int(sock_stats["allocate_calls"]),
                            "request_bytes_per_step": int(req_avg * TENANTS * WAVES_PER_STEP),
                            "socket_gbps_per_rank": sock_decisions / world / r_sock["dt"] * req_avg / 1e9,
                            "where_the_time_goes_rank0": {
                                "per_request_us": r_sock["mine"] / (sock_steps * WAVES_PER_STEP * TENANTS) * 1e6,
                                "server_recv_of_body_us": sock_stats["recv_body_ns"] / max(sock_stats["allocate_calls"], 1) / 1e3,
                                "server_scan_decide_encode_us": sock_stats["allocate_ns"] / max(sock_stats["allocate_calls"], 1) / 1e3,
                                "note": "the body is copied twice by the kernel (client send, server recv); "
                                        "recv() of the body overlaps the client's send, the rest is the "
                                        "decision and the round trip"}},
            "single_queue": single_queue,
            "gpu_launches": int(la[0]),
            "roofline": {"bound": "hbm", "achieved": live_gbps, "peak": peak, "unit": "GB/s",
                         "frac": live_gbps / peak, "peak_source": which + " (MEASURED_PEAKS.json hbm_gbs: "
                         "torch copy_, read+write)", "frac_of_nominal_8000": live_gbps / 8000.0,
                         "traffic": (traffic_ratio * 2 * la[1] / max(la[7], 1)) if traffic_ratio else None,
                         "traffic_source": "NOT measured in this run: DRAM bytes / algorithmic bytes = %s from "
                                           "the stored ncu --set full capture (%s), scaled to this launch pair" % (
                                               traffic_ratio, traffic.get("source")),
                         "algorithmic_bytes_per_launch_pair": 2 * la[1] / max(la[7], 1),
                         "kernels": {"write_gbps": la[1] / (la[4] * 1e-3) / 1e9 if la[4] else 0,
                                     "verify_gbps": la[1] / (la[5] * 1e-3) / 1e9 if la[5] else 0},
                         "alone": alone_low, "alone_hash_pattern": alone_hash,
                         "load_pass_ceiling": "profiles/r02_read_ceiling.txt: the TMA engine pulling the same "
                                              "slices into shared memory with no compare at all reads 7.37-7.38 "
                                              "TB/s on this pool; the load+compare pass runs at that rate",
                         "how": "sum of slice bytes x 2 / (CUDA-event write ms + verify ms) over every "
                                "probe batch launched inside the timed region, on the probe's stream"},
            "uuid_source": uuid_source,
            "probe": {"batches": int(la[7]), "launches_per_rank": la[0] / world, "arena_bytes": arena_bytes,
                      "bytes_probed": int(la[1]), "bytes_granted": int(la[2]),
                      "coverage": la[1] / la[2] if la[2] else 0, "grants_dropped": int(la[3]),
                      "bad_slices": int(la[6]), "seeds": "fresh per sweep (gs_probe_next_seed)",
                      "write_rate_target_gbps": 7600,
                      "decisions_per_s_if_every_grant_waited_for_its_sweep":
                          world * TENANTS / ((la[4] + la[5]) / max(la[7], 1) * 1e-3) if live_ms else None,
                      "size_curve": size_curve,
                      "note": "grants outrun HBM by design: a wave grants ~%d GiB, one sweep of it takes "
                              "~%.0f ms, so repeated grants of a range coalesce and coverage is the "
                              "fraction actually swept; the last line is what the rate would be if "
                              "Allocate blocked on the sweep (it must not: parity)" % (
                                  sum(sizes), (la[4] + la[5]) / max(la[7], 1))},
            "list_and_watch": {"devices": int(table.device_count), "payload_bytes": payload,
                               "borrow_segments_ms": law_borrow * 1e3, "encode_copy_ms": law_copy * 1e3,
                               "table_build_ms": law_build * 1e3,
                               "encode_copy_gbps": payload / law_copy / 1e9},
            "clocks": clocks, "host": host_info(), "placement": placement,
            "per_rank": {"columns": ["abi_ms_per_step", "abi_decision_clock_ms_per_step", "api_ms_per_step",
                                     "socket_ms_per_step", "memcpy_gbps_all_ranks_at_once"],
                         "rows": per_rank_all},
        }
        memcpy = host_memcpy_gbps()
        scan_gbps = (decisions / world) / r_abi["dt"] * req_avg / 1e9
        conc = [r[4] for r in per_rank_all]
        line["host_roofline"] = {
            "memcpy_gbps_1_thread_alone": memcpy, "memcpy_gbps_per_rank_all_at_once": statistics.mean(conc),
            "memcpy_gbps_sum_all_ranks": sum(conc),
            "request_bytes_gbps_per_rank": scan_gbps, "request_bytes_gbps_all_ranks": scan_gbps * world,
            "frac_of_memcpy": scan_gbps / memcpy,
            "note": "Allocate is host work: every request byte is touched once by the id scan (one "
                    "length-prefix hop per ~51-byte id). No GPU roofline applies to it; the yardsticks are "
                    "this box's single-thread memcpy rate alone and with all ranks copying at once (the "
                    "memory system's share per rank)."}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], ref_sock = cpu_baseline(pods_k8s, wires, uuid, units)
            line["over_socket"]["reference"] = ref_sock
            line["over_socket"]["vs_reference"] = line["over_socket"]["value"] / ref_sock["value"]
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if dist is not None:
        dist.destroy_process_group()


def run_single_queue(args, n_gpus, gpus, lib, api, L, Probe, ProbeScheduler, C):
    """ONE process, ONE plugin, ONE lock over the N-GPU node (allocate.go:60-61, server.go:43-57):
    an N x 183 359-id table, a lister list of 64 N pods, one request stream with the tenants of
    all N GPUs interleaved, N probe engines fed from it. A step holds as many decisions as a
    per-shard step, whatever N is."""
    uuids, units = [], UNITS_B200_MIB
    for g in range(n_gpus):
        if g in gpus:
            uuids.append(gpus[g][0])
            units = gpus[0][1] if 0 in gpus else units
        else:
            uuids.append(synthetic_uuid(g))
    pods_k8s, stream, sizes = build_node_workload(n_gpus, uuids, units)
    pods = [api.Pod.from_k8s(p) for p in pods_k8s]
    table = api.DeviceTable(uuids, units, api.MiB)
    plug = api.NvidiaDevicePlugin(table, NODE)
    pod_list = api.PodList(pods)
    views = pod_list.views
    probes, scheds, arenas, gslices = [], [], [], []
    for g in range(n_gpus):
        p = Probe(g)
        arena, _nb, sl = make_arena(p, sizes[g])
        probes.append(p)
        arenas.append(arena)
        gslices.append(sl)
    resp = C.create_string_buffer(4096)
    res = L.AllocResult()
    waves = max(1, WAVES_PER_STEP // n_gpus)

    def step(scheds):
        for _ in range(waves):
            L.check(lib.gs_plugin_set_pods(plug.h, views, len(pods)))
            for g, k, pod_index, w in stream:
                rc = lib.gs_plugin_allocate(plug.h, w, len(w), resp, 4096, C.byref(res))
                if rc != 0 or res.kind != L.GS_ALLOC_MATCHED or res.pod_index != pod_index or res.gpu_index != g:
                    raise RuntimeError("single queue: unexpected decision rc %d kind %d pod %d gpu %d (want pod %d gpu %d)"
                                       % (rc, res.kind, res.pod_index, res.gpu_index, pod_index, g))
                scheds[g].grant(gslices[g][k][0], gslices[g][k][1], None, k)
            for s in scheds:
                s.pump()

    warm = [ProbeScheduler(p) for p in probes]
    for _ in range(max(1, args.warmup // 2)):
        step(warm)
    for s in warm:
        s.drain()
    scheds = [ProbeScheduler(p) for p in probes]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(scheds)
    t_dec = time.perf_counter() - t0
    for s in scheds:
        s.drain()
    dt = time.perf_counter() - t0
    decisions = args.steps * waves * len(stream)
    per_gpu = [s.gbps() for s in scheds]
    out = {"value": decisions / dt, "unit": "decisions/s", "n_gpus": n_gpus,
           "fake_ids": int(table.device_count), "pods_in_list": len(pods), "steps": args.steps,
           "waves_per_step": waves, "decisions_per_step": waves * len(stream), "timed_region_s": dt,
           "decision_clock_s": t_dec, "request_bytes_avg": sum(len(w) for *_x, w in stream) / len(stream),
           "probe_gbps_per_gpu": per_gpu, "probe_batches": sum(s.batches for s in scheds),
           "gpu_launches": sum(s.launches for s in scheds), "bad_slices": sum(s.bad for s in scheds),
           "what": "one process, one gs_plugin (one lock) over the whole node, one request stream for all "
                   "GPUs' tenants, one probe engine per GPU fed from it: the reference's shape; compare with "
                   "the reference arm's single_queue at the same N"}
    for p, a in zip(probes, arenas):
        p.free(a)
        p.close()
    return out


def probe_size_curve(probe, base, arena_bytes, L):
    """Probe rate against slice size: one slice alone, and 64 slices of that size in one batch
    (what the scheduler makes of many small grants). An L2 scrub is added when the batch is
    smaller than 512 MiB, as the scheduler does, and counts as time. Both passes, CUDA events,
    each form repeated for 0.2 s, median of the second half."""
    MiB = 1 << 20
    rows = []
    for nbytes in (16 * MiB, 64 * MiB, 256 * MiB, 1024 * MiB, 2048 * MiB, 8192 * MiB, 40960 * MiB):
        if nbytes > arena_bytes:
            break
        row = {"slice_bytes": nbytes}
        for label, count in (("alone", 1), ("batch_of_64", 64)):
            if count * nbytes > arena_bytes:
                continue
            passes = L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY
            if count * nbytes < 512 * MiB:
                passes |= L.GS_PROBE_SCRUB
            rates, t0 = [], time.perf_counter()
            while len(rates) < 6 or time.perf_counter() - t0 < 0.2:    # long enough for the clocks to settle
                sl = [(base + k * nbytes, nbytes, probe.next_seed(), k) for k in range(count)]
                v, tm = probe.run(sl, passes)
                if not all(x.clean for x in v):
                    raise RuntimeError("size curve: dirty slice")
                rates.append(tm.gbps())
            row[label + "_gbps"] = statistics.median(rates[len(rates) // 2:])
        rows.append(row)
    return rows


def cpu_baseline(pods_k8s, wires, uuid, units, budget_s=10.0):
    """oracle/ref_model.cpp on this box's host cores: a bounded sample of the same waves, called
    in-process and over the socket."""
    node = RefNode([uuid], units, pods_k8s)
    waves, dt = node.run_for(wires, budget_s)
    sw, sdt = node.over_socket(wires, budget_s / 2)
    return ({"value": waves * TENANTS / dt, "unit": "decisions/s", "cores": 1, "kind": "port",
             "sample": "%d waves of the same 8 tenant requests in %.1f s; one thread (the reference "
                       "serialises Allocate under m.Lock())" % (waves, dt)},
            {"value": sw * TENANTS / sdt, "unit": "decisions/s",
             "sample": "%d waves in %.1f s, the reference model served by the same transport" % (sw, sdt)})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

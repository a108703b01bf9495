#!/usr/bin/env python
"""bench.py -- the gpushare hot path on N B200s (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Metric (BASELINE.json): Allocate decisions/s at MiB granularity, with the slice probe's
HBM GB/s as the kernel roofline. One rank = one GPU's shard of the node: its 183 359
fake-device ids, the tenant pods the scheduler extender bound to it, its probe engine.
Nothing is shared between shards (no collective on the data path): weak scaling.

A step = WAVES_PER_STEP admission waves. One wave = the lister's pod list is refreshed
(64 pods: 8 pending tenants bound to this GPU, sizes 1-40 GiB as BASELINE config 4 draws
them, plus 56 already-assigned pods) and kubelet's 8 AllocateRequests arrive, each carrying
n x 1024 device ids (n GiB at MiB granularity, 51 KB - 2 MB of wire bytes). Every decision
= decode request + candidate filter/sort/match + response and patch encode + patch applied
(SURVEY 8(d) metric 1). Every grant is queued for the slice probe, which sweeps exactly the
granted range on the GPU beside the decisions and never gates them (probe.ProbeScheduler).

  value : decisions/s at the bare C ABI (gs_plugin_allocate, caller-owned buffers), probe
          engine running on HBM-resident ranges.
  e2e   : the same through the package's public API (NvidiaDevicePlugin.Allocate: bytes
          in, bytes out) with the probe scheduler, result read-back included.
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
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NODE = "node-a"
UNITS_B200_MIB = 183359          # nvidia-smi / NVML total of a B200 in MiB
TENANTS = 8                      # config 4: 64 tenants over 8 GPUs
PODS_IN_LIST = 64
WAVES_PER_STEP = 32
SEED = 20261106


# ---------------------------------------------------------------- workload ------------
def tenant_sizes_gib(rank):
    """BASELINE config 4 per-GPU share: 8 sizes uniform in 1..40 GiB, redrawn until they fit
    one B200 (sum <= 160 GiB so the probed ranges fit beside the CUDA context)."""
    import random
    rng = random.Random(SEED + rank)
    if os.environ.get("GS_BENCH_SMALL"):               # CPU tests: MiB-sized tenants
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


def build_workload(rank, uuid, units):
    """-> (pods as k8s dicts, request wires in arrival order, sizes GiB)."""
    import numpy as np
    sizes = tenant_sizes_gib(rank)
    per = 16 if os.environ.get("GS_BENCH_SMALL") else 1024   # units per size step
    rng = np.random.default_rng(SEED + 1000 * rank)
    free = rng.permutation(units)                      # ids drawn without replacement
    wires, off = [], 0
    for gib in sizes:
        n = gib * per
        ids = ["%s-_-%d" % (uuid, j) for j in free[off:off + n]]
        off += n
        wires.append(encode_allocate_request([ids]))
    pods = []
    for k, gib in enumerate(sizes):
        pods.append({"metadata": {"uid": "tenant-%d-%d" % (rank, k), "name": "tenant-%d" % k,
                                  "namespace": "default",
                                  "annotations": {"ALIYUN_COM_GPU_MEM_IDX": "0",
                                                  "ALIYUN_COM_GPU_MEM_ASSUME_TIME": str(10**18 + 977 * k),
                                                  "ALIYUN_COM_GPU_MEM_ASSIGNED": "false"}},
                     "spec": {"nodeName": NODE, "containers": [
                         {"name": "main", "resources": {"limits": {"aliyun.com/gpu-mem": str(gib * per)}}}]},
                     "status": {"phase": "Pending"}})
    for k in range(PODS_IN_LIST - TENANTS):
        pods.append({"metadata": {"uid": "bg-%d-%d" % (rank, k), "name": "running-%d" % k,
                                  "namespace": "default",
                                  "annotations": {"ALIYUN_COM_GPU_MEM_IDX": "0",
                                                  "ALIYUN_COM_GPU_MEM_ASSUME_TIME": str(10**17 + k),
                                                  "ALIYUN_COM_GPU_MEM_ASSIGNED": "true"}},
                     "spec": {"nodeName": NODE, "containers": [
                         {"name": "main", "resources": {"limits": {"aliyun.com/gpu-mem": str(1024 * (1 + k % 7))}}}]},
                     "status": {"phase": "Pending"}})
    return pods, wires, sizes


# ---------------------------------------------------------------- clocks --------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        out = self.p.communicate()[0]
        sm, mx, reasons = [], 0, set()
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [x for x in sm if x > 0.5 * mx] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------- reference arm -------
def run_reference(args, rank, world):
    """The reference algorithm (oracle/ref_model.cpp) on the host cores, same workload."""
    from oracle import ref_cpp
    if rank != 0:
        return
    uuid = "GPU-%08x-0000-0000-0000-%012x" % (0, 0)
    pods, wires, sizes = build_workload(0, uuid, UNITS_B200_MIB)
    ref = ref_cpp.RefPlugin([uuid], UNITS_B200_MIB, "MiB", NODE)
    # bounded sample: a reference step is REF_WAVES waves of the same 8 requests
    ref_waves = max(1, WAVES_PER_STEP // 8)
    import ctypes as C
    lib, h = ref.l, ref.h
    buf = C.create_string_buffer(1 << 16)
    n = C.c_size_t()

    ref.set_pods_k8s(pods)
    ref.save_pods()

    def step():
        for _ in range(ref_waves):
            ref.restore_pods()          # the lister's answer, re-installed in C++ (no ctypes cost)
            for w in wires:
                kind = lib.ref_allocate(h, w, len(w), 1, buf, len(buf), C.byref(n))
                assert kind == 0, kind
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    decisions = args.steps * ref_waves * TENANTS
    t1 = time.perf_counter()
    for _ in range(3):
        ref.get_devices()
    law_build = (time.perf_counter() - t1) / 3
    t1 = time.perf_counter()
    for _ in range(3):
        nbytes = lib.ref_list_and_watch(h, None, 0)
    law_send = (time.perf_counter() - t1) / 3
    v = decisions / dt
    print(json.dumps({
        "impl": "reference", "metric": "allocate_decisions_per_s", "value": v, "unit": "decisions/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(sizes, ref_waves),
        "cpu_baseline": {"value": v, "unit": "decisions/s", "cores": 1, "kind": "port",
                         "sample": "%d decisions (%d waves x 8 tenant requests per step); one thread: "
                                   "the reference holds one lock across Allocate (allocate.go:60)" %
                                   (decisions, ref_waves)},
        "e2e": {"value": v, "unit": "decisions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "list_and_watch": {"devices": UNITS_B200_MIB, "payload_bytes": nbytes,
                           "get_devices_ms": law_build * 1e3, "marshal_ms": law_send * 1e3},
        "host": host_info(),
        "note": "reference algorithm restated in C++ (Go binary unavailable on this box); C++ is "
                "likely faster than the Go original, so this baseline is conservative",
    }))


def workload_config(sizes, waves):
    return {"workload": "BASELINE configs[2]/[3] per-GPU shard: 1xB200 at MiB granularity "
                        "(183359 fake ids), 8 tenant pods %s GiB + 56 assigned pods in the lister's "
                        "list, requests of n*1024 ids; probe of each granted range" % sizes,
            "waves_per_step": waves, "decisions_per_step": waves * TENANTS,
            "request_bytes_avg": None, "l2": "probed ranges (>= 100 GiB per batch) exceed the 126 MB L2",
            "parallelism": "one process per GPU, no collective; each rank decides for its own GPU's shard of the "
                           "node (one kubelet queue feeding one plugin would run at the 1-GPU rate)"}


def host_memcpy_gbps():
    """Single-thread memcpy of a buffer larger than the L3 (bytes copied / s, reads + writes not
    double counted): the host-side yardstick for the request scan and the payload copy."""
    import ctypes
    n = 256 << 20
    a, b = ctypes.create_string_buffer(n), ctypes.create_string_buffer(n)
    ctypes.memmove(b, a, n)
    best = 0.0
    for _ in range(3):
        t0 = time.perf_counter()
        ctypes.memmove(b, a, n)
        best = max(best, n / (time.perf_counter() - t0) / 1e9)
    return best


def host_info():
    model = ""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return {"nproc": os.cpu_count(), "cpu": model}


# ---------------------------------------------------------------- our arm -------------
def run_ours(args, rank, world, local_rank):
    import ctypes as C

    import torch

    from gpushare_b200 import _lib as L
    from gpushare_b200 import api
    from gpushare_b200.probe import Probe, ProbeScheduler

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the gpushare path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner)
    # are sent to stderr until the line is written
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    lib = L.load()
    # NVML enumerate (real driver): this rank's GPU
    uuid, units, ordinal = None, UNITS_B200_MIB, local_rank
    try:
        with api.Nvml() as nv:
            infos = nv.devices()
            ords = nv.map_cuda_ordinals(infos)
            for info, o in zip(infos, ords):
                if o == local_rank:
                    uuid = info.uuid.decode()
                    units = lib.gs_units_from_raw(info.total_bytes >> 20, L.GS_UNIT_MIB)
    except L.GpushareError:
        pass
    if uuid is None:
        uuid = "GPU-%08x-0000-0000-0000-%012x" % (rank, rank)

    pods_k8s, wires, sizes = build_workload(rank, uuid, units)
    pods = [api.Pod.from_k8s(p) for p in pods_k8s]
    table = api.DeviceTable([uuid], units, api.MiB)
    plug = api.NvidiaDevicePlugin(table, NODE)
    pod_list = api.PodList(pods)         # the lister's output, laid out once
    views = pod_list.views

    probe = Probe(ordinal)
    free, _total = probe.mem_info()
    want = sum(sizes) << 30
    arena_bytes = min(want, free - (6 << 30))
    scale = arena_bytes / want
    arena = probe.alloc(arena_bytes)
    slices, off = [], 0
    for k, gib in enumerate(sizes):
        nb = (int((gib << 30) * scale) >> 20) << 20     # exactly the granted range (MiB units)
        slices.append((arena + off, nb, SEED + k, k))
        off += nb

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
                sched.grant(*slices[k])
            sched.pump()

    def step_api(sched):
        """e2e leg: public Python API, bytes in / bytes out."""
        for _ in range(WAVES_PER_STEP):
            plug.set_pods(pod_list)
            for k, w in enumerate(wires):
                d = plug.Allocate(w)
                if d.kind != "matched" or d.pod_index != k:
                    raise RuntimeError("unexpected decision " + repr(d))
                sched.grant(*slices[k])
            sched.pump()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(step_fn):
        sched = ProbeScheduler(probe)
        for _ in range(args.warmup):
            step_fn(sched)
        sched.drain()
        sched = ProbeScheduler(probe)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_fn(sched)
        sched.drain()                      # the in-flight probe batch belongs to the job
        barrier()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt, sched

    sampler = ClockSampler(local_rank) if rank == 0 else None
    dt_abi, s_abi = timed(step_abi)
    dt_api, s_api = timed(step_api)

    # third leg (SURVEY 8(d) metric 1(ii)): the same job with kubelet's bytes arriving over the
    # unix socket -- the library's HTTP/2 + gRPC transport on both ends, the client standing in
    # for kubelet. The serving thread makes the decision; this thread only queues the sweep.
    import tempfile

    from gpushare_b200.native_transport import NativeClient, NativeDevicePluginServer
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
                sched.grant(*slices[k])
            sched.pump()

    dt_sock, _s_sock = timed(step_socket)
    sock_stats = srv.stats()
    cl.close()
    srv.Stop()
    try:
        os.rmdir(sockdir)
    except OSError:
        pass
    clocks = sampler.stop() if sampler else None

    # probe kernels alone (device-resident ranges, CUDA events on the probe stream)
    probe_runs = []
    for _ in range(3):
        _v, tm = probe.run(slices)
        probe_runs.append(tm)
    best = min(probe_runs, key=lambda t: t.write_ms + t.verify_ms)

    # ListAndWatch payload for this shard
    t1 = time.perf_counter()
    for _ in range(5):
        payload = table.payload_size()
        segs = table.payload_segments()
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

    decisions = args.steps * WAVES_PER_STEP * TENANTS * world
    totals = {}
    for name, s in (("abi", s_abi), ("api", s_api)):
        vals = [s.launches, s.bytes_probed, s.bytes_granted, s.dropped, s.write_ms, s.verify_ms, s.bad,
                s.batches]
        if dist is not None:
            t = torch.tensor(vals, dtype=torch.float64, device="cuda")
            dist.all_reduce(t)
            vals = t.tolist()
        totals[name] = vals

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak, which = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
        try:
            traffic_ratio = json.load(open(os.path.join(ROOT, "profiles", "probe_traffic.json")))[
                "traffic_over_algorithmic"]
        except (OSError, KeyError, ValueError):
            traffic_ratio = None
        la = totals["abi"]
        live_ms = la[4] + la[5]
        live_gbps = 2 * la[1] / (live_ms * 1e-3) / 1e9 if live_ms else 0.0
        launches_per_rank = la[0] / world
        req_avg = sum(len(w) for w in wires) / len(wires)
        cfg = workload_config(sizes, WAVES_PER_STEP)
        cfg["request_bytes_avg"] = req_avg
        cfg["probe_arena_bytes"] = arena_bytes
        line = {
            "metric": "allocate_decisions_per_s", "value": decisions / dt_abi, "unit": "decisions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt_abi / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
            "e2e": {"value": decisions / dt_api, "unit": "decisions/s",
                    "h2d_bytes_per_step": int(totals["api"][0] / world / args.steps * 2584),
                    "d2h_bytes_per_step": int(totals["api"][7] / world / args.steps * 32 * TENANTS),
                    "note": "request and response bytes live on the host (the kubelet side); the GPU "
                            "leg's host->device traffic is the slice descriptors passed as kernel "
                            "parameters, device->host the 32-byte verdict per slice"},
            "over_socket": {"value": decisions / dt_sock, "unit": "decisions/s",
                            "transport": "gs_server_* <- unix socket <- gs_grpc_client_* (HTTP/2 + gRPC "
                                         "in the library on both ends; the client stands in for kubelet)",
                            "allocate_calls_served_rank0": int(sock_stats["allocate_calls"]),
                            "request_bytes_per_step": int(req_avg * TENANTS * WAVES_PER_STEP)},
            "gpu_launches": int(la[0]),
            "roofline": {"bound": "hbm", "achieved": live_gbps, "peak": peak, "unit": "GB/s",
                         "frac": live_gbps / peak, "peak_source": which + " (MEASURED_PEAKS.json hbm_gbs: "
                         "torch copy_, read+write)", "frac_of_nominal_8000": live_gbps / 8000.0,
                         "traffic": (traffic_ratio * 2 * la[1] / max(la[7], 1)) if traffic_ratio else None,
                         "traffic_source": "DRAM bytes / algorithmic bytes = %s from the ncu --set full "
                                           "capture in profiles/probe_traffic.json, scaled to this "
                                           "launch pair" % traffic_ratio,
                         "algorithmic_bytes_per_launch_pair": 2 * la[1] / max(la[7], 1),
                         "kernels": {"write_gbps": la[1] / (la[4] * 1e-3) / 1e9 if la[4] else 0,
                                     "verify_gbps": la[1] / (la[5] * 1e-3) / 1e9 if la[5] else 0},
                         "alone": {"write_gbps": best.bytes / best.write_ms / 1e6,
                                   "verify_gbps": best.bytes / best.verify_ms / 1e6,
                                   "combined_gbps": best.gbps()},
                         "how": "sum of slice bytes x 2 / (CUDA-event write ms + verify ms) over every "
                                "probe batch launched inside the timed region, on the probe's stream"},
            "probe": {"batches": int(la[7]), "launches_per_rank": launches_per_rank,
                      "bytes_probed": int(la[1]), "bytes_granted": int(la[2]),
                      "coverage": la[1] / la[2] if la[2] else 0, "grants_dropped": int(la[3]),
                      "bad_slices": int(la[6]),
                      "write_rate_target_gbps_rank0": probe.write_rate()[0],
                      "decisions_per_s_if_every_grant_waited_for_its_sweep":
                          world * TENANTS / ((la[4] + la[5]) / max(la[7], 1) * 1e-3) if live_ms else None,
                      "note": "grants outrun HBM by design: a wave grants ~%d GiB, one sweep of it takes "
                              "~%.0f ms, so repeated grants of a range coalesce and coverage is the "
                              "fraction actually swept; the last line is what the rate would be if "
                              "Allocate blocked on the sweep (it must not: parity)" % (
                                  sum(sizes), (la[4] + la[5]) / max(la[7], 1))},
            "list_and_watch": {"devices": int(table.device_count), "payload_bytes": payload,
                               "borrow_segments_ms": law_borrow * 1e3, "encode_copy_ms": law_copy * 1e3,
                               "table_build_ms": law_build * 1e3,
                               "encode_copy_gbps": payload / law_copy / 1e9},
            "clocks": clocks, "host": host_info(),
        }
        memcpy = host_memcpy_gbps()
        scan_gbps = (decisions / world) / dt_abi * req_avg / 1e9
        line["host_roofline"] = {
            "memcpy_gbps_1_thread": memcpy, "request_bytes_gbps_per_rank": scan_gbps,
            "frac_of_memcpy": scan_gbps / memcpy,
            "note": "Allocate is host work: every request byte is touched once by the id scan (one "
                    "length-prefix hop per ~51-byte id). No GPU roofline applies to it; the yardstick is "
                    "this box's single-thread memcpy rate."}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(pods_k8s, wires, uuid, units)
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    probe.free(arena)
    probe.close()
    if dist is not None:
        dist.destroy_process_group()


def cpu_baseline(pods_k8s, wires, uuid, units, budget_s=12.0):
    """oracle/ref_model.cpp on this box's host cores: a bounded sample of the same waves."""
    import ctypes as C

    from oracle import ref_cpp
    ref = ref_cpp.RefPlugin([uuid], units, "MiB", NODE)
    lib, h = ref.l, ref.h
    buf = C.create_string_buffer(1 << 16)
    n = C.c_size_t()
    ref.set_pods_k8s(pods_k8s)
    ref.save_pods()
    waves, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget_s:
        ref.restore_pods()
        for w in wires:
            kind = lib.ref_allocate(h, w, len(w), 1, buf, len(buf), C.byref(n))
            assert kind == 0
        waves += 1
    dt = time.perf_counter() - t0
    return {"value": waves * TENANTS / dt, "unit": "decisions/s", "cores": 1, "kind": "port",
            "sample": "%d waves of the same 8 tenant requests in %.1f s; one thread (the reference "
                      "serialises Allocate under m.Lock())" % (waves, dt)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
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

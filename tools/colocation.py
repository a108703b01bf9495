"""BASELINE config 4: 64 tenant pods with mixed 1-40 GiB slices on 8 B200s, all slices probed
concurrently -- one process, one probe engine (stream) per GPU, no collective.

    python tools/colocation.py [--gpus N] [--scale 1.0] > profiles/r01_colocation_8gpu.json

Tenants: sizes uniform in 1..40 GiB (seed 1), first-fit bin-packed onto the GPUs with at most 8
per GPU and at most 160 GiB per GPU (standing in for the scheduler extender), as SURVEY 8(d)
describes. Every tenant's Allocate goes through the C ABI at MiB granularity; every granted range
is then swept on its GPU, all GPUs at once. One tenant's range is corrupted between the passes to
show that exactly that tenant -- and nobody else -- is reported.
"""
import argparse
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200 import api  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402


def encode_request(ids):
    out = bytearray()
    inner = bytearray()
    for i in ids:
        b = i.encode()
        inner += b"\x0a" + bytes([len(b)]) + b
    n, hdr = len(inner), bytearray(b"\x0a")
    while n >= 0x80:
        hdr.append((n & 0x7F) | 0x80)
        n >>= 7
    hdr.append(n)
    out += hdr + inner
    return bytes(out)


def place(n_gpus, n_tenants, seed=1, per_gpu_cap=160, per_gpu_max=8):
    """Sizes are drawn uniformly in 1..40 GiB; a draw that does not fit any GPU (first fit,
    <= per_gpu_max tenants and <= per_gpu_cap GiB each) is rejected and redrawn, as SURVEY 8(d)
    says. Bounded: gives up after 10 000 draws and returns what was placed."""
    rng = random.Random(seed)
    load, count, tenants = [0] * n_gpus, [0] * n_gpus, []
    for _ in range(10000):
        if len(tenants) >= n_tenants:
            break
        size = rng.randint(1, 40)
        for g in range(n_gpus):
            if count[g] < per_gpu_max and load[g] + size <= per_gpu_cap:
                tenants.append((g, size))
                load[g] += size
                count[g] += 1
                break
    return tenants


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--scale", type=float, default=1.0, help="shrink every slice (smoke runs)")
    a = ap.parse_args()
    with api.Nvml() as nv:
        infos = nv.devices()
        ords = nv.map_cuda_ordinals(infos)
    gpus = [(i, o) for i, o in zip(infos, ords) if o >= 0]
    if a.gpus:
        gpus = gpus[:a.gpus]
    n = len(gpus)
    table = api.DeviceTable.from_gpus((L.GpuInfo * n)(*[g[0] for g in gpus]), api.MiB)
    units = table.units_per_gpu
    tenants = place(n, 8 * n)
    pods = [api.Pod("u%d" % k, "default", "tenant-%d" % k, "node-a", gib * 1024, str(g),
                    str(10**18 + 13 * k), "false") for k, (g, gib) in enumerate(tenants)]
    plug = api.NvidiaDevicePlugin(table, "node-a")
    plug.set_pods(pods)
    rng = random.Random(2)
    t0 = time.perf_counter()
    decisions = []
    for k, (g, gib) in enumerate(tenants):
        ids = ["%s-_-%d" % (gpus[rng.randrange(n)][0].uuid.decode(), rng.randrange(units))
               for _ in range(gib * 1024)]
        d = plug.Allocate(encode_request(ids))
        assert d.kind == "matched" and d.pod_index == k and d.gpu_index == g, (k, d.kind, d.gpu_index)
        decisions.append(d)
    alloc_s = time.perf_counter() - t0

    probes = [Probe(o) for _i, o in gpus]
    arenas, slices = [], [[] for _ in range(n)]
    for g in range(n):
        want = sum(int((gib << 30) * a.scale) & ~0xFFFFF for gg, gib in tenants if gg == g)
        arenas.append(probes[g].alloc(max(want, 1 << 20)))
        off = 0
        for k, (gg, gib) in enumerate(tenants):
            if gg == g:
                nb = int((gib << 30) * a.scale) & ~0xFFFFF
                slices[g].append((arenas[g] + off, nb, 1000 + k, k))
                off += nb

    def sweep(passes):
        t = time.perf_counter()
        for g in range(n):
            probes[g].submit(slices[g], passes)          # asynchronous: all GPUs run at once
        out = [probes[g].wait() for g in range(n)]
        return out, time.perf_counter() - t

    sweep(L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY)           # warm
    res, wall = sweep(L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY)
    per_gpu = [{"gpu": g, "tenants": len(slices[g]), "bytes": tm.bytes, "write_gbps": tm.bytes / tm.write_ms / 1e6,
                "verify_gbps": tm.bytes / tm.verify_ms / 1e6, "combined_gbps": tm.gbps()}
               for g, (_v, tm) in enumerate(res)]
    assert all(v.clean for vs, _tm in res for v in vs)
    total = sum(p["bytes"] for p in per_gpu)
    # corrupt one tenant's range; only that tenant may be reported
    victim_gpu = n - 1
    victim = slices[victim_gpu][len(slices[victim_gpu]) // 2]
    probes[victim_gpu].inject(victim[0], (victim[1] // 2) & ~3, 0x00100000)
    res2, _ = sweep(L.GS_PROBE_VERIFY)
    flagged = [s[3] for g in range(n) for s, v in zip(slices[g], res2[g][0]) if not v.clean]
    for g in range(n):
        probes[g].free(arenas[g])
        probes[g].close()
    print(json.dumps({
        "config": "BASELINE configs[3]: %d GPUs, %d tenants, sizes 1-40 GiB, <= 8 per GPU, scale %.3f" % (
            n, len(tenants), a.scale),
        "tenants_gib": [gib for _g, gib in tenants], "placement": [g for g, _ in tenants],
        "allocate": {"decisions": len(tenants), "seconds": alloc_s,
                     "note": "C ABI through the Python API, MiB granularity, ids built in Python (not timed apart)"},
        "probe": {"per_gpu": per_gpu, "bytes_swept": total, "wall_s_both_passes_all_gpus": wall,
                  "aggregate_gbps_wall": 2 * total / wall / 1e9,
                  "min_gpu_combined_gbps": min(p["combined_gbps"] for p in per_gpu)},
        "fault_injection": {"victim_tenant": victim[3], "flagged_tenants": flagged},
    }, indent=1))
    assert flagged == [victim[3]], flagged


if __name__ == "__main__":
    main()

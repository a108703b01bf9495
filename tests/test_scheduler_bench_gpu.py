"""Probe scheduler and the bench contract line on a real GPU."""
import json
import os
import subprocess
import sys

import pytest

from gpushare_b200 import _lib as L
from tests.conftest import ROOT

pytestmark = pytest.mark.gpu
MiB = 1 << 20


def test_scheduler_coalesces_drains_and_reports_bad_slices():
    from gpushare_b200.probe import Probe, ProbeScheduler
    bad = []
    with Probe(0) as p:
        arena = p.alloc(64 * MiB)
        try:
            s = ProbeScheduler(p, on_bad=lambda tag, v: bad.append((tag, v.mismatches)))
            for wave in range(50):                       # the same 4 ranges granted 50 times
                for k in range(4):
                    s.grant(arena + k * 16 * MiB, 16 * MiB, 7 + k, k)
                s.pump()
            s.drain()
            assert s.inflight is None and not s.pending
            assert s.bytes_granted == 50 * 64 * MiB
            assert 0 < s.bytes_probed <= s.bytes_granted and s.bytes_probed % (16 * MiB) == 0
            assert s.coalesced > 0 and s.dropped == 0 and s.bad == 0
            assert s.launches == 3 * s.batches           # 16 MiB slices: store, L2 scrub, load
            assert s.gbps() > 100
            # a verify-only sweep over a range someone scribbled on reports through on_bad
            p.inject(arena + 16 * MiB, 4096, 0x1)
            v, _ = p.run([(arena + 16 * MiB, 16 * MiB, 8, 1)], L.GS_PROBE_VERIFY)
            assert v[0].mismatches == 1
            s2 = ProbeScheduler(p, on_bad=lambda tag, v: bad.append((tag, v.mismatches)))
            s2.grant(arena, 16 * MiB, 7, 0)
            s2.drain()
            assert s2.bad == 0 and bad == []             # a fresh write+verify is clean again
        finally:
            p.free(arena)


def test_bench_line_contract():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "2", "--warmup", "3",
                          "--no-cpu-baseline"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    r = json.loads(out.stdout.strip().splitlines()[-1])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "roofline",
                "clocks"):
        assert key in r, key
    assert r["metric"] == "allocate_decisions_per_s" and r["n_gpus"] == 1 and r["steps"] == 2
    assert r["value"] > 1000 and r["e2e"]["value"] > 500 and r["gpu_launches"] >= 2
    assert r["roofline"]["bound"] == "hbm" and r["roofline"]["achieved"] > 3000
    assert r["probe"]["bad_slices"] == 0 and "workload" in r["config"]
    assert r["over_socket"]["value"] > 500 and r["over_socket"]["allocate_calls_served_rank0"] >= 2 * 256


def test_colocation_tool_small_scale():
    """BASELINE config 4 in miniature on however many GPUs are visible: every tenant allocated,
    all ranges swept concurrently, the one corrupted tenant (and only it) reported."""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "colocation.py"), "--scale", "0.01"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    r = json.loads(out.stdout)
    assert r["fault_injection"]["flagged_tenants"] == [r["fault_injection"]["victim_tenant"]]
    assert len(r["probe"]["per_gpu"]) >= 1 and r["allocate"]["decisions"] == len(r["tenants_gib"])


def _start_daemon(tmp_path, pods, extra_env=None, interval=1, extra_args=()):
    import time
    (tmp_path / "pods.json").write_text(json.dumps({"items": pods}))
    env = dict(os.environ, NODE_NAME="node-a", PYTHONPATH=ROOT, **(extra_env or {}))
    exe = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "bin", "gpushare-device-plugin-v2")
    proc = subprocess.Popen([exe, "--v=1", "--memory-unit=GiB", "--socket-dir", str(tmp_path),
                             "--pod-source-file", str(tmp_path / "pods.json"), "--no-register", "--probe",
                             "--health-check", "--health-probe-interval=%d" % interval] + list(extra_args),
                            env=env, stderr=subprocess.PIPE, text=True)
    sock = str(tmp_path / "aliyungpushare.sock")
    deadline = time.time() + 60
    while not os.path.exists(sock) and time.time() < deadline:
        time.sleep(0.1)
    return proc, sock


def _stop_daemon(proc):
    import signal
    proc.send_signal(signal.SIGTERM)
    try:
        proc.wait(30)
    except subprocess.TimeoutExpired:
        proc.kill()
    return proc.stderr.read()


def _tenant(name, gib, assume):
    return {"metadata": {"uid": "u-" + name, "name": name, "namespace": "default", "annotations": {
        "ALIYUN_COM_GPU_MEM_IDX": "0", "ALIYUN_COM_GPU_MEM_ASSUME_TIME": str(assume),
        "ALIYUN_COM_GPU_MEM_ASSIGNED": "false"}},
        "spec": {"nodeName": "node-a", "containers": [
            {"resources": {"limits": {"aliyun.com/gpu-mem": str(gib)}}}]}, "status": {"phase": "Pending"}}


def _request(uuid, n):
    ids = [(uuid + "-_-%d" % j).encode() for j in range(n)]
    inner = b"".join(b"\x0a" + bytes([len(i)]) + i for i in ids)
    ln = len(inner)
    var = bytearray()
    while ln >= 0x80:
        var.append((ln & 0x7F) | 0x80)
        ln >>= 7
    var.append(ln)
    return b"\x0a" + bytes(var) + inner


def test_daemon_on_real_gpu_probes_at_allocate_and_health_time(tmp_path):
    """The shipped daemon (the native binary) against the real NVML and GPU: enumerate, serve,
    Allocate for pods bound to GPU 0, sweep the granted ranges -- grants that arrive within the
    gather window (stretched to 1.5 s here so the test does not depend on timing) share ONE launch
    pair -- and sweep free memory on the health timer."""
    import time

    import grpc

    from gpushare_b200 import api
    from gpushare_b200.grpc_harness import KubeletStub
    with api.Nvml() as nv:
        uuid = nv.devices()[0].uuid.decode()
    proc, sock = _start_daemon(tmp_path, [_tenant("big", 80, 1), _tenant("small-a", 2, 2), _tenant("small-b", 3, 3)],
                               interval=3, extra_args=["--probe-gather-ms=1500"])
    try:
        ch = KubeletStub.connect(sock)
        grpc.channel_ready_future(ch).result(timeout=30)
        call = KubeletStub.method(ch, "Allocate")
        for n in (80, 2, 3):
            resp = call(_request(uuid, n))
            assert b"ALIYUN_COM_GPU_MEM_IDX" in resp and b"no-gpu-has" not in resp
        time.sleep(5.5)                                  # the gather window, the sweep, then the health timer
        ch.close()
    finally:
        err = _stop_daemon(proc)
    assert proc.returncode == 0, err[-2000:]
    assert "probe: gpu 0, %d bytes granted to pod 0 queued for a sweep" % (80 << 30) in err
    assert "probe: gpu 0, 3 slice(s), %d bytes in one launch pair" % (85 << 30) in err, err[-3000:]
    assert "in 3 slice(s), 2 launches" in err                       # one store pass, one load pass for all three
    assert "health probe: gpu 0 swept" in err and "0 mismatching words" in err
    assert "mismatching words" in err and " 0 mismatching words" in err
    import re
    assert all(int(m) == 0 for m in re.findall(r"(\d+) mismatching words", err)), err[-3000:]


def test_daemon_health_sweep_catches_a_store_pass_that_did_nothing(tmp_path):
    """Every sweep of the daemon draws a fresh seed, so a store kernel that silently does nothing
    cannot hide behind what the previous sweep left in the same memory. Here the engine drops
    every store pass after its second (the sentinel's and the first health sweep's run): the next
    health sweep must find (nearly) every word wrong and the daemon must take the GPU out of
    service. With one constant seed per sweep (as round 1 had it) that sweep read back clean."""
    import re
    import time

    import grpc

    from gpushare_b200.grpc_harness import KubeletStub
    proc, sock = _start_daemon(tmp_path, [], {"GS_PROBE_TEST_SKIP_STORES": "3"})
    try:
        ch = KubeletStub.connect(sock)
        grpc.channel_ready_future(ch).result(timeout=30)
        time.sleep(4.5)                                  # at least three health ticks
        stream = KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")
        first = next(iter(stream))
        stream.cancel()
        ch.close()
    finally:
        err = _stop_daemon(proc)
    assert proc.returncode == 0, err[-2000:]
    sweeps = [(int(b), int(m)) for b, m in re.findall(r"health probe: gpu 0 swept (\d+) bytes .* (\d+) mismatching words", err)]
    assert len(sweeps) >= 2, err[-3000:]
    assert sweeps[0][1] == 0                              # stores ran: clean
    assert sweeps[1][1] > 0.99 * (sweeps[1][0] // 16)     # stores dropped: the fresh seed exposes it
    assert b"Unhealthy" in first and b"\x12\x07Healthy" not in first   # every device of the GPU went unhealthy


def test_scheduler_sweep_catches_a_store_pass_that_did_nothing():
    """The same property one level down, through the C ABI: ProbeScheduler sweeps a range, then the
    store pass is disabled by the test hook and the SAME range is granted again. The second sweep
    must report every word (its seed is new; the data is the first sweep's)."""
    from gpushare_b200.probe import Probe, ProbeScheduler
    bad = []
    with Probe(0) as p:
        arena = p.alloc(64 * MiB)
        try:
            s = ProbeScheduler(p, on_bad=lambda tag, v: bad.append((tag, v.mismatches)))
            s.grant(arena, 64 * MiB, None, 5)
            s.drain()
            assert bad == [] and s.bytes_probed == 64 * MiB
            p.test_skip_stores(True)
            s.grant(arena, 64 * MiB, None, 5)
            s.drain()
            assert bad == [(5, 64 * MiB // 16)]
            p.test_skip_stores(False)
            s.grant(arena, 64 * MiB, None, 5)
            s.drain()
            assert len(bad) == 1                            # stores back: clean again
            a, b = p.next_seed(), p.next_seed()
            assert a != b
        finally:
            p.free(arena)


def test_daemon_health_sweeps_run_on_all_gpus_at_once(tmp_path):
    """A node's health-time sweep takes far less than the sum of its GPUs' sweeps: every GPU has
    its own worker thread, the ticks fall on the same instant, and allocation, sweep and release of
    one GPU overlap with those of the others. Measured on 8 B200s: the GPUs of a round are done
    362 ... 734 ms after the tick (the driver serialises about 50 ms of each GPU's 176 GiB
    allocation; run one after the other they would take 8 x 362 ms). Needs two GPUs or more."""
    import re
    import time

    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU: nothing to overlap")
    proc, sock = _start_daemon(tmp_path, [], interval=3)
    try:
        time.sleep(11)                                   # three ticks
    finally:
        err = _stop_daemon(proc)
    assert proc.returncode == 0, err[-2000:]
    done = [(int(g), float(ms)) for g, ms in re.findall(r"health probe: gpu (\d+) done (\d+) ms after the tick", err)]
    assert len(done) >= 2 * n, err[-3000:]
    last_round = dict(done[-n:])                         # the last tick: one line per GPU
    assert sorted(last_round) == list(range(n)), done
    fastest, slowest = min(last_round.values()), max(last_round.values())
    # run one after the other the slowest would finish n times later than the fastest; allow a
    # quarter of one GPU's time per further GPU for what the driver serialises
    assert slowest < fastest * (1 + 0.25 * (n - 1)) + 150, last_round
    swept = re.findall(r"health probe: gpu \d+ swept (\d+) bytes .* (\d+) mismatching words", err)
    assert swept and all(int(m) == 0 and int(b) > 100 << 30 for b, m in swept)

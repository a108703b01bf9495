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


@pytest.mark.parametrize("program", ["python", "binary"])
def test_daemon_on_real_gpu_probes_at_allocate_and_health_time(tmp_path, program):
    """The daemon against the real NVML and GPU: enumerate, serve, Allocate for a pod bound to
    GPU 0, sweep the granted range, and sweep free memory on the health timer."""
    import signal
    import time

    import grpc

    from gpushare_b200 import api
    from gpushare_b200.grpc_harness import KubeletStub
    with api.Nvml() as nv:
        uuid = nv.devices()[0].uuid.decode()
    pods = {"items": [{"metadata": {"uid": "u1", "name": "tenant", "namespace": "default", "annotations": {
        "ALIYUN_COM_GPU_MEM_IDX": "0", "ALIYUN_COM_GPU_MEM_ASSUME_TIME": "1",
        "ALIYUN_COM_GPU_MEM_ASSIGNED": "false"}},
        "spec": {"nodeName": "node-a", "containers": [
            {"resources": {"limits": {"aliyun.com/gpu-mem": "2"}}}]}, "status": {"phase": "Pending"}}]}
    (tmp_path / "pods.json").write_text(json.dumps(pods))
    env = dict(os.environ, NODE_NAME="node-a", PYTHONPATH=ROOT)
    exe = [sys.executable, "-m", "gpushare_b200.daemon"] if program == "python" else [
        os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "bin", "gpushare-device-plugin-v2")]
    proc = subprocess.Popen(exe + ["--v=1", "--memory-unit=GiB",
                             "--socket-dir", str(tmp_path), "--pod-source-file", str(tmp_path / "pods.json"),
                             "--no-register", "--probe", "--health-check", "--health-probe-interval=1"],
                            env=env, stderr=subprocess.PIPE, text=True)
    try:
        sock = str(tmp_path / "aliyungpushare.sock")
        deadline = time.time() + 60
        while not os.path.exists(sock) and time.time() < deadline:
            time.sleep(0.1)
        ch = KubeletStub.connect(sock)
        grpc.channel_ready_future(ch).result(timeout=30)
        req = b"\x0a" + bytes([2 * (2 + len(uuid) + 4)]) + b"".join(
            b"\x0a" + bytes([len(uuid) + 4]) + (uuid + "-_-%d" % j).encode() for j in (1, 2))
        resp = KubeletStub.method(ch, "Allocate")(req)
        assert b"ALIYUN_COM_GPU_MEM_IDX" in resp and b"no-gpu-has" not in resp
        time.sleep(3.5)                                  # let the health timer fire
        ch.close()
    finally:
        proc.send_signal(signal.SIGTERM)
        try:
            proc.wait(30)
        except subprocess.TimeoutExpired:
            proc.kill()
    err = proc.stderr.read()
    assert proc.returncode == 0, err[-2000:]
    assert "probe: gpu 0, 2147483648 bytes granted to pod 0 queued for a sweep" in err
    assert "health probe: gpu 0 swept" in err and "0 mismatching words" in err

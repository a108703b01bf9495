"""The daemon's command line (cmd/nvidia/main.go:15-26) and an end-to-end run against the
fake NVML over a real socket: BASELINE config 1."""
import json
import os
import signal
import subprocess
import sys
import tempfile
import time

import grpc
import pytest

from gpushare_b200 import daemon
from gpushare_b200.grpc_harness import KubeletStub
from oracle import ref_model as R
from tests.conftest import FAKE_NVML, ROOT

NATIVE_BIN = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "bin", "gpushare-device-plugin-v2")
# the same program twice: daemon.py over either transport, and the native binary (daemon_cpp/)
PROGRAMS = {"python-native": [sys.executable, "-m", "gpushare_b200.daemon", "--transport", "native"],
            "python-grpcio": [sys.executable, "-m", "gpushare_b200.daemon", "--transport", "grpcio"],
            "binary": [NATIVE_BIN]}


def test_flag_names_and_defaults_match_reference():
    a = daemon.parse_args([])
    assert (a.mps, a.health_check, a.memory_unit, a.query_kubelet) == (False, False, "GiB", False)
    assert (a.kubelet_address, a.kubelet_port, a.client_cert, a.client_key, a.token, a.timeout) == \
        ("0.0.0.0", 10250, "", "", "", 10)
    # the DaemonSet's arguments (device-plugin-ds.yaml:29-33) and Go's single-dash form
    a = daemon.parse_args(["-logtostderr", "--v=5", "--memory-unit=GiB"])
    assert a.logtostderr and a.v == 5
    a = daemon.parse_args(["-memory-unit", "MiB", "-health-check", "-query-kubelet=false", "--mps=true"])
    assert a.memory_unit == "MiB" and a.health_check and not a.query_kubelet and a.mps


@pytest.mark.parametrize("program", sorted(PROGRAMS))
def test_daemon_config1_end_to_end(golden_cases, program):
    c = next(x for x in golden_cases if x["name"] == "G1-binpack-1")
    with tempfile.TemporaryDirectory(prefix="gsd") as d:
        pods = os.path.join(d, "pods.json")
        json.dump({"items": c["pods"]}, open(pods, "w"))
        kubelet = KubeletStub(os.path.join(d, "kubelet.sock"))
        kubelet.start()
        env = dict(os.environ, NODE_NAME=c["node"], FAKE_NVML_GPUS="8", FAKE_NVML_MEM_MIB="16384",
                   PYTHONPATH=ROOT)
        proc = subprocess.Popen(PROGRAMS[program] + ["-logtostderr", "--v=5",
                                 "--memory-unit=GiB", "--nvml-lib", FAKE_NVML, "--socket-dir", d,
                                 "--pod-source-file", pods, "--health-check"],
                                env=env, stderr=subprocess.PIPE, text=True)
        try:
            assert kubelet.wait_registered(30), proc.stderr.read() if proc.poll() is not None else ""
            ch = KubeletStub.connect(os.path.join(d, "aliyungpushare.sock"))
            grpc.channel_ready_future(ch).result(timeout=10)
            devs = R.decode_list_and_watch(next(KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")))
            assert len(devs) == 128 and devs[0][0].endswith("-_-0")
            st = c["steps"][0]
            resp = KubeletStub.method(ch, "Allocate")(R.encode_allocate_request(st["request"]))
            assert R.decode_allocate_response(resp) == st["expect_envs"]
            ch.close()
        finally:
            proc.send_signal(signal.SIGTERM)
            try:
                proc.wait(20)
            except subprocess.TimeoutExpired:
                proc.kill()
            kubelet.stop()
        assert proc.returncode == 0


@pytest.mark.parametrize("program", ["python-native", "binary"])
def test_daemon_requires_node_name(program):
    env = {k: v for k, v in os.environ.items() if k != "NODE_NAME"}
    env["PYTHONPATH"] = ROOT
    out = subprocess.run(PROGRAMS[program] + ["--nvml-lib", FAKE_NVML],
                         env=env, capture_output=True, text=True, timeout=60)
    assert out.returncode != 0 and "Please set env NODE_NAME" in out.stderr


@pytest.mark.parametrize("program", ["python-native", "binary"])
def test_supervisor_reregisters_on_kubelet_restart_and_sighup(golden_cases, program):
    """gpumanager.go:83-96: kubelet.sock re-created => tear down, re-enumerate, register again;
    SIGHUP => the same."""
    c = next(x for x in golden_cases if x["name"] == "G1-binpack-1")
    with tempfile.TemporaryDirectory(prefix="gsd") as d:
        pods = os.path.join(d, "pods.json")
        json.dump({"items": c["pods"]}, open(pods, "w"))
        kubelet = KubeletStub(os.path.join(d, "kubelet.sock"))
        kubelet.start()
        memfile = os.path.join(d, "mem_mib")
        open(memfile, "w").write("4096")
        env = dict(os.environ, NODE_NAME=c["node"], FAKE_NVML_GPUS="2", FAKE_NVML_MEM_FILE=memfile,
                   PYTHONPATH=ROOT)
        proc = subprocess.Popen(PROGRAMS[program] + ["--v=1", "--nvml-lib", FAKE_NVML,
                                 "--socket-dir", d, "--pod-source-file", pods],
                                env=env, stderr=subprocess.PIPE, text=True)
        try:
            assert kubelet.wait_registered(30)
            # kubelet restarts: its socket disappears and comes back
            kubelet.stop()
            time.sleep(0.3)
            kubelet2 = KubeletStub(os.path.join(d, "kubelet.sock"))
            kubelet2.start()
            assert kubelet2.wait_registered(15), "no re-registration after kubelet.sock was re-created"
            n_before = len(kubelet2.registered)
            # H7: the unit count is a process global of the reference (nvidia.go:12-15,67-69), set from
            # GPU 0 once and kept across restarts: GPUs that report twice the memory after the restart
            # are still advertised with 4 units each
            open(memfile, "w").write("8192")
            proc.send_signal(signal.SIGHUP)
            deadline = time.time() + 15
            while len(kubelet2.registered) == n_before and time.time() < deadline:
                time.sleep(0.05)
            assert len(kubelet2.registered) >= n_before + 1, "no re-registration after SIGHUP"
            ch = KubeletStub.connect(os.path.join(d, "aliyungpushare.sock"))
            grpc.channel_ready_future(ch).result(timeout=10)
            devs = R.decode_list_and_watch(next(KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")))
            assert len(devs) == 8                        # 2 GPUs x 4 GiB, as before the restart
            ch.close()
            kubelet2.stop()
        finally:
            proc.send_signal(signal.SIGTERM)
            try:
                proc.wait(20)
            except subprocess.TimeoutExpired:
                proc.kill()
        assert proc.returncode == 0
        err = proc.stderr.read()
        assert "kubelet.sock created, restarting." in err and "Received SIGHUP, restarting." in err


@pytest.mark.parametrize("program", ["python-native", "binary"])
def test_daemon_health_events_reach_the_watch_stream(program):
    """BASELINE config 5 through a whole daemon: XID 79 on GPU 3 re-advertises that GPU's 16
    devices as Unhealthy, XID 31 is ignored (nvidia.go:134), an event without a device takes
    every GPU down (nvidia.go:138-143); Allocate keeps answering throughout."""
    from gpushare_b200.native_transport import NativeClient
    with tempfile.TemporaryDirectory(prefix="gsd") as d:
        pods = os.path.join(d, "pods.json")
        json.dump({"items": []}, open(pods, "w"))
        fifo = os.path.join(d, "xid.fifo")
        os.mkfifo(fifo)
        env = dict(os.environ, NODE_NAME="node-a", FAKE_NVML_GPUS="8", FAKE_NVML_MEM_MIB="16384",
                   FAKE_NVML_EVENT_FIFO=fifo, PYTHONPATH=ROOT)
        proc = subprocess.Popen(PROGRAMS[program] + ["--v=1", "--nvml-lib", FAKE_NVML, "--socket-dir", d,
                                                      "--pod-source-file", pods, "--health-check", "--no-register"],
                                env=env, stderr=subprocess.PIPE, text=True)
        try:
            sock = os.path.join(d, "aliyungpushare.sock")
            deadline = time.time() + 30
            while not os.path.exists(sock) and time.time() < deadline:
                time.sleep(0.05)
            with NativeClient(sock, timeout_ms=10000) as cl, open(fifo, "w") as events:
                sid = cl.open_stream("ListAndWatch")
                first = R.decode_list_and_watch(cl.next_message(sid))
                assert len(first) == 128 and all(h == "Healthy" for _, h in first)
                time.sleep(0.3)                         # the health thread is up by now
                events.write("3 79\n")
                events.flush()
                t0 = time.time()
                second = R.decode_list_and_watch(cl.next_message(sid, timeout_ms=5000))
                assert time.time() - t0 < 2.5
                assert [i for i, h in second if h == "Unhealthy"] == [i for i, _ in first[48:64]]
                events.write("1 31\n")                 # an application error: not the GPU's fault
                events.flush()
                with pytest.raises(TimeoutError):
                    cl.next_message(sid, timeout_ms=1500)
                assert b"no-gpu-has-1GiB-to-run" in cl.call("Allocate", R.encode_allocate_request([["x"]]))
                events.write("-1 48\n")
                events.flush()
                third = R.decode_list_and_watch(cl.next_message(sid, timeout_ms=5000))
                assert all(h == "Unhealthy" for _, h in third)
        finally:
            proc.send_signal(signal.SIGTERM)
            try:
                proc.wait(20)
            except subprocess.TimeoutExpired:
                proc.kill()
        assert proc.returncode == 0, proc.stderr.read()[-2000:]

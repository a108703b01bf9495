"""The kubelet-facing surface end to end over a unix socket: registration,
the five RPCs, a streamed ListAndWatch with a health event, the 4 MiB limit (H12)."""
import os
import tempfile
import threading
import time

import grpc
import pytest

from gpushare_b200 import api
from gpushare_b200.grpc_harness import DevicePluginServer, KubeletStub, encode_register_request
from oracle import ref_model as R
from tests.helpers import product_for_case


@pytest.fixture
def sockdir():
    with tempfile.TemporaryDirectory(prefix="gs") as d:
        yield d


def test_register_and_rpcs(golden_cases, sockdir):
    c = next(x for x in golden_cases if x["name"] == "G1-binpack-1")
    table, plug = product_for_case(c)
    kubelet = KubeletStub(os.path.join(sockdir, "kubelet.sock"))
    kubelet.start()
    srv = DevicePluginServer(plug, os.path.join(sockdir, "aliyungpushare.sock"))
    try:
        srv.Serve(kubelet.socket)
        assert kubelet.wait_registered(5)
        # RegisterRequest{version, endpoint, resource_name} as server.go:166-170
        fields = {f: v for f, _w, v in R._fields(kubelet.registered[0])}
        assert fields == {1: b"v1beta1", 2: b"aliyungpushare.sock", 3: b"aliyun.com/gpu-mem"}
        assert kubelet.registered[0] == encode_register_request("v1beta1", "aliyungpushare.sock",
                                                                "aliyun.com/gpu-mem")
        ch = KubeletStub.connect(srv.socket)
        assert KubeletStub.method(ch, "GetDevicePluginOptions")(b"") == b""
        assert KubeletStub.method(ch, "PreStartContainer")(b"\x0a\x03abc") == b""
        with pytest.raises(grpc.RpcError):
            KubeletStub.method(ch, "GetPreferredAllocation")(b"")
        alloc = KubeletStub.method(ch, "Allocate")
        for st in c["steps"]:
            resp = alloc(R.encode_allocate_request(st["request"]))
            assert R.decode_allocate_response(resp) == st["expect_envs"]
        # ListAndWatch: full list, then one re-send after an XID on GPU 3
        stream = KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")
        first = R.decode_list_and_watch(next(stream))
        assert len(first) == 128 and all(h == "Healthy" for _, h in first)
        assert first[17] == ("GPU-00000001-0000-0000-0000-000000000001-_-1", "Healthy")
        plug.unhealthy(3)
        second = R.decode_list_and_watch(next(stream))
        assert [h for i, h in second if i.startswith("GPU-00000003")] == ["Unhealthy"] * 16
        assert sum(h == "Unhealthy" for _, h in second) == 16
        stream.cancel()
        ch.close()
    finally:
        srv.Stop()
        kubelet.stop()
    assert not os.path.exists(srv.socket)                     # cleanup() of server.go:203-209


def test_stale_socket_is_replaced_and_register_failure_stops(sockdir):
    table = api.DeviceTable(["GPU-a"], 4)
    plug = api.NvidiaDevicePlugin(table, "n")
    path = os.path.join(sockdir, "aliyungpushare.sock")
    open(path, "w").close()                                   # stale file in the way
    srv = DevicePluginServer(plug, path)
    with pytest.raises(Exception):
        srv.Serve(os.path.join(sockdir, "no-kubelet.sock"))    # Register fails => Stop
    assert srv.server is None and plug.stopped and not os.path.exists(path)


def test_mib_payload_exceeds_default_grpc_limit(sockdir):
    """H12: one B200 at MiB granularity is 11.3 MB; a stock client (4 MiB) refuses it, a
    client with the limit lifted gets every device."""
    uuid = "GPU-%08x-0000-0000-0000-%012x" % (0, 0)
    table = api.DeviceTable([uuid], 183359, "MiB")
    plug = api.NvidiaDevicePlugin(table, "n")
    srv = DevicePluginServer(plug, os.path.join(sockdir, "aliyungpushare.sock"))
    srv.Start()
    try:
        ch = KubeletStub.connect(srv.socket, unlimited=False)
        with pytest.raises(grpc.RpcError) as e:
            next(KubeletStub.method(ch, "ListAndWatch", stream=True)(b""))
        assert e.value.code() == grpc.StatusCode.RESOURCE_EXHAUSTED
        ch.close()
        ch = KubeletStub.connect(srv.socket)
        t0 = time.time()
        stream = KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")
        msg = next(stream)
        dt = time.time() - t0
        assert len(msg) == table.payload_size() == sum(13 + 43 + len(str(j)) for j in range(183359))
        assert msg[:2] == b"\x0a\x37" and msg.count(b"Healthy") == 183359
        assert dt < 5
        stream.cancel()
        ch.close()
    finally:
        srv.Stop()


def test_concurrent_allocates_are_serialised(sockdir):
    """allocate.go:60: pod claiming is serialised -- 8 clients racing for 8 equal pods get
    8 distinct pods."""
    n = 8
    pods = [api.Pod("u%d" % k, "default", "p%d" % k, "n", 2, str(k), str(100 + k), "false")
            for k in range(n)]
    table = api.DeviceTable(["GPU-%d" % i for i in range(8)], 16)
    plug = api.NvidiaDevicePlugin(table, "n")
    plug.set_pods(pods)
    srv = DevicePluginServer(plug, os.path.join(sockdir, "aliyungpushare.sock"))
    srv.Start()
    got = []
    try:
        def client():
            ch = KubeletStub.connect(srv.socket)
            resp = KubeletStub.method(ch, "Allocate")(R.encode_allocate_request([["a", "b"]]))
            got.append(R.decode_allocate_response(resp)[0]["ALIYUN_COM_GPU_MEM_IDX"])
            ch.close()
        ts = [threading.Thread(target=client) for _ in range(n)]
        [t.start() for t in ts]
        [t.join(10) for t in ts]
    finally:
        srv.Stop()
    assert sorted(got) == [str(k) for k in range(n)]

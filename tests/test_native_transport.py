"""The native gRPC transport (csrc/hpack.h, h2conn.h, grpc_server.cpp, grpc_client.cpp):
RFC 7541 known answers for the header codec, then interoperability in all four pairings of
{grpcio, native} x {client, server} over real unix sockets, on the reference's own surface
(Register, the five RPCs, a streamed ListAndWatch with a health event)."""
import os
import subprocess
import tempfile
import threading
import time

import grpc
import pytest

from gpushare_b200 import _lib as L
from gpushare_b200 import api
from gpushare_b200.grpc_harness import DevicePluginServer, KubeletStub, encode_register_request
from gpushare_b200.native_transport import NativeClient, NativeDevicePluginServer, RpcError
from oracle import ref_model as R
from tests.helpers import product_for_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "csrc")


@pytest.fixture
def sockdir():
    with tempfile.TemporaryDirectory(prefix="gs") as d:
        yield d


def test_hpack_rfc7541_known_answers(tmp_path):
    """Appendix C examples (integers, literals, the request series with and without Huffman,
    the response series with evictions) plus malformed blocks, under ASan/UBSan."""
    exe = str(tmp_path / "hpack_kat")
    subprocess.run(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined",
                    "-fno-sanitize-recover=undefined", "-I" + CSRC,
                    os.path.join(ROOT, "tests", "cpp", "hpack_kat.cpp"), "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("ok  ") == 18 and "FAIL" not in r.stdout
    # The block with all 256 byte values Huffman-coded, as our table encodes it, goes through an
    # independent HPACK implementation: nghttp2's inflater (a system library here; test peer only)
    # must decode it to the same bytes.
    block = bytes.fromhex(next(l for l in r.stdout.splitlines() if l.startswith("block ")).split()[1])
    want = bytes((i * 7 + 3) & 0xFF for i in range(256))
    import ctypes as C
    import ctypes.util
    path = ctypes.util.find_library("nghttp2") or "libnghttp2.so.14"
    try:
        ng = C.CDLL(path)
    except OSError:
        pytest.skip("libnghttp2 not on this machine")

    class NV(C.Structure):
        _fields_ = [("name", C.POINTER(C.c_uint8)), ("value", C.POINTER(C.c_uint8)), ("namelen", C.c_size_t),
                    ("valuelen", C.c_size_t), ("flags", C.c_uint8)]
    infl = C.c_void_p()
    assert ng.nghttp2_hd_inflate_new(C.byref(infl)) == 0
    ng.nghttp2_hd_inflate_hd2.restype = C.c_ssize_t
    ng.nghttp2_hd_inflate_hd2.argtypes = [C.c_void_p, C.POINTER(NV), C.POINTER(C.c_int), C.c_char_p, C.c_size_t, C.c_int]
    nv, flags = NV(), C.c_int()
    used = ng.nghttp2_hd_inflate_hd2(infl, C.byref(nv), C.byref(flags), block, len(block), 1)
    assert used == len(block) and flags.value & 2                      # NGHTTP2_HD_INFLATE_EMIT
    assert bytes(nv.name[:nv.namelen]) == b"x-bin" and bytes(nv.value[:nv.valuelen]) == want
    ng.nghttp2_hd_inflate_del.argtypes = [C.c_void_p]
    ng.nghttp2_hd_inflate_del(infl)


def _golden(golden_cases, name="G1-binpack-1"):
    return next(x for x in golden_cases if x["name"] == name)


def test_grpcio_client_against_native_server(golden_cases, sockdir):
    """What kubelet does, with grpc-core as the client (Huffman strings, a dynamic table,
    BDP pings): registration, every RPC, the watch stream, stop."""
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    kubelet = KubeletStub(os.path.join(sockdir, "kubelet.sock"))
    kubelet.start()
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "aliyungpushare.sock"))
    assert not srv.python_allocate
    try:
        srv.Serve(kubelet.socket)
        assert kubelet.wait_registered(5)
        assert kubelet.registered[0] == encode_register_request("v1beta1", "aliyungpushare.sock",
                                                                "aliyun.com/gpu-mem")
        ch = KubeletStub.connect(srv.socket)
        assert KubeletStub.method(ch, "GetDevicePluginOptions")(b"") == b""
        assert KubeletStub.method(ch, "PreStartContainer")(b"\x0a\x03abc") == b""
        with pytest.raises(grpc.RpcError) as e:
            KubeletStub.method(ch, "GetPreferredAllocation")(b"")
        assert e.value.code() == grpc.StatusCode.UNIMPLEMENTED and e.value.details() == "implement me"
        with pytest.raises(grpc.RpcError) as e:
            KubeletStub.method(ch, "NoSuchMethod")(b"")
        assert e.value.code() == grpc.StatusCode.UNIMPLEMENTED
        assert e.value.details() == "unknown method NoSuchMethod for service v1beta1.DevicePlugin"
        with pytest.raises(grpc.RpcError) as e:
            ch.unary_unary("/v1beta1.Registration/Register")(b"")
        assert e.value.details() == "unknown service v1beta1.Registration"
        alloc = KubeletStub.method(ch, "Allocate")
        for st in c["steps"]:                        # same socket, same answers as the fixture
            resp = alloc(R.encode_allocate_request(st["request"]))
            assert R.decode_allocate_response(resp) == st["expect_envs"]
        with pytest.raises(grpc.RpcError) as e:      # bytes that are not an AllocateRequest
            alloc(b"\x0a\xff\xff\xff\xff\x0f")
        assert e.value.code() == grpc.StatusCode.INTERNAL
        stream = KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")
        first = R.decode_list_and_watch(next(stream))
        assert len(first) == 128 and all(h == "Healthy" for _, h in first)
        # Allocate keeps working on the same connection while the stream is open
        assert R.decode_allocate_response(alloc(R.encode_allocate_request(c["steps"][-1]["request"]))) \
            == c["steps"][-1]["expect_envs"]
        plug.unhealthy(3)
        second = R.decode_list_and_watch(next(stream))
        assert [h for i, h in second if i.startswith("GPU-00000003")] == ["Unhealthy"] * 16
        assert sum(h == "Unhealthy" for _, h in second) == 16
        st = srv.stats()
        assert st["law_messages"] == 2 and st["allocate_calls"] == len(c["steps"]) + 2
        assert st["protocol_errors"] == 0
        # Stop: the stream ends with status OK (ListAndWatch returns nil on m.stop)
        srv.Stop()
        assert list(stream) == []
        assert stream.code() == grpc.StatusCode.OK
        ch.close()
    finally:
        srv.Stop()
        kubelet.stop()
    assert not os.path.exists(srv.socket)


def test_native_client_against_grpcio_server(golden_cases, sockdir):
    """The native client against grpc-core as the server: our Register against the kubelet
    stub, and the plugin surface served by the grpcio harness."""
    lib = L.load()
    kubelet = KubeletStub(os.path.join(sockdir, "kubelet.sock"))
    kubelet.start()
    try:
        assert lib.gs_register(kubelet.socket.encode(), b"aliyungpushare.sock", b"aliyun.com/gpu-mem", 5000) == 0
        assert kubelet.registered == [encode_register_request("v1beta1", "aliyungpushare.sock",
                                                              "aliyun.com/gpu-mem")]
        # a method the kubelet stub does not serve: the status comes back, not a hang
        with NativeClient(kubelet.socket) as cl:
            with pytest.raises(RpcError) as e:
                cl.call("Nope", b"", service="v1beta1.Registration")
            assert e.value.code == 12
    finally:
        kubelet.stop()
    t0 = time.time()
    assert lib.gs_register(os.path.join(sockdir, "absent.sock").encode(), b"e", b"r", 300) == L.GS_E_TIMEOUT
    assert 0.25 < time.time() - t0 < 3

    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    srv = DevicePluginServer(plug, os.path.join(sockdir, "aliyungpushare.sock"))
    srv.Start()
    try:
        with NativeClient(srv.socket) as cl:
            assert cl.call("GetDevicePluginOptions") == b""
            for st in c["steps"]:
                assert R.decode_allocate_response(cl.call("Allocate", R.encode_allocate_request(st["request"]))) \
                    == st["expect_envs"]
            with pytest.raises(RpcError) as e:
                cl.call("GetPreferredAllocation")
            assert (e.value.code, e.value.message) == (12, "implement me")
            sid = cl.open_stream("ListAndWatch")
            assert len(R.decode_list_and_watch(cl.next_message(sid))) == 128
            assert cl.call("PreStartContainer", b"\x0a\x01x") == b""      # unary beside the stream
            plug.unhealthy(1)
            second = R.decode_list_and_watch(cl.next_message(sid))
            assert sum(h == "Unhealthy" for _, h in second) == 16
            with pytest.raises(TimeoutError):
                cl.next_message(sid, timeout_ms=100)
            cl.cancel(sid)
    finally:
        srv.Stop()


def test_native_both_sides_large_messages(sockdir):
    """One B200 at MiB granularity: an 11.3 MB ListAndWatchResponse out (flow control, frame
    splitting, borrowed segments) and a 2 MB AllocateRequest in, byte-exact."""
    uuid = "GPU-0f9b9cf4-2b1c-4a3e-9c1d-5e2f6a7b8c9d"
    units = 183359
    table = api.DeviceTable([uuid], units, api.MiB)
    plug = api.NvidiaDevicePlugin(table, "node-a")
    pods = []
    for k, gib in enumerate((40, 1)):
        pods.append({"metadata": {"uid": "u%d" % k, "namespace": "default", "name": "t%d" % k, "annotations": {
            "ALIYUN_COM_GPU_MEM_IDX": "0", "ALIYUN_COM_GPU_MEM_ASSUME_TIME": str(1000 + k),
            "ALIYUN_COM_GPU_MEM_ASSIGNED": "false"}},
            "spec": {"nodeName": "node-a", "containers": [{"name": "c", "resources": {"limits": {
                "aliyun.com/gpu-mem": str(gib * 1024)}}}]}, "status": {"phase": "Pending"}})
    plug.set_pods([api.Pod.from_k8s(p) for p in pods])
    seen = []
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"),
                                   on_decision=lambda *a: seen.append(a))
    srv.Start()
    try:
        with NativeClient(srv.socket) as cl:
            sid = cl.open_stream("ListAndWatch")
            want = table.list_and_watch_bytes()
            assert len(want) == 11257148
            got = cl.next_message(sid, timeout_ms=20000)
            assert got == want
            ids = ["%s-_-%d" % (uuid, j) for j in range(40 * 1024)]
            req = R.encode_allocate_request([ids])
            assert len(req) > 2_000_000
            envs = R.decode_allocate_response(cl.call("Allocate", req))
            assert envs == [{"NVIDIA_VISIBLE_DEVICES": "0", "ALIYUN_COM_GPU_MEM_IDX": "0",
                             "ALIYUN_COM_GPU_MEM_POD": "40960", "ALIYUN_COM_GPU_MEM_CONTAINER": "40960",
                             "ALIYUN_COM_GPU_MEM_DEV": str(units)}]
            assert seen == [("matched", 0, 0, 40960)]
            assert plug.get_pod(0).ann_assigned == "true"
            # a health change while nothing is in flight: exactly one fresh list
            plug.unhealthy(0)
            second = cl.next_message(sid, timeout_ms=20000)
            assert second == table.list_and_watch_bytes() and len(second) == len(want) + 2 * units
            with pytest.raises(TimeoutError):
                cl.next_message(sid, timeout_ms=200)
        # the same list through grpc-core's flow control (64 KiB windows growing by BDP)
        ch = KubeletStub.connect(srv.socket)
        stream = KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")
        assert next(stream) == table.list_and_watch_bytes()
        stream.cancel()
        ch.close()
    finally:
        srv.Stop()
    # a request beyond the limit is refused with a status, and the connection survives
    plug2 = api.NvidiaDevicePlugin(table, "node-a")
    small = NativeDevicePluginServer(plug2, os.path.join(sockdir, "q.sock"), max_request_bytes=4096)
    small.Start()
    try:
        with NativeClient(small.socket) as cl:
            with pytest.raises(RpcError) as e:
                cl.call("Allocate", req)
            assert e.value.code == 8 and "larger than max" in e.value.message
            assert cl.call("GetDevicePluginOptions") == b""
        ch = KubeletStub.connect(small.socket)
        with pytest.raises(grpc.RpcError) as e:
            KubeletStub.method(ch, "Allocate")(req)
        assert e.value.code() == grpc.StatusCode.RESOURCE_EXHAUSTED
        assert KubeletStub.method(ch, "GetDevicePluginOptions")(b"") == b""
        ch.close()
    finally:
        small.Stop()


def test_large_and_oddly_shaped_requests_over_the_socket_answer_like_in_process(sockdir):
    """Large Allocate bodies (64 KiB and up: several recv()s, the multi-chain id scan) over the
    socket. Whatever the request looks like -- kubelet's shape, several containers, ids of 128+
    bytes, foreign fields, a container that runs past the end, trailing junk -- the answer must be
    the one the in-process call gives on the same bytes: same envs and claim, or the same refusal."""
    import random
    uuid = "GPU-0f9b9cf4-2b1c-4a3e-9c1d-5e2f6a7b8c9d"
    units = 183359
    rng = random.Random(64)

    def pod(k, units_):
        return {"metadata": {"uid": "u%d" % k, "namespace": "default", "name": "t%d" % k, "annotations": {
            "ALIYUN_COM_GPU_MEM_IDX": "0", "ALIYUN_COM_GPU_MEM_ASSUME_TIME": str(1000 + k),
            "ALIYUN_COM_GPU_MEM_ASSIGNED": "false"}},
            "spec": {"nodeName": "node-a", "containers": [{"name": "c", "resources": {"limits": {
                "aliyun.com/gpu-mem": str(units_)}}}]}, "status": {"phase": "Pending"}}

    def ids(n, width=0):
        return [("%s-_-%d" % (uuid, rng.randrange(units))).ljust(width, "x") for _ in range(n)]

    def varint(n):
        out = bytearray()
        while n >= 0x80:
            out.append((n & 0x7F) | 0x80)
            n >>= 7
        out.append(n)
        return bytes(out)

    cases = []                                            # (name, wire)
    cases.append(("one container, 3000 ids", R.encode_allocate_request([ids(3000)])))
    cases.append(("exactly one piece", None))             # filled below: body of exactly 64 KiB
    cases.append(("three containers", R.encode_allocate_request([ids(2000), ids(1), ids(4096)])))
    cases.append(("empty container between", R.encode_allocate_request([ids(1500), [], ids(1500)])))
    cases.append(("ids of 130 bytes", R.encode_allocate_request([ids(1200, 130)])))
    cases.append(("one long id among short", R.encode_allocate_request([ids(1000) + ids(1, 300) + ids(1000)])))
    base = R.encode_allocate_request([ids(2500)])
    cases.append(("foreign varint field after", base + b"\x10\x07"))
    cases.append(("foreign bytes field before", b"\x12\x03abc" + base))
    cases.append(("trailing junk", base + b"\xff"))
    cases.append(("container runs past the end", b"\x0a" + varint(len(base) + 10) + base[1 + len(varint(len(base) - 1 - len(varint(len(base))))):]))
    inner = b"".join(b"\x0a" + bytes([len(i)]) + i.encode() for i in ids(2000))
    cases.append(("foreign field inside a container", b"\x0a" + varint(len(inner) + 2) + inner[:50000] + b"\x18\x01" + inner[50000:]))
    cases.append(("id length runs past its container", b"\x0a" + varint(len(inner)) + inner[:-1] + b"\x7f"))
    cases.append(("non-canonical container length", b"\x0a" + varint(len(inner))[:-1] + bytes([varint(len(inner))[-1] | 0x80, 0x00]) + inner))
    # a body of exactly 64 KiB with the 5-byte gRPC prefix: 1 258 records of 52 bytes and
    # one of 111 make 65 527 bytes of ids, + tag + 3-byte length = 65 531
    n_ids = 1259
    body = R.encode_allocate_request([ids(1258, 50) + ids(1, 109)])
    assert len(body) + 5 == 65536
    cases[1] = ("exactly one piece", body)

    table = api.DeviceTable([uuid], units, api.MiB)
    outcomes = []
    for name, wire in cases:
        assert len(wire) + 5 >= 65536, name               # a large body
        twin_pods = [api.Pod.from_k8s(pod(k, u)) for k, u in enumerate((3000, 6097, 1200, 2001, 1300, 2500, n_ids))]
        served, local = api.NvidiaDevicePlugin(table, "node-a"), api.NvidiaDevicePlugin(table, "node-a")
        served.set_clock(9)
        local.set_clock(9)
        served.set_pods(twin_pods)
        local.set_pods(twin_pods)
        srv = NativeDevicePluginServer(served, os.path.join(sockdir, "s.sock"))
        srv.Start()
        try:
            with NativeClient(srv.socket) as cl:
                try:
                    got = ("ok", cl.call("Allocate", wire))
                except RpcError as e:
                    got = ("status", e.code)
            try:
                d = local.Allocate(wire)
                want = ("ok", d.response)
            except L.GpushareError as e:
                want = ("status", 13 if e.status == L.GS_E_WIRE else 2)
            assert got == want, name
            assert [served.get_pod(i).ann_assigned for i in range(7)] == \
                [local.get_pod(i).ann_assigned for i in range(7)], name
            outcomes.append((name, got[0], "true" in [served.get_pod(i).ann_assigned for i in range(7)]))
        finally:
            srv.Stop()
    claimed = [n for n, kind, took in outcomes if kind == "ok" and took]
    refused = [n for n, kind, _t in outcomes if kind == "status"]
    assert len(claimed) >= 7 and len(refused) >= 3, outcomes   # both kinds of answer were exercised


def test_python_allocate_behind_native_server(golden_cases, sockdir):
    """A plugin whose Allocate is Python (the cluster plugin, the daemon's probe hook) is
    called back with the request bytes; its failures become statuses, not dead connections."""
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    decisions = []
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"), on_allocate=decisions.append)
    assert srv.python_allocate
    srv.Start()
    try:
        with NativeClient(srv.socket) as cl:
            for st in c["steps"]:
                assert R.decode_allocate_response(cl.call("Allocate", R.encode_allocate_request(st["request"]))) \
                    == st["expect_envs"]
            assert [d.kind for d in decisions][:3] == ["matched"] * 3
            with pytest.raises(RpcError) as e:
                cl.call("Allocate", b"\x0a\xff\xff\xff\xff\x0f")
            assert e.value.code == 13
            plug.Allocate = lambda request: 1 / 0
            with pytest.raises(RpcError) as e:
                cl.call("Allocate", b"")
            assert e.value.code == 2
            assert cl.call("GetDevicePluginOptions") == b""
    finally:
        srv.Stop()


def test_slow_python_allocate_does_not_hold_up_the_watch_stream(golden_cases, sockdir):
    """An Allocate callback that sits on the network (lister ladders) runs on the server's
    worker: health re-advertisements and other RPCs go out meanwhile; Allocates stay in order."""
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    inner = plug.Allocate
    order = []

    def slow(request):
        order.append(len(request))
        time.sleep(0.8)
        return inner(request)
    plug.Allocate = slow
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"))
    srv.Start()
    try:
        watcher = NativeClient(srv.socket)
        sid = watcher.open_stream("ListAndWatch")
        watcher.next_message(sid)
        results = {}

        def caller(name, req):
            with NativeClient(srv.socket) as cl:
                results[name] = R.decode_allocate_response(cl.call("Allocate", req))
        r1 = R.encode_allocate_request(c["steps"][0]["request"])
        r2 = R.encode_allocate_request(c["steps"][1]["request"]) + b""
        t1 = threading.Thread(target=caller, args=("first", r1))
        t1.start()
        time.sleep(0.15)
        t2 = threading.Thread(target=caller, args=("second", r2))
        t2.start()
        time.sleep(0.1)
        t0 = time.time()
        plug.unhealthy(0)
        msg = watcher.next_message(sid, timeout_ms=3000)
        assert time.time() - t0 < 0.65 and msg.count(b"Unhealthy") == 16    # not behind the callback
        assert watcher.call("GetDevicePluginOptions") == b""
        assert time.time() - t0 < 0.7
        t1.join(10)
        t2.join(10)
        assert results["first"] == c["steps"][0]["expect_envs"]
        assert results["second"] == c["steps"][1]["expect_envs"]
        assert len(order) == 2
        watcher.close()
    finally:
        srv.Stop()


def test_many_calls_and_two_clients(golden_cases, sockdir):
    """Several hundred calls on one connection (HPACK dynamic table churn on the grpc-core
    side), two connections at once, and a client that vanishes mid-stream."""
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"))
    srv.Start()
    try:
        req = R.encode_allocate_request(c["steps"][0]["request"])
        ch = KubeletStub.connect(srv.socket)
        alloc = KubeletStub.method(ch, "Allocate")
        first = alloc(req, metadata=(("x-trace", "0"),))
        errors = []

        def hammer():
            try:
                with NativeClient(srv.socket) as cl:
                    for _ in range(300):
                        R.decode_allocate_response(cl.call("Allocate", req))
            except Exception as e:                      # noqa: BLE001
                errors.append(e)
        t = threading.Thread(target=hammer)
        t.start()
        for i in range(300):
            r = alloc(req, metadata=(("x-trace", str(i)), ("x-long", "v" * (i % 200))))
            assert len(R.decode_allocate_response(r)) == len(R.decode_allocate_response(first)) == 1
        t.join(30)
        assert not errors and not t.is_alive()
        # a client that opens the watch stream and disappears
        cl = NativeClient(srv.socket)
        sid = cl.open_stream("ListAndWatch")
        cl.next_message(sid)
        cl.close()
        assert KubeletStub.method(ch, "GetDevicePluginOptions")(b"") == b""
        plug.unhealthy(0)                               # nobody is listening: must not wedge
        assert KubeletStub.method(ch, "GetDevicePluginOptions")(b"") == b""
        assert srv.stats()["protocol_errors"] == 0
        ch.close()
    finally:
        srv.Stop()


def test_server_rejects_garbage_and_survives(golden_cases, sockdir):
    """Bytes that are not HTTP/2, a bad preface, an oversized header frame, a truncated
    frame: the connection is dropped, the server keeps serving others."""
    import socket as pysock
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"))
    srv.Start()
    preface = b"PRI * HTTP/2.0\r\n\r\nSM\r\n\r\n"
    try:
        for junk in (b"GET / HTTP/1.1\r\nHost: x\r\n\r\n",
                     preface + b"\x00\x00\x00\x04\x00\x00\x00\x00\x00" + b"\x01\xff\xff\x01\x05\x00\x00\x00\x01" + b"x" * 16,
                     preface + b"\x00\x00\x05\x01\x04\x00\x00\x00\x01\x80\x80\x80\x80\x80",   # HPACK: index 0
                     preface + b"\x00\x00\x04\x08\x00\x00\x00\x00\x00\x00\x00\x00\x00",       # WINDOW_UPDATE 0
                     preface + b"\x00\x00\x08\x06\x00\x00\x00\x00\x01" + b"12345678",          # PING on a stream
                     preface + b"\x00\x00\x03\x00\x00\x00\x00\x00\x00abc",                    # DATA on stream 0
                     preface + b"\x00\x00\x03\x00\x00\x00\x00\x00\x05abc",                    # DATA on an idle stream
                     preface + b"\x00\x40\x00"):                                               # truncated header
            s = pysock.socket(pysock.AF_UNIX, pysock.SOCK_STREAM)
            s.settimeout(5)
            s.connect(srv.socket)
            s.sendall(junk)
            s.shutdown(pysock.SHUT_WR)
            while s.recv(65536):                        # server says what it has to say, then closes
                pass
            s.close()
        with NativeClient(srv.socket) as cl:
            assert cl.call("GetDevicePluginOptions") == b""
        assert srv.stats()["connections"] >= 9
    finally:
        srv.Stop()


def test_stream_cap_is_enforced(golden_cases, sockdir):
    """SETTINGS_MAX_CONCURRENT_STREAMS (128) is advertised and enforced: the 129th open watch
    stream is refused with RST_STREAM, the others and the connection are untouched."""
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"))
    srv.Start()
    try:
        with NativeClient(srv.socket) as cl:
            healthy = table.payload_size()
            sids = [cl.open_stream("ListAndWatch") for _ in range(128)]
            for sid in sids:
                assert len(cl.next_message(sid)) == healthy
            extra = cl.open_stream("ListAndWatch")
            with pytest.raises(RpcError) as e:
                cl.next_message(extra)
            assert e.value.code == 13                      # reset by the peer: REFUSED_STREAM
            cl.cancel(sids[0])
            again = cl.open_stream("ListAndWatch")         # a slot is free again
            assert len(cl.next_message(again)) == healthy
            plug.unhealthy(2)
            for sid in sids[1:] + [again]:                 # every watcher gets the one re-send
                assert len(cl.next_message(sid)) == healthy + 2 * 16
    finally:
        srv.Stop()


def test_abi_argument_checks(sockdir):
    lib = L.load()
    import ctypes as C
    h = C.c_void_p()
    assert lib.gs_server_start(None, None, C.byref(h)) == L.GS_E_INVAL
    assert lib.gs_server_stop(None) == L.GS_E_INVAL
    assert lib.gs_grpc_client_connect(None, 10, C.byref(h)) == L.GS_E_INVAL
    assert lib.gs_grpc_client_connect(("x" * 200).encode(), 10, C.byref(h)) == L.GS_E_INVAL
    assert lib.gs_register(None, b"a", b"b", 10) == L.GS_E_INVAL
    lib.gs_grpc_client_close(None)
    table = api.DeviceTable(["GPU-a"], 4, api.GiB)
    plug = api.NvidiaDevicePlugin(table, "n")
    cfg = L.ServerConfig()
    cfg.socket_path = os.path.join(sockdir, "no", "such", "dir", "s.sock").encode()
    assert lib.gs_server_start(plug.h, C.byref(cfg), C.byref(h)) == L.GS_E_STATE


def _build_transport_test(tmp_path, name, kind):
    from tests.conftest import build_sanitized
    return build_sanitized(kind, os.path.join(ROOT, "tests", "cpp", "fuzz_transport.cpp"), tmp_path / name)


def test_transport_fuzz_under_address_and_ub_sanitizers(tmp_path):
    """4000 mutated client sessions against a live server and 1000 mutated server sessions
    against the client, ASan + UBSan, with well-formed calls checked in between."""
    exe = _build_transport_test(tmp_path, "fuzz_transport", "asan")
    out = subprocess.run([exe, "fuzz", "4000"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "TRANSPORT FUZZ OK" in out.stdout, (out.stdout + out.stderr)[-3000:]


def test_transport_under_thread_sanitizer(tmp_path):
    """Two Allocate callers, a ListAndWatch reader, a health marker and Stop at once: no race,
    and the watch stream ends with status OK."""
    exe = _build_transport_test(tmp_path, "race_transport", "tsan")
    out = subprocess.run([exe, "race", "2"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "TRANSPORT RACE OK" in out.stdout \
        and "WARNING: ThreadSanitizer" not in out.stderr, (out.stdout + out.stderr)[-4000:]


def _frame(ftype, flags, stream, payload=b""):
    return len(payload).to_bytes(3, "big") + bytes([ftype, flags]) + stream.to_bytes(4, "big") + payload


def _lit(name, value):                                   # HPACK literal without indexing, new name
    return b"\x00" + bytes([len(name)]) + name + bytes([len(value)]) + value


def _read_frames(sock):
    buf = b""
    while True:
        chunk = sock.recv(65536)
        if not chunk:
            break
        buf += chunk
    frames, i = [], 0
    while i + 9 <= len(buf):
        n = int.from_bytes(buf[i:i + 3], "big")
        frames.append((buf[i + 3], buf[i + 4], int.from_bytes(buf[i + 5:i + 9], "big"), buf[i + 9:i + 9 + n]))
        i += 9 + n
    return frames


def test_frames_grpc_core_never_sends(golden_cases, sockdir):
    """CONTINUATION, padded and prioritised HEADERS, padded DATA split over several frames, a
    PRIORITY frame, an unknown frame type, a PING: RFC 7540 says all of it is legal, and the
    call must be answered like any other."""
    import socket as pysock
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"))
    srv.Start()
    try:
        req = R.encode_allocate_request(c["steps"][0]["request"])
        body = b"\x00" + len(req).to_bytes(4, "big") + req
        block = (b"\x83\x86" + b"\x04" + bytes([len(b"/v1beta1.DevicePlugin/Allocate")]) + b"/v1beta1.DevicePlugin/Allocate"
                 + _lit(b"content-type", b"application/grpc") + _lit(b"te", b"trailers")
                 + _lit(b"x-filler", b"f" * 100))
        pad = b"\x00" * 7
        headers_payload = bytes([len(pad)]) + b"\x80\x00\x00\x00\x10" + block[:20] + pad    # PADDED | PRIORITY
        session = (b"PRI * HTTP/2.0\r\n\r\nSM\r\n\r\n" + _frame(4, 0, 0)
                   + _frame(2, 0, 1, b"\x00\x00\x00\x00\x05")                               # PRIORITY
                   + _frame(0x42, 0xff, 0, b"extension")                                    # unknown type
                   + _frame(6, 0, 0, b"pingpong")
                   + _frame(1, 0x08 | 0x20, 1, headers_payload)                             # no END_HEADERS
                   + _frame(9, 0, 1, block[20:60]) + _frame(9, 0x04, 1, block[60:])         # CONTINUATION x2
                   + _frame(0, 0x08, 1, b"\x03" + body[:10] + b"\x00" * 3)                  # padded DATA
                   + _frame(0, 0, 1, b"")                                                   # empty DATA
                   + _frame(0, 0x01, 1, body[10:]))                                         # END_STREAM
        s = pysock.socket(pysock.AF_UNIX, pysock.SOCK_STREAM)
        s.settimeout(5)
        s.connect(srv.socket)
        s.sendall(session)
        s.shutdown(pysock.SHUT_WR)
        frames = _read_frames(s)
        s.close()
        assert (6, 1, 0, b"pingpong") in frames                                             # PING ack
        data = b"".join(p for t, f, sid, p in frames if t == 0 and sid == 1)
        assert data[0] == 0 and int.from_bytes(data[1:5], "big") == len(data) - 5
        assert R.decode_allocate_response(data[5:]) == c["steps"][0]["expect_envs"]
        trailers = [p for t, f, sid, p in frames if t == 1 and sid == 1 and f & 0x01]
        assert len(trailers) == 1 and trailers[0].endswith(b"\x0bgrpc-status\x010")
        assert not any(t == 7 and int.from_bytes(p[4:8], "big") != 0 for t, f, sid, p in frames)   # no error GOAWAY
        # CONTINUATION interrupted by another frame is a connection error (RFC 7540 6.10)
        s = pysock.socket(pysock.AF_UNIX, pysock.SOCK_STREAM)
        s.settimeout(5)
        s.connect(srv.socket)
        s.sendall(b"PRI * HTTP/2.0\r\n\r\nSM\r\n\r\n" + _frame(4, 0, 0) + _frame(1, 0, 1, block[:20])
                  + _frame(6, 0, 0, b"12345678"))
        s.shutdown(pysock.SHUT_WR)
        frames = _read_frames(s)
        s.close()
        goaway = [p for t, f, sid, p in frames if t == 7]
        assert goaway and int.from_bytes(goaway[0][4:8], "big") == 1                        # PROTOCOL_ERROR
        # trailers on a stream whose request is already complete (an open watch stream): a stream
        # error, and the connection keeps serving
        law = b"\x83\x86" + b"\x04" + bytes([len(b"/v1beta1.DevicePlugin/ListAndWatch")]) + \
            b"/v1beta1.DevicePlugin/ListAndWatch" + _lit(b"content-type", b"application/grpc")
        opts = b"\x83\x86" + b"\x04" + bytes([len(b"/v1beta1.DevicePlugin/GetDevicePluginOptions")]) + \
            b"/v1beta1.DevicePlugin/GetDevicePluginOptions" + _lit(b"content-type", b"application/grpc")
        s = pysock.socket(pysock.AF_UNIX, pysock.SOCK_STREAM)
        s.settimeout(5)
        s.connect(srv.socket)
        s.sendall(b"PRI * HTTP/2.0\r\n\r\nSM\r\n\r\n" + _frame(4, 0, 0)
                  + _frame(1, 0x04, 1, law) + _frame(0, 0x01, 1, b"\x00\x00\x00\x00\x00")
                  + _frame(1, 0x04 | 0x01, 1, _lit(b"x-late", b"trailer"))
                  + _frame(1, 0x04, 3, opts) + _frame(0, 0x01, 3, b"\x00\x00\x00\x00\x00"))
        s.shutdown(pysock.SHUT_WR)
        frames = _read_frames(s)
        s.close()
        assert any(t == 3 and sid == 1 and int.from_bytes(p, "big") == 5 for t, f, sid, p in frames)   # STREAM_CLOSED
        assert any(t == 1 and sid == 3 and p.endswith(b"\x0bgrpc-status\x010") for t, f, sid, p in frames)
    finally:
        srv.Stop()


def test_kubelet_dials_back_inside_register(golden_cases, sockdir):
    """kubelet's Register handler connects to the plugin's socket (blocking dial,
    GetDevicePluginOptions) before it answers; the plugin must be serving while its own
    Register call is still outstanding (server.go:240-244 registers after Start for that reason)."""
    from concurrent import futures
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    plugin_sock = os.path.join(sockdir, "aliyungpushare.sock")
    seen = {}

    def register(request, context):
        ch = grpc.insecure_channel("unix://" + plugin_sock)
        grpc.channel_ready_future(ch).result(timeout=5)
        seen["options"] = KubeletStub.method(ch, "GetDevicePluginOptions")(b"", timeout=5)
        stream = KubeletStub.method(ch, "ListAndWatch", stream=True)(b"")
        seen["devices"] = len(R.decode_list_and_watch(next(stream)))
        stream.cancel()
        ch.close()
        seen["request"] = request
        return b""
    kubelet = grpc.server(futures.ThreadPoolExecutor(max_workers=2))
    kubelet.add_generic_rpc_handlers((grpc.method_handlers_generic_handler("v1beta1.Registration", {
        "Register": grpc.unary_unary_rpc_method_handler(register, lambda b: b, lambda b: b)}),))
    kubelet_sock = os.path.join(sockdir, "kubelet.sock")
    kubelet.add_insecure_port("unix://" + kubelet_sock)
    kubelet.start()
    srv = NativeDevicePluginServer(plug, plugin_sock)
    try:
        srv.Serve(kubelet_sock)
        assert seen["options"] == b"" and seen["devices"] == 128
        assert seen["request"] == encode_register_request("v1beta1", "aliyungpushare.sock", "aliyun.com/gpu-mem")
    finally:
        srv.Stop()
        kubelet.stop(grace=0)


def test_tiny_flow_control_windows(golden_cases, sockdir):
    """A peer that grants the list a thousand bytes at a time (SETTINGS_INITIAL_WINDOW_SIZE =
    1000, then WINDOW_UPDATEs on the stream; the connection window runs dry too and is topped up):
    the server never sends a byte more than it was granted, and the pieces add up to the list."""
    import socket as pysock
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"))
    srv.Start()
    try:
        want = table.list_and_watch_bytes()
        law = b"\x83\x86" + b"\x04" + bytes([len(b"/v1beta1.DevicePlugin/ListAndWatch")]) + \
            b"/v1beta1.DevicePlugin/ListAndWatch" + _lit(b"content-type", b"application/grpc")
        s = pysock.socket(pysock.AF_UNIX, pysock.SOCK_STREAM)
        s.settimeout(5)
        s.connect(srv.socket)
        s.sendall(b"PRI * HTTP/2.0\r\n\r\nSM\r\n\r\n" + _frame(4, 0, 0, b"\x00\x04" + (1000).to_bytes(4, "big"))
                  + _frame(1, 0x04, 1, law) + _frame(0, 0x01, 1, b"\x00\x00\x00\x00\x00"))
        buf, data, granted_stream, granted_conn = b"", b"", 1000, 65535

        def pump(until):
            nonlocal buf, data
            while not until():
                chunk = s.recv(65536)
                assert chunk, "server closed early"
                buf += chunk
                while len(buf) >= 9 and len(buf) >= 9 + int.from_bytes(buf[:3], "big"):
                    n = int.from_bytes(buf[:3], "big")
                    ftype, sid, payload = buf[3], int.from_bytes(buf[5:9], "big"), buf[9:9 + n]
                    buf = buf[9 + n:]
                    if ftype == 0 and sid == 1:
                        data += payload
        rounds = 0
        while len(data) < 5 + len(want):
            pump(lambda: len(data) >= min(granted_stream, granted_conn, 5 + len(want)))
            assert len(data) <= min(granted_stream, granted_conn), "sent beyond the window"
            time.sleep(0.02)
            s.settimeout(0.05)
            try:
                extra = s.recv(65536)                      # nothing may arrive while the window is shut
                buf += extra
                pump(lambda: True)
            except (TimeoutError, pysock.timeout):
                pass
            s.settimeout(5)
            assert len(data) <= min(granted_stream, granted_conn)
            s.sendall(_frame(8, 0, 1, (700).to_bytes(4, "big")))
            granted_stream += 700
            rounds += 1
        assert data[0] == 0 and int.from_bytes(data[1:5], "big") == len(want) and data[5:] == want
        assert rounds >= 9
        s.close()
    finally:
        srv.Stop()


def test_connection_cap(golden_cases, sockdir):
    """At most 64 connections are held (kubelet opens one): further ones are closed at once,
    and a slot frees up when a client leaves."""
    c = _golden(golden_cases)
    table, plug = product_for_case(c)
    srv = NativeDevicePluginServer(plug, os.path.join(sockdir, "p.sock"))
    srv.Start()
    try:
        clients = [NativeClient(srv.socket) for _ in range(64)]
        assert all(cl.call("GetDevicePluginOptions") == b"" for cl in clients)
        with pytest.raises(L.GpushareError):
            NativeClient(srv.socket, timeout_ms=500)
        clients.pop().close()
        time.sleep(0.3)
        late = NativeClient(srv.socket)
        assert late.call("GetDevicePluginOptions") == b""
        late.close()
        for cl in clients:
            cl.close()
    finally:
        srv.Stop()

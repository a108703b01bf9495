"""The product's wire bytes against an independent protobuf implementation.

Everything else in the suite decodes the product's AllocateResponse / ListAndWatchResponse with
the oracle's own reader (oracle/ref_model.py:_fields), i.e. codec and checker share an author.
Here the deviceplugin v1beta1 schema (k8s.io/kubelet v0.28.3 pkg/apis/deviceplugin/v1beta1/
api.proto, the package the reference imports at server.go:15 and fills at server.go:165-181 and
allocate.go:116-124) is declared to Google's protobuf runtime through descriptor_pb2, and that
runtime parses what the product wrote and serialises what the product must read.

The schema is restated from the published api.proto (not on disk here: the reference's go.mod
names the module, go.mod:14, and vendors nothing); field numbers and types are the ones the
reference's calls depend on.
"""
import os
import tempfile
from concurrent import futures

import pytest
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory, unknown_fields

from gpushare_b200 import api
from oracle import ref_model as R
from tests.helpers import product_for_case, product_step

F = descriptor_pb2.FieldDescriptorProto
STR, BOOL, I32, I64, MSG = F.TYPE_STRING, F.TYPE_BOOL, F.TYPE_INT32, F.TYPE_INT64, F.TYPE_MESSAGE


def _schema():
    fd = descriptor_pb2.FileDescriptorProto(name="deviceplugin_v1beta1_api.proto", package="v1beta1",
                                            syntax="proto3")

    def msg(name, *fields):
        m = fd.message_type.add(name=name)
        for fname, number, ftype, repeated, type_name in fields:
            f = m.field.add(name=fname, number=number, type=ftype,
                            label=F.LABEL_REPEATED if repeated else F.LABEL_OPTIONAL)
            if type_name:
                f.type_name = ".v1beta1." + type_name
        return m

    def str_map(m, fname, number):
        entry = m.nested_type.add(name="".join(p.capitalize() for p in fname.split("_")) + "Entry")
        entry.options.map_entry = True
        entry.field.add(name="key", number=1, type=STR, label=F.LABEL_OPTIONAL)
        entry.field.add(name="value", number=2, type=STR, label=F.LABEL_OPTIONAL)
        m.field.add(name=fname, number=number, type=MSG, label=F.LABEL_REPEATED,
                    type_name=".v1beta1.%s.%s" % (m.name, entry.name))

    msg("DevicePluginOptions", ("pre_start_required", 1, BOOL, False, None),
        ("get_preferred_allocation_available", 2, BOOL, False, None))
    msg("RegisterRequest", ("version", 1, STR, False, None), ("endpoint", 2, STR, False, None),
        ("resource_name", 3, STR, False, None), ("options", 4, MSG, False, "DevicePluginOptions"))
    msg("Empty")
    msg("NUMANode", ("ID", 1, I64, False, None))
    msg("TopologyInfo", ("nodes", 1, MSG, True, "NUMANode"))
    msg("Device", ("ID", 1, STR, False, None), ("health", 2, STR, False, None),
        ("topology", 3, MSG, False, "TopologyInfo"))
    msg("ListAndWatchResponse", ("devices", 1, MSG, True, "Device"))
    msg("PreStartContainerRequest", ("devices_ids", 1, STR, True, None))
    msg("PreStartContainerResponse")
    msg("ContainerAllocateRequest", ("devices_ids", 1, STR, True, None))
    msg("AllocateRequest", ("container_requests", 1, MSG, True, "ContainerAllocateRequest"))
    msg("Mount", ("container_path", 1, STR, False, None), ("host_path", 2, STR, False, None),
        ("read_only", 3, BOOL, False, None))
    msg("DeviceSpec", ("container_path", 1, STR, False, None), ("host_path", 2, STR, False, None),
        ("permissions", 3, STR, False, None))
    msg("CDIDevice", ("name", 1, STR, False, None))
    car = msg("ContainerAllocateResponse")
    str_map(car, "envs", 1)
    car.field.add(name="mounts", number=2, type=MSG, label=F.LABEL_REPEATED, type_name=".v1beta1.Mount")
    car.field.add(name="devices", number=3, type=MSG, label=F.LABEL_REPEATED, type_name=".v1beta1.DeviceSpec")
    str_map(car, "annotations", 4)
    car.field.add(name="cdi_devices", number=5, type=MSG, label=F.LABEL_REPEATED, type_name=".v1beta1.CDIDevice")
    msg("AllocateResponse", ("container_responses", 1, MSG, True, "ContainerAllocateResponse"))
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return {m.name: message_factory.GetMessageClass(pool.FindMessageTypeByName("v1beta1." + m.name))
            for m in fd.message_type}


PB = _schema()


def pb_envs(resp_bytes):
    """AllocateResponse bytes -> [env dict per container], by Google's parser; asserts that only
    `envs` is populated (allocate.go:116-124 sets nothing else)."""
    r = PB["AllocateResponse"]()
    r.ParseFromString(resp_bytes)
    assert len(unknown_fields.UnknownFieldSet(r)) == 0
    out = []
    for c in r.container_responses:
        assert len(unknown_fields.UnknownFieldSet(c)) == 0
        assert not c.mounts and not c.devices and not c.annotations and not c.cdi_devices
        out.append(dict(c.envs))
    return out


def test_allocate_response_parses_with_google_protobuf(golden_cases):
    """Every step of every golden case: the product's response, parsed by the protobuf runtime,
    equals the oracle's decoded view of the same bytes and the case's expected envs, and
    re-serialising the parsed message gives back a message that parses to the same thing."""
    steps = 0
    for c in golden_cases:
        table, plug = product_for_case(c)
        for st in c["steps"]:
            d_envs, _kind, _patched = product_step(c, table, plug, st["request"])
            assert d_envs == st["expect_envs"]
        # the same walk again, this time keeping the raw bytes
        table, plug = product_for_case(c)
        for st in c["steps"]:
            wire = R.encode_allocate_request(st["request"])
            if c["fail_list"] or c["fail_patch"] == "always":
                d = api.allocate_error_response(table, c["node"], wire, api.L.GS_ALLOC_ERR_PODS)
            else:
                d = plug.Allocate(wire)
            envs = pb_envs(d.response)
            assert envs == R.decode_allocate_response(d.response)
            if not (c["fail_list"] or c["fail_patch"] == "always"):
                assert envs == st["expect_envs"], c["name"]
            again = PB["AllocateResponse"]()
            again.ParseFromString(d.response)
            assert pb_envs(again.SerializeToString()) == envs
            steps += 1
    assert steps >= 47


def test_allocate_request_from_google_protobuf_is_scanned_exactly():
    """AllocateRequest built and serialised by the protobuf runtime (what kubelet's gogo
    marshaller produces for this shape is the same canonical encoding): gs_request_scan must
    count exactly the ids per container, for empty containers, long ids and many containers."""
    import random
    rng = random.Random(7)
    for trial in range(60):
        req = PB["AllocateRequest"]()
        shape = []
        for _ in range(rng.randint(0, 9)):
            c = req.container_requests.add()
            n = rng.choice([0, 1, 2, 17, 300, 5000]) if trial % 3 else rng.randint(0, 40)
            for j in range(n):
                c.devices_ids.append("GPU-%08x-aaaa-bbbb-cccc-%012x-_-%d" % (trial, j, rng.randint(0, 183358))
                                     if rng.random() < 0.97 else "x" * rng.choice([0, 1, 127, 128, 300, 20000]))
            shape.append(n)
        wire = req.SerializeToString()
        per, total = api.request_scan(wire)
        assert (total, list(per)) == (sum(shape), shape)
        assert R.decode_allocate_request(wire) == [list(c.devices_ids) for c in req.container_requests]
        # and the oracle's encoder writes the very same bytes as the runtime
        assert R.encode_allocate_request([list(c.devices_ids) for c in req.container_requests]) == wire


@pytest.mark.parametrize("unit,raw,n_gpus", [("GiB", 16384, 8), ("MiB", 2048, 3), ("GiB", 183359, 1)])
def test_list_and_watch_parses_with_google_protobuf(unit, raw, n_gpus):
    """ListAndWatchResponse bytes in all three health states (server.go:180-197): all healthy,
    mixed (one fake device, then one whole GPU), all unhealthy -- parsed by the protobuf runtime,
    equal to the oracle's device list, ids in (gpu asc, j asc) order (nvidia.go:72-80)."""
    uuids = ["GPU-%08x-0000-0000-0000-%012x" % (i, i) for i in range(n_gpus)]
    table = api.DeviceTable.from_raw(uuids, raw, unit)
    g = R.Globals(unit)
    oracle = R.NvidiaDevicePlugin(g, [(u, raw) for u in uuids], "n", R.PodSource([]))

    def check():
        wire = table.list_and_watch_bytes()
        r = PB["ListAndWatchResponse"]()
        r.ParseFromString(wire)
        got = [(d.ID, d.health) for d in r.devices]
        assert all(not d.HasField("topology") for d in r.devices)
        assert got == oracle.ListAndWatchResponse()
        assert got == R.decode_list_and_watch(wire)
        assert len(r.SerializeToString()) == len(wire)          # nothing but canonical fields
        return got
    got = check()
    assert len(got) == n_gpus * table.units_per_gpu and all(h == "Healthy" for _id, h in got)
    victim = oracle.devs[min(3, len(oracle.devs) - 1)]
    table.mark_unhealthy_id(victim[0])
    victim[1] = "Unhealthy"
    got = check()
    assert sum(h == "Unhealthy" for _id, h in got) == 1
    last = n_gpus - 1
    table.mark_unhealthy_gpu(last)
    for d in oracle.devs:
        if d[0].startswith(uuids[last] + "-_-"):
            d[1] = "Unhealthy"
    got = check()
    for gpu in range(n_gpus):
        table.mark_unhealthy_gpu(gpu)
    for d in oracle.devs:
        d[1] = "Unhealthy"
    got = check()
    assert all(h == "Unhealthy" for _id, h in got)


def test_options_and_prestart_parse_with_google_protobuf():
    table = api.DeviceTable.from_raw(["GPU-a"], 16384, "GiB")
    plug = api.NvidiaDevicePlugin(table, "n")
    o = PB["DevicePluginOptions"]()
    o.ParseFromString(plug.GetDevicePluginOptions())
    assert (o.pre_start_required, o.get_preferred_allocation_available) == (False, False)   # server.go:93-95
    p = PB["PreStartContainerResponse"]()
    req = PB["PreStartContainerRequest"](devices_ids=["GPU-a-_-0"])
    p.ParseFromString(plug.PreStartContainer(req.SerializeToString()))
    assert p.SerializeToString() == b""                                                      # server.go:199-201


def test_register_request_parses_with_google_protobuf():
    """gs_register (server.go:158-177) against a grpcio stand-in for kubelet whose handler parses
    the request with the protobuf runtime: version, endpoint, resource name, options unset."""
    import grpc

    from gpushare_b200 import _lib as L
    lib = L.load()
    seen = []

    def register(req, _ctx):
        seen.append(req)
        return PB["Empty"]()

    handler = grpc.method_handlers_generic_handler("v1beta1.Registration", {
        "Register": grpc.unary_unary_rpc_method_handler(
            register, request_deserializer=PB["RegisterRequest"].FromString,
            response_serializer=lambda m: m.SerializeToString())})
    d = tempfile.mkdtemp(prefix="gsreg")
    sock = os.path.join(d, "kubelet.sock")
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=2))
    server.add_generic_rpc_handlers((handler,))
    server.add_insecure_port("unix://" + sock)
    server.start()
    try:
        rc = lib.gs_register(sock.encode(), b"aliyungpushare.sock", b"aliyun.com/gpu-mem", 5000)
        assert rc == L.GS_OK
        assert len(seen) == 1
        r = seen[0]
        assert (r.version, r.endpoint, r.resource_name) == ("v1beta1", "aliyungpushare.sock", "aliyun.com/gpu-mem")
        assert not r.HasField("options")                                                    # server.go:166-170
        want = PB["RegisterRequest"](version="v1beta1", endpoint="aliyungpushare.sock",
                                     resource_name="aliyun.com/gpu-mem").SerializeToString()
        assert r.SerializeToString() == want
    finally:
        server.stop(0)
        try:
            os.unlink(sock)
        except OSError:
            pass
        os.rmdir(d)


def test_native_server_answers_a_protobuf_speaking_grpcio_client(golden_cases):
    """The whole boundary with generated-style stubs on the kubelet side: a grpcio channel whose
    (de)serialisers are the protobuf runtime's classes talks to the library's own server on a unix
    socket: options, one ListAndWatch message, Allocate for G1 -- decoded envs as the case expects."""
    import grpc

    from gpushare_b200.native_transport import NativeDevicePluginServer
    case = next(c for c in golden_cases if c["name"] == "G1-binpack-1")
    table, plug = product_for_case(case)
    d = tempfile.mkdtemp(prefix="gspb")
    srv = NativeDevicePluginServer(plug, os.path.join(d, "aliyungpushare.sock"))
    srv.Start()
    try:
        ch = grpc.insecure_channel("unix://" + srv.socket, options=[("grpc.max_receive_message_length", -1)])
        ser = lambda m: m.SerializeToString()                                               # noqa: E731
        opts = ch.unary_unary("/v1beta1.DevicePlugin/GetDevicePluginOptions", request_serializer=ser,
                              response_deserializer=PB["DevicePluginOptions"].FromString)(PB["Empty"](), timeout=10)
        assert not opts.pre_start_required
        stream = ch.unary_stream("/v1beta1.DevicePlugin/ListAndWatch", request_serializer=ser,
                                 response_deserializer=PB["ListAndWatchResponse"].FromString)(PB["Empty"]())
        first = next(iter(stream))
        assert len(first.devices) == 8 * 16 and first.devices[17].ID == case["gpus"][1][0] + "-_-1"
        assert {dv.health for dv in first.devices} == {"Healthy"}
        req = PB["AllocateRequest"]()
        for ids in case["steps"][0]["request"]:
            req.container_requests.add().devices_ids.extend(ids)
        resp = ch.unary_unary("/v1beta1.DevicePlugin/Allocate", request_serializer=ser,
                              response_deserializer=PB["AllocateResponse"].FromString)(req, timeout=10)
        assert [dict(c.envs) for c in resp.container_responses] == case["steps"][0]["expect_envs"]
        stream.cancel()
        ch.close()
    finally:
        srv.Stop()
        try:
            os.rmdir(d)
        except OSError:
            pass

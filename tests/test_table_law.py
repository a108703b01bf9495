"""Device model and ListAndWatch payload against the oracle (A1-A5)."""
import ctypes as C
import random

import pytest

from gpushare_b200 import api
from oracle import ref_model as R


def uuid(i):
    return "GPU-%08x-0000-0000-0000-%012x" % (i, i)


def oracle_devs(n, raw, unit):
    g = R.Globals(unit)
    return R.getDevices(g, [(uuid(i), raw) for i in range(n)])


@pytest.mark.parametrize("n,raw,unit", [(8, 16384, "GiB"), (1, 183359, "GiB"), (2, 1100, "MiB"),
                                        (3, 1023, "GiB"), (0, 0, "GiB"), (8, 183359, "GiB")])
def test_ids_and_payload_match_oracle(n, raw, unit):
    devs, names = oracle_devs(n, raw, unit)
    t = api.DeviceTable.from_raw([uuid(i) for i in range(n)], raw, unit)
    assert t.gpu_count == len(names) == n
    assert t.device_count == len(devs)
    wire = t.list_and_watch_bytes()
    assert len(wire) == t.payload_size()
    assert wire == R.encode_list_and_watch(devs)            # same order, same bytes
    assert R.decode_list_and_watch(wire) == [tuple(d) for d in devs]
    per = t.units_per_gpu
    for k in range(0, len(devs), max(1, len(devs) // 50)):
        g, j = divmod(k, per)
        assert t.fake_id(g, j) == devs[k][0]
        assert t.lookup(devs[k][0]) == (g, j)
        assert t.extract_real_device_id(devs[k][0]) == R.extractRealDeviceID(devs[k][0]) == uuid(g)


def test_payload_sizes_from_survey():
    """13 + len(ID) bytes per healthy device."""
    t = api.DeviceTable.from_raw([uuid(i) for i in range(8)], 16384, "GiB")
    assert t.payload_size() == sum(13 + len("%s-_-%d" % (uuid(g), j)) for g in range(8) for j in range(16))
    t = api.DeviceTable.from_raw([uuid(0)], 183359, "MiB")
    assert t.device_count == 183359
    want = sum(13 + 43 + len(str(j)) for j in range(183359))
    assert t.payload_size() == want and len(t.list_and_watch_bytes()) == want
    assert 61.3 < want / 183359 < 61.5                      # 61.39 B/device at MiB on B200


def test_mib_full_node_payload_is_byte_exact():
    """BASELINE config 3 shape on 2 GPUs (the 8-GPU payload is 90 MB; same code path)."""
    devs, _ = oracle_devs(2, 183359, "MiB")
    t = api.DeviceTable.from_raw([uuid(0), uuid(1)], 183359, "MiB")
    assert t.list_and_watch_bytes() == R.encode_list_and_watch(devs)


def test_segments_concatenate_to_payload():
    t = api.DeviceTable.from_raw([uuid(i) for i in range(4)], 16384, "GiB")
    segs = t.payload_segments()
    assert len(segs) == 4
    assert b"".join(C.string_at(a, n) for a, n in segs) == t.list_and_watch_bytes()
    t.mark_unhealthy_gpu(2)
    t.mark_unhealthy_device(0, 5)
    segs = t.payload_segments()
    assert b"".join(C.string_at(a, n) for a, n in segs) == t.list_and_watch_bytes()


def test_lookup_rejects_foreign_ids():
    t = api.DeviceTable.from_raw([uuid(0), uuid(1)], 16384, "GiB")
    assert t.lookup(uuid(1) + "-_-15") == (1, 15)
    for bad in [uuid(1) + "-_-16", uuid(2) + "-_-0", uuid(1) + "-_-", uuid(1) + "-_-01",
                uuid(1) + "_-_1", uuid(1), "", "-_-3", uuid(1) + "-_-+1", uuid(1) + "-_-1 ",
                uuid(1) + "-_-99999999999999999999999"]:
        assert t.lookup(bad) is None, bad


def test_extract_real_device_id_is_first_separator():
    t = api.DeviceTable(["GPU-a"], 1)
    assert t.extract_real_device_id("GPU-a-_-1-_-2") == "GPU-a"
    assert t.extract_real_device_id("no-separator") == "no-separator"
    assert t.extract_real_device_id("") == ""
    assert R.extractRealDeviceID("GPU-a-_-1-_-2") == "GPU-a"


def test_device_name_by_index():
    t = api.DeviceTable.from_raw([uuid(i) for i in range(3)], 4096, "GiB")
    assert t.device_name_by_index(2) == (uuid(2), True)
    assert t.device_name_by_index(3) == ("", False)
    assert t.device_name_by_index(2**63) == ("", False)
    assert t.device_index_by_name(uuid(1)) == 1 and t.device_index_by_name("GPU-x") is None


def test_duplicate_uuid_rejected():
    with pytest.raises(api.L.GpushareError):
        api.DeviceTable(["GPU-a", "GPU-a"], 4)


def test_health_matches_oracle_rule():
    """watchXIDs rule (nvidia.go:134-150) + ListAndWatch resend (server.go:187-190): final
    device states equal the oracle's; the number of sends differs by design (H13)."""
    n, raw = 4, 8192
    g = R.Globals("GiB")
    src = R.PodSource([])
    oracle = R.NvidiaDevicePlugin(g, [(uuid(i), raw) for i in range(n)], "n", src)
    t = api.DeviceTable.from_raw([uuid(i) for i in range(n)], raw, "GiB")
    plug = api.NvidiaDevicePlugin(t, "n")
    gen0 = t.generation
    for ev_gpu, xid in [(1, 31), (3, 79), (1, 43), (None, 45)]:
        sends = oracle.unhealthy_event(None if ev_gpu is None else uuid(ev_gpu), xid)
        if xid not in (31, 43, 45):
            plug.unhealthy(-1 if ev_gpu is None else ev_gpu)
            assert sends == (n * 8 if ev_gpu is None else 8)
        assert R.decode_list_and_watch(t.list_and_watch_bytes()) == oracle.ListAndWatchResponse()
    assert t.generation == gen0 + 1                           # one coalesced re-advertise
    assert not t.device_healthy(3, 0) and t.device_healthy(2, 7)
    sends = oracle.unhealthy_event("", 62)                    # no UUID => every device
    plug.unhealthy(-1)
    assert sends == 32
    assert R.decode_list_and_watch(t.list_and_watch_bytes()) == oracle.ListAndWatchResponse()
    assert len(t.list_and_watch_bytes()) == t.payload_size()


def test_single_fake_device_unhealthy_like_reference():
    """server.go:189 marks ONE fake device; mixed segments still encode exactly."""
    rng = random.Random(3)
    devs, _ = oracle_devs(2, 300, "MiB")
    t = api.DeviceTable.from_raw([uuid(0), uuid(1)], 300, "MiB")
    plug = api.NvidiaDevicePlugin(t, "n")
    for k in rng.sample(range(len(devs)), 40):
        devs[k][1] = "Unhealthy"
        plug.unhealthy_device(devs[k][0])
    wire = t.list_and_watch_bytes()
    assert wire == R.encode_list_and_watch(devs) and len(wire) == t.payload_size()
    t.reset_health()
    assert all(h == "Healthy" for _, h in R.decode_list_and_watch(t.list_and_watch_bytes()))


def test_list_and_watch_stream_and_stop():
    import threading
    t = api.DeviceTable.from_raw([uuid(0), uuid(1)], 2048, "GiB")
    plug = api.NvidiaDevicePlugin(t, "n")
    got = []

    def consume():
        for msg in plug.ListAndWatch(poll_ms=50):
            got.append(R.decode_list_and_watch(msg))
    th = threading.Thread(target=consume)
    th.start()
    import time
    time.sleep(0.1)
    plug.unhealthy(1)
    time.sleep(0.2)
    plug.Stop()
    th.join(2)
    assert not th.is_alive() and plug.stopped
    assert len(got) == 2
    assert [h for _, h in got[0]] == ["Healthy"] * 4
    assert [h for _, h in got[1]] == ["Healthy", "Healthy", "Unhealthy", "Unhealthy"]


def test_empty_rpcs_and_preferred_allocation():
    t = api.DeviceTable(["GPU-a"], 1)
    plug = api.NvidiaDevicePlugin(t, "n")
    assert plug.GetDevicePluginOptions() == b"" and plug.PreStartContainer(b"\x0a\x01x") == b""
    with pytest.raises(NotImplementedError):
        plug.GetPreferredAllocation()


def test_translate_memory_units():
    assert api.translate_memory_units("MiB") == ("MiB", False)
    assert api.translate_memory_units("GiB") == ("GiB", False)
    assert api.translate_memory_units("Gi") == ("GiB", True)
    assert R.translatememoryUnits("Gi") == ("GiB", True)


@pytest.mark.parametrize("s", ["0", "-0", "+3", "007", "-12", "9223372036854775807",
                               "-9223372036854775808", "", "+", " 1", "1_0", "0x1",
                               "9223372036854775808", "-9223372036854775809", "1.0"])
def test_c_atoi_matches_oracle(s):
    try:
        want = R.go_atoi(s)
    except ValueError:
        with pytest.raises(api.L.GpushareError):
            api.go_atoi(s)
    else:
        assert api.go_atoi(s) == want


@pytest.mark.parametrize("s", ["2", "2000m", "1500m", "1m", "0", "1Ki", "1.5Gi", "2e3", "2E3",
                               "1E", "100n", "-1500m", "0.5", "12e-1", ".5", "5.", "", "abc",
                               "1K", "1ki", "1e", "e3", "--1", "1Gii", "9223372036854775807",
                               "9223372036854775808", "8Ei", "7Ei", "1e18", "1e19", "0.000000001",
                               "123456789012345678901234567890m", ".", "+", "-", "-.", ".m", "+Ki",
                               "1.G", "1.5.2", "1e+", "++1", "e", "E", "Ei"])
def test_c_quantity_matches_oracle(s):
    try:
        want = R.quantity_value(s)
    except ValueError:
        with pytest.raises(api.L.GpushareError):
            api.quantity_value(s)
    else:
        assert api.quantity_value(s) == want


def test_c_parse_uint_matches_oracle():
    for s in ["0", "18446744073709551615", "18446744073709551616", "", "+1", "-1", "1e3", "00012"]:
        try:
            want = R.go_parse_uint(s)
        except ValueError:
            with pytest.raises(api.L.GpushareError):
                api.go_parse_uint(s)
        else:
            assert api.go_parse_uint(s) == want


def test_long_uuids_cross_the_one_byte_varint_boundary():
    """Ids longer than 114 characters make the Device message longer than 127 bytes: two-byte
    length prefixes in the payload, one-byte ones no longer valid. Still byte-exact."""
    uuids = ["GPU-" + "x" * 91, "GPU-" + "y" * 60]
    g = R.Globals("GiB")
    devs, _ = R.getDevices(g, [(u, 12288) for u in uuids])
    t = api.DeviceTable.from_raw(uuids, 12288, "GiB")
    assert t.list_and_watch_bytes() == R.encode_list_and_watch(devs)
    t.mark_unhealthy_gpu(0)
    t.mark_unhealthy_device(1, 11)
    for d in devs[:12] + [devs[23]]:
        d[1] = "Unhealthy"
    wire = t.list_and_watch_bytes()
    assert wire == R.encode_list_and_watch(devs) and len(wire) == t.payload_size()
    assert t.lookup(uuids[0] + "-_-11") == (0, 11)
    import pytest
    with pytest.raises(api.L.GpushareError):
        api.DeviceTable(["G" * 96], 1)

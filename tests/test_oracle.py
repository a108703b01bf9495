"""The oracle against the hand-derived golden vectors, and its Go-semantics helpers."""
import random

import pytest

from oracle import ref_model as R
from tests.helpers import oracle_for_case


def test_oracle_matches_golden(golden_cases):
    assert len(golden_cases) >= 40
    for c in golden_cases:
        src, plug = oracle_for_case(c)
        for k, st in enumerate(c["steps"]):
            before = len(src.patches)
            envs = plug.Allocate(st["request"])
            assert envs == st["expect_envs"], (c["name"], k)
            if st["expect_patch"]:
                ns, name, body = src.patches[before]
                assert (ns, name) == (st["expect_patch"]["namespace"], st["expect_patch"]["name"])
                assert body == st["expect_patch"]["body"].encode()
            else:
                assert len(src.patches) == before, (c["name"], k)


def test_pattern_kat(oracle_lib, pattern_kat):
    import ctypes
    for v in pattern_kat:
        w = (ctypes.c_uint32 * 4)()
        oracle_lib.gs_oracle_pattern_word(ctypes.c_uint64(v["i"]), ctypes.c_uint32(v["seed"]), w)
        assert list(w) == v["words"]
    # an independent restatement of the written definition (csrc/probe.cu):
    # background(seed) = r0..r3, word(i) = {r0 ^ lo(i), r1 ^ hi(i), r2, r3}
    M = 0xFFFFFFFF

    def background(seed):
        x = (seed * 0x9E3779B1 + 0x7F4A7C15) & M
        x ^= x >> 15
        x = (x * 0x2C1B3C6D) & M
        x ^= x >> 12
        r1 = (x * 0x85EBCA77 + 1) & M
        r2 = (r1 * 0xC2B2AE3D + 1) & M
        return x, r1, r2, (r2 * 0x27D4EB2F + 1) & M
    assert background(0) == (0x98E34241, 0xFA961638, 0x9BCD5B59, 0x52427858)       # by hand
    for v in pattern_kat:
        r = background(v["seed"])
        assert v["words"] == [r[0] ^ (v["i"] & M), r[1] ^ (v["i"] >> 32), r[2], r[3]]
    assert pattern_kat[0]["words"] == [0x98E34241, 0xFA961638, 0x9BCD5B59, 0x52427858]   # i = 0, seed = 0
    # every word of a sweep carries its own address; a different seed changes every lane
    assert len({tuple(v["words"]) for v in pattern_kat}) == len(pattern_kat)


def test_pattern_kat_hash(oracle_lib, pattern_kat_hash):
    """GS_PROBE_PATTERN_HASH: the oracle against the vectors, and both against an independent
    restatement of the written definition (csrc/probe_kernels.cuh)."""
    import ctypes
    M = 0xFFFFFFFF

    def word(i, seed):
        lo, hi = i & M, i >> 32
        h = (lo * 0x9E3779B1 + ((hi ^ seed) * 0x85EBCA77) + seed) & M
        h ^= h >> 15
        h = (h * 0x2C1B3C6D) & M
        h ^= h >> 12
        rot = ((h << 13) | (h >> 19)) & M
        return [h, h ^ M, (h * 0x297A2D39 + lo) & M, rot ^ hi ^ 0xC2B2AE3D]
    for v in pattern_kat_hash:
        w = (ctypes.c_uint32 * 4)()
        assert oracle_lib.gs_oracle_pattern_word_p(ctypes.c_uint64(v["i"]), ctypes.c_uint32(v["seed"]),
                                                   ctypes.c_uint32(1), w) == 0
        assert list(w) == v["words"] == word(v["i"], v["seed"])
        assert v["words"][0] ^ v["words"][1] == M                   # both polarities in one word
    assert pattern_kat_hash[0]["words"] == [0, M, 0, 0xC2B2AE3D]     # i = 0, seed = 0: by hand
    w = (ctypes.c_uint32 * 4)(1, 2, 3, 4)
    assert oracle_lib.gs_oracle_pattern_word_p(ctypes.c_uint64(0), ctypes.c_uint32(0), ctypes.c_uint32(2), w) == -1
    assert list(w) == [1, 2, 3, 4]                                  # unknown pattern: untouched


def test_oracle_pattern_check_counts(oracle_lib):
    import ctypes
    n = 1024
    buf = (ctypes.c_uint32 * (4 * n))()
    oracle_lib.gs_oracle_pattern_fill(buf, ctypes.c_uint64(10), ctypes.c_uint64(n), ctypes.c_uint32(3))
    buf[4 * 100 + 2] ^= 0x00FF0000
    buf[4 * 7 + 1] ^= 0x1
    mm, fb, fl = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
    oracle_lib.gs_oracle_pattern_check(buf, ctypes.c_uint64(10), ctypes.c_uint64(n), ctypes.c_uint32(3),
                                       ctypes.byref(mm), ctypes.byref(fb), ctypes.byref(fl))
    assert (mm.value, fb.value, fl.value) == (2, (10 + 7) * 16 + 4, 9)


@pytest.mark.parametrize("s,want", [("0", 0), ("-0", 0), ("+3", 3), ("007", 7), ("-12", -12),
                                    ("9223372036854775807", 2**63 - 1),
                                    ("-9223372036854775808", -2**63)])
def test_go_atoi_ok(s, want):
    assert R.go_atoi(s) == want


@pytest.mark.parametrize("s", ["", "+", "-", " 1", "1 ", "1_0", "0x1", "1.0", "٣",
                               "9223372036854775808", "-9223372036854775809", "++1"])
def test_go_atoi_err(s):
    with pytest.raises(ValueError):
        R.go_atoi(s)


def test_go_parse_uint():
    assert R.go_parse_uint("18446744073709551615") == 2**64 - 1
    for bad in ["", "+1", "-1", "18446744073709551616", "1e3", " 1"]:
        with pytest.raises(ValueError):
            R.go_parse_uint(bad)


@pytest.mark.parametrize("s,want", [("2", 2), (2, 2), ("2000m", 2), ("1500m", 2), ("1m", 1),
                                    ("0", 0), ("1Ki", 1024), ("1.5Gi", 1610612736),
                                    ("2e3", 2000), ("2E3", 2000), ("1E", 10**18),
                                    ("100n", 1), ("-1500m", -2), ("0.5", 1), ("12e-1", 2),
                                    (".5", 1), ("5.", 5), (".", 0), ("+", 0), ("-", 0), ("E", 0),
                                    ("Ei", 0), ("1.G", 10**9)])
def test_quantity_value(s, want):
    assert R.quantity_value(s) == want


@pytest.mark.parametrize("s", ["", "abc", "1K", "1ki", "1 Gi", "1e", "--1", "1Gii", "1.5.2", "1e+"])
def test_quantity_value_err(s):
    with pytest.raises(ValueError):
        R.quantity_value(s)


def test_go_sort_small_is_insertion_with_reversed_ties():
    keys = [5, 3, 5, 3, 1, 5]
    idx = list(range(len(keys)))
    R.go_sort_less_swap(len(idx), lambda i, j: keys[idx[i]] <= keys[idx[j]],
                        lambda i, j: idx.__setitem__(slice(None), _swapped(idx, i, j)))
    assert [(keys[k], k) for k in idx] == [(1, 4), (3, 3), (3, 1), (5, 5), (5, 2), (5, 0)]


def _swapped(lst, i, j):
    out = list(lst)
    out[i], out[j] = out[j], out[i]
    return out


def test_go_sort_large_distinct_sorts():
    rng = random.Random(20261106)
    for n in (13, 14, 50, 64, 200, 1000):
        for shape in ("random", "ascending", "descending", "organ"):
            keys = rng.sample(range(10 * n), n)
            if shape == "ascending":
                keys.sort()
            elif shape == "descending":
                keys.sort(reverse=True)
            elif shape == "organ":
                keys.sort()
                keys = keys[::2] + keys[1::2][::-1]
            a = list(keys)

            def swap(i, j):
                a[i], a[j] = a[j], a[i]
            R.go_sort_less_swap(n, lambda i, j: a[i] <= a[j], swap)
            assert a == sorted(keys), (n, shape)


def test_wire_roundtrip():
    reqs = [["a-_-1", "b-_-22"], [], ["c" * 200]]
    assert R.decode_allocate_request(R.encode_allocate_request(reqs)) == reqs
    devs = [("GPU-x-_-0", "Healthy"), ("GPU-x-_-1", "Unhealthy")]
    assert R.decode_list_and_watch(R.encode_list_and_watch(devs)) == devs


def test_get_devices_order_and_h7():
    g = R.Globals("GiB")
    devs, names = R.getDevices(g, [("GPU-a", 16500), ("GPU-b", 81920)])
    assert g.gpuMemory == 16 and len(devs) == 32
    assert devs[0] == ["GPU-a-_-0", "Healthy"] and devs[16] == ["GPU-b-_-0", "Healthy"]
    assert names == {"GPU-a": 0, "GPU-b": 1}
    # H7: the global survives a "restart" with different hardware
    devs2, _ = R.getDevices(g, [("GPU-c", 81920)])
    assert len(devs2) == 16


def test_watch_xids_rule():
    g = R.Globals("GiB")
    devs, _ = R.getDevices(g, [("GPU-a", 2048), ("GPU-b", 2048)])
    assert R.watchXIDs_event(devs, "GPU-b", 79) == [2, 3]
    assert R.watchXIDs_event(devs, "GPU-b", 31) == []
    assert R.watchXIDs_event(devs, "GPU-b", 43) == []
    assert R.watchXIDs_event(devs, "GPU-b", 45) == []
    assert R.watchXIDs_event(devs, "", 79) == [0, 1, 2, 3]
    assert R.watchXIDs_event(devs, None, 48) == [0, 1, 2, 3]

"""Allocate parity: the C ABI against the golden vectors and against the oracle.

Envs are compared as decoded maps, never as wire bytes (H3); the patch body is
compared byte for byte with the injected clock (H4).
"""
import random

import pytest

from gpushare_b200 import api
from oracle import ref_model as R
from tests.helpers import oracle_for_case, patch_body, product_for_case, product_step


def test_product_matches_golden(golden_cases):
    for c in golden_cases:
        table, plug = product_for_case(c)
        for k, st in enumerate(c["steps"]):
            envs, kind, patched = product_step(c, table, plug, st["request"])
            assert envs == st["expect_envs"], (c["name"], k, envs)
            assert kind == st["expect_kind"], (c["name"], k, kind)
            want = st["expect_patch"]["name"] if st["expect_patch"] else None
            if c["fail_patch"] != "always":
                assert patched == want, (c["name"], k)


def test_patch_body_bytes(golden_cases):
    c = next(x for x in golden_cases if x["name"] == "G1-binpack-1")
    table, plug = product_for_case(c)
    pods = [plug.get_pod(i) for i in range(plug.pod_count())]
    d = api.allocate(table, c["node"], R.encode_allocate_request(c["steps"][0]["request"]), pods,
                     c["clock"])
    assert d.patch == c["steps"][0]["expect_patch"]["body"].encode() == patch_body(c["clock"])
    d = api.allocate(table, c["node"], R.encode_allocate_request(c["steps"][0]["request"]), pods,
                     -5)
    assert d.patch == patch_body(-5)


def test_patch_applied_to_pod_source(golden_cases):
    c = next(x for x in golden_cases if x["name"] == "G1-binpack-1")
    table, plug = product_for_case(c)
    plug.Allocate(R.encode_allocate_request(c["steps"][0]["request"]))
    p0 = plug.get_pod(0)
    assert p0.ann_assigned == "true" and p0.ann_assume_time == str(c["clock"])
    assert plug.get_pod(1).ann_assigned == "false"


def _random_case(rng, n_pods, n_gpus, unit, ties):
    node = "node-a"
    times = [str(rng.randrange(1, 10**6)) for _ in range(n_pods)]
    if ties:
        times = [str(rng.randrange(1, max(2, n_pods // 3))) for _ in range(n_pods)]
    pods = []
    for k in range(n_pods):
        ann = {}
        r = rng.random()
        if r < 0.85:
            ann["ALIYUN_COM_GPU_MEM_IDX"] = rng.choice(
                [str(rng.randrange(n_gpus)), str(rng.randrange(n_gpus)), "x", "-1", str(n_gpus),
                 "+0", "01"])
        if rng.random() < 0.9:
            ann["ALIYUN_COM_GPU_MEM_ASSUME_TIME"] = rng.choice([times[k], times[k], times[k], "bad", ""])
        if rng.random() < 0.92:
            ann["ALIYUN_COM_GPU_MEM_ASSIGNED"] = rng.choice(["false"] * 6 + ["true", "False"])
        limits = [rng.choice([1, 2, 2, 3, "4", "1500m", None]) for _ in range(rng.choice([1, 1, 2, 3]))]
        containers = [{"name": "c%d" % i, "resources": ({} if l is None else
                                                         {"limits": {"aliyun.com/gpu-mem": l}})}
                      for i, l in enumerate(limits)]
        pods.append({"metadata": {"uid": "u%d" % (k if rng.random() < 0.93 else rng.randrange(n_pods)),
                                  "name": "p%d" % k, "namespace": rng.choice(["default", "team-b"]),
                                  "annotations": ann},
                     "spec": {"nodeName": rng.choice([node] * 9 + ["node-b"]), "containers": containers},
                     "status": {"phase": "Pending"}})
    gpus = [["GPU-%08x-0000-0000-0000-%012x" % (i, i), 16384] for i in range(n_gpus)]
    return {"name": "rand", "gpus": gpus, "unit": unit, "node": node,
            "cgpu_disable": rng.random() < 0.3, "clock": rng.randrange(1, 2**62), "pods": pods,
            "fail_list": False, "fail_patch": None}


def _random_request(rng, gpus):
    reqs = []
    for _ in range(rng.choice([1, 1, 1, 2, 3])):
        n = rng.choice([0, 1, 2, 2, 3, 4, 5])
        reqs.append(["%s-_-%d" % (rng.choice(gpus)[0], rng.randrange(16)) for _ in range(n)])
    return reqs


@pytest.mark.parametrize("n_pods,ties", [(0, False), (3, False), (8, True), (12, True),
                                          (13, False), (40, False), (64, False), (200, False)])
def test_random_sequences_match_oracle(n_pods, ties):
    """Seeded random pod sets and request sequences: every response and every patch
    target equals the oracle's. Tie order is only exercised for n <= 12 (H2)."""
    rng = random.Random(20261106 + n_pods)
    for trial in range(25 if n_pods <= 64 else 4):
        c = _random_case(rng, n_pods, rng.choice([1, 2, 8]), rng.choice(["GiB", "MiB"]), ties)
        table, plug = product_for_case(c)
        src, oracle = oracle_for_case(c)
        for step in range(min(2 * n_pods + 3, 40)):
            reqs = _random_request(rng, c["gpus"])
            before = len(src.patches)
            want = oracle.Allocate(reqs)
            envs, kind, patched = product_step(c, table, plug, reqs)
            assert envs == want, (n_pods, trial, step, kind)
            want_patched = src.patches[before][1] if len(src.patches) > before else None
            assert patched == want_patched, (n_pods, trial, step)


@pytest.mark.parametrize("n_pods,ties", [(5, True), (20, True), (64, True), (150, False)])
def test_plugin_candidate_cache_equals_stateless_allocate(n_pods, ties):
    """gs_plugin filters the installed list once (getCandidatePods) and drops a pod when it claims
    it; gs_allocate filters the list it is handed on every call. Both must decide alike at every
    step of a claim sequence -- same kind, same pod, same response bytes -- with duplicate UIDs,
    other nodes' pods and assume-time ties (any n: the two paths sort the very same sequence)."""
    rng = random.Random(977 + n_pods)
    for trial in range(12):
        c = _random_case(rng, n_pods, rng.choice([1, 2, 8]), rng.choice(["GiB", "MiB"]), ties)
        table, plug = product_for_case(c)
        for step in range(min(2 * n_pods + 3, 60)):
            wire = R.encode_allocate_request(_random_request(rng, c["gpus"]))
            pods_now = [plug.get_pod(i) for i in range(plug.pod_count())]
            want = api.allocate(table, c["node"], wire, pods_now, c["clock"], c["cgpu_disable"])
            got = plug.Allocate(wire)
            assert (got.kind, got.pod_index, got.gpu_index, got.response) == \
                (want.kind, want.pod_index, want.gpu_index, want.response), (n_pods, trial, step)


def test_ties_above_12_agree_with_oracle_port():
    """n > 12 with equal assume times: product and oracle both follow the published
    pdqsort, so they agree with each other; agreement with real Go is unpinned (H2)."""
    rng = random.Random(7)
    for n in (13, 20, 50, 64):
        c = _random_case(rng, n, 8, "GiB", ties=True)
        table, plug = product_for_case(c)
        src, oracle = oracle_for_case(c)
        for step in range(n):
            reqs = [["x"] * rng.choice([1, 2, 3, 4])]
            assert product_step(c, table, plug, reqs)[0] == oracle.Allocate(reqs), (n, step)


def test_request_scan_and_malformed_wire():
    reqs = [["a" * 44] * 1024, [], ["b" * 200, ""]]
    wire = R.encode_allocate_request(reqs)
    counts, total = api.request_scan(wire)
    assert counts == [1024, 0, 2] and total == 1026
    # unknown fields of every wire type are skipped, as a protobuf decoder would
    extra = (b"\x10\x05" + b"\x19" + b"\0" * 8 + b"\x25" + b"\0" * 4 + b"\x1a\x03abc")
    inner = b"\x0a\x01x" + b"\x10\x7f" + b"\x0a\x02yz"
    wire2 = extra + b"\x0a" + bytes([len(inner)]) + inner
    assert api.request_scan(wire2) == ([2], 2)
    table = api.DeviceTable(["GPU-a"], 16, "GiB")
    for bad in (b"\x0a\x05ab", b"\x0a", b"\x0a\x03\x0a\x05a", b"\x0b\x00", b"\xff" * 11):
        with pytest.raises(api.L.GpushareError) as e:
            api.allocate(table, "n", bad, [])
        assert e.value.status == api.L.GS_E_WIRE


def test_empty_request_no_containers():
    table = api.DeviceTable(["GPU-a", "GPU-b"], 16, "GiB")
    d = api.allocate(table, "n", b"", [])
    assert d.response == b"" and d.n_containers == 0 and d.kind == "err-no-match"


def test_response_buffer_too_small_reports_need():
    import ctypes as C
    lib = api.L.load()
    table = api.DeviceTable(["GPU-a", "GPU-b"], 16, "GiB")
    wire = R.encode_allocate_request([["x", "y"]])
    keep = []
    ctx = api.L.AllocCtx(table.h, api._str("n", keep), 0, 0)
    res = api.L.AllocResult()
    small = C.create_string_buffer(8)
    patch = C.create_string_buffer(256)
    rc = lib.gs_allocate(C.byref(ctx), wire, len(wire), None, 0, small, 8, patch, 256, C.byref(res))
    assert rc == api.L.GS_E_NOSPACE and res.resp_len > 8
    big = C.create_string_buffer(res.resp_len)
    rc = lib.gs_allocate(C.byref(ctx), wire, len(wire), None, 0, big, res.resp_len, patch, 256,
                         C.byref(res))
    assert rc == api.L.GS_OK
    assert R.decode_allocate_response(big.raw[:res.resp_len])[0]["ALIYUN_COM_GPU_MEM_IDX"] == "-1"


def test_large_mib_request_counts():
    """BASELINE config 3 sizes: 1 024 .. 40 960 ids per request at MiB granularity."""
    rng = random.Random(1)
    uuids = ["GPU-%08x-0000-0000-0000-%012x" % (i, i) for i in range(8)]
    table = api.DeviceTable(uuids, 183359, "MiB")
    for n in (1024, 4096, 16384, 40960):
        ids = ["%s-_-%d" % (uuids[rng.randrange(8)], rng.randrange(183359)) for _ in range(n)]
        wire = R.encode_allocate_request([ids])
        pod = api.Pod("u", "default", "tenant", "n", n, "5", "1", "false")
        d = api.allocate(table, "n", wire, [pod], 42)
        envs = R.decode_allocate_response(d.response)
        assert envs == [{"NVIDIA_VISIBLE_DEVICES": "5", "ALIYUN_COM_GPU_MEM_IDX": "5",
                         "ALIYUN_COM_GPU_MEM_POD": str(n), "ALIYUN_COM_GPU_MEM_CONTAINER": str(n),
                         "ALIYUN_COM_GPU_MEM_DEV": "183359"}]


def test_multichain_scan_is_exact_on_hostile_ids():
    """Containers above 1 MiB are counted by several interleaved chains that start at guessed
    record boundaries; ids full of bytes that LOOK like boundaries must not fool it."""
    rng = random.Random(99)
    for trial in range(24):
        ids = []
        for _ in range(rng.choice([300, 45000, 70000])):
            kind = rng.random()
            if kind < 0.4:
                ids.append("GPU-%08x-_-%d" % (rng.randrange(2**32), rng.randrange(200000)))
            elif kind < 0.8:      # fake records inside the id: "\n" + length byte + filler
                ids.append("".join(rng.choice(["\n", chr(rng.randrange(1, 100)), "\n\x05abcde", "x"])
                                   for _ in range(rng.randrange(0, 40))))
            elif kind < 0.9:
                ids.append("")
            else:
                ids.append("y" * rng.choice([127, 128, 300]))   # two-byte length prefixes
        reqs = [ids, ["tail"]]
        wire = R.encode_allocate_request(reqs)
        if trial % 3 == 0:        # an unknown varint field in the middle of the container
            inner = b"".join(R._ld(1, i.encode()) for i in ids[:len(ids) // 2]) + b"\x10\x07" + \
                b"".join(R._ld(1, i.encode()) for i in ids[len(ids) // 2:])
            wire = R._ld(1, inner) + R._ld(1, R._ld(1, b"tail"))
        counts, total = api.request_scan(wire)
        assert counts == [len(ids), 1] and total == len(ids) + 1, trial


def test_request_scan_fuzz_agrees_with_oracle_decoder():
    """Mutated wire bytes: the product either rejects them (GS_E_WIRE) exactly when the oracle's
    independent protobuf reader does, or counts the same ids per container."""
    rng = random.Random(4242)
    base = R.encode_allocate_request([["GPU-aaaa-_-%d" % k for k in range(300)], ["x" * 130, ""], []])
    agree_ok = agree_bad = 0
    for trial in range(3000):
        b = bytearray(base[:rng.choice([len(base), rng.randrange(1, len(base))])])
        for _ in range(rng.choice([0, 1, 1, 2, 5])):
            op = rng.random()
            i = rng.randrange(len(b))
            if op < 0.5:
                b[i] = rng.choice([0x0A, 0x00, 0x7F, 0x80, 0xFF, 0x12, 0x0B, 0x0D, 0x09, rng.randrange(256)])
            elif op < 0.75:
                del b[i]
                if not b:
                    b = bytearray(b"\x0a")
            else:
                b.insert(i, rng.randrange(256))
        wire = bytes(b)
        try:
            want = [len(c) for c in _strict_decode(wire)]
        except ValueError:
            want = None
        try:
            got, total = api.request_scan(wire)
        except api.L.GpushareError as e:
            assert e.status == api.L.GS_E_WIRE
            got = None
        assert got == want, (trial, wire[:40])
        agree_ok += want is not None
        agree_bad += want is None
    assert agree_ok > 300 and agree_bad > 300          # both outcomes are exercised


def _strict_decode(buf):
    """AllocateRequest -> per-container lists of raw id bytes, raising ValueError on anything a
    protobuf reader must refuse (truncation, over-long varints, group wire types)."""
    out = []
    for f, w, c in R._fields(buf):
        if f == 1 and w == 2:
            out.append([v for f2, w2, v in R._fields(c) if f2 == 1 and w2 == 2])
    return out


def test_stateless_allocate_is_thread_safe():
    """gs_allocate keeps per-thread scratch only: 8 threads deciding different pod lists at once
    get the answers they would get alone."""
    import threading
    table = api.DeviceTable(["GPU-%d" % i for i in range(8)], 16)
    errors = []

    def worker(t):
        rng = random.Random(t)
        pods = [api.Pod("u%d-%d" % (t, k), "default", "p%d" % k, "n", 1 + (k % 5), str((t + k) % 8),
                        str(rng.randrange(10**6)), "false") for k in range(5 + 9 * t)]
        for _ in range(200):
            n = 1 + rng.randrange(5)
            d = api.allocate(table, "n", R.encode_allocate_request([["i"] * n]), pods, 1)
            cands = sorted((int(p.ann_assume_time), -i, i) for i, p in enumerate(pods) if p.gpu_mem_limit == n)
            want = cands[0][2] if cands else -1
            if len(pods) <= 12 or len({c[0] for c in cands}) == len(cands):
                if d.pod_index != want:
                    errors.append((t, n, d.pod_index, want))
    ts = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
    [t.start() for t in ts]
    [t.join(60) for t in ts]
    assert not errors, errors[:3]

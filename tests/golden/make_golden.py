"""Writes tests/golden/allocate_cases.json and tests/golden/pattern_kat.json.

The reference has no golden vectors for this path (its only test asserts
nothing), and it cannot be run here (Go, no toolchain). So every expected value below
was derived BY HAND from the cited reference lines and typed in as a literal; this
script then checks that oracle/ref_model.py reproduces each literal before writing
the file. Parity stays "unpinned" in the strict sense: nothing here came out of a run
of the reference.

    python tests/golden/make_golden.py        # rewrites the two JSON files
"""
import ctypes
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_model as R  # noqa: E402

NODE = "node-a"
CLOCK = 1700000000123456789


def uuid(i):
    """GPU-%08x-0000-0000-0000-%012x from the index."""
    return "GPU-%08x-0000-0000-0000-%012x" % (i, i)


def gpus(n, raw_mib):
    return [[uuid(i), raw_mib] for i in range(n)]


def pod(name, limits, idx="0", assume="1000", assigned="false", node=NODE, uid=None,
        ns="default", extra_ann=None, init_limits=None):
    ann = {}
    if idx is not None:
        ann["ALIYUN_COM_GPU_MEM_IDX"] = idx
    if assume is not None:
        ann["ALIYUN_COM_GPU_MEM_ASSUME_TIME"] = assume
    if assigned is not None:
        ann["ALIYUN_COM_GPU_MEM_ASSIGNED"] = assigned
    ann.update(extra_ann or {})
    containers = []
    for k, lim in enumerate(limits):
        res = {} if lim is None else {"limits": {"aliyun.com/gpu-mem": lim}}
        containers.append({"name": "%s-c%d" % (name, k), "resources": res})
    spec = {"nodeName": node, "containers": containers}
    if init_limits:
        spec["initContainers"] = [{"name": "init", "resources": {
            "limits": {"aliyun.com/gpu-mem": x}}} for x in init_limits]
    return {"metadata": {"uid": uid or ("uid-" + name), "name": name, "namespace": ns,
                         "annotations": ann},
            "spec": spec, "status": {"phase": "Pending"}}


def ids(*pairs):
    """Device ids as kubelet would send them: any healthy ids, NOT tied to the GPU the
    scheduler extender chose (the reference never reads them, H9)."""
    return ["%s-_-%d" % (uuid(g), j) for g, j in pairs]


def ok(vis, idx, pod_total, container, dev, cgpu=False):
    e = {"NVIDIA_VISIBLE_DEVICES": vis, "ALIYUN_COM_GPU_MEM_IDX": idx,
         "ALIYUN_COM_GPU_MEM_POD": pod_total, "ALIYUN_COM_GPU_MEM_CONTAINER": container,
         "ALIYUN_COM_GPU_MEM_DEV": dev}
    if cgpu:
        e["CGPU_DISABLE"] = "true"
    return e


def err(total, unit, container, dev):
    # allocate.go:30-34
    return ok("no-gpu-has-%s%s-to-run" % (total, unit), "-1", total, container, dev)


PATCH = ('{"metadata":{"annotations":{"ALIYUN_COM_GPU_MEM_ASSIGNED":"true",'
         '"ALIYUN_COM_GPU_MEM_ASSUME_TIME":"%d"}}}' % CLOCK)


def step(request, envs, kind, patched=None):
    return {"request": request, "expect_envs": envs, "expect_kind": kind,
            "expect_patch": None if patched is None else
            {"namespace": "default", "name": patched, "body": PATCH}}


def case(name, why, gpus_, unit, pods, steps, cgpu=False, fail_list=False, fail_patch=None,
         unpinned=False):
    return {"name": name, "why": why, "gpus": gpus_, "unit": unit, "node": NODE,
            "cgpu_disable": cgpu, "clock": CLOCK, "pods": pods, "steps": steps,
            "fail_list": fail_list, "fail_patch": fail_patch, "tie_order_unpinned": unpinned}


CASES = []

# G1 -- demo/binpack-1/binpack-1.yaml:25,38-43: StatefulSet, 3 replicas x gpu-mem 2,
# BASELINE config 1 (8 x 16 GiB, GiB unit => 16 ids per GPU, DEV=16).
req2 = [ids((5, 3), (2, 7))]
CASES.append(case(
    "G1-binpack-1", "3 replicas claimed oldest first, 4th request finds nobody",
    gpus(8, 16384), "GiB",
    [pod("binpack-1-0", ["2"], assume="1000"), pod("binpack-1-1", ["2"], assume="2000"),
     pod("binpack-1-2", ["2"], assume="3000")],
    [step(req2, [ok("0", "0", "2", "2", "16")], "matched", "binpack-1-0"),
     step(req2, [ok("0", "0", "2", "2", "16")], "matched", "binpack-1-1"),
     step(req2, [ok("0", "0", "2", "2", "16")], "matched", "binpack-1-2"),
     step(req2, [err("2", "GiB", "2", "16")], "err-no-match")]))

# demo/binpack-1/job.yaml:11-15: Job gpu-job x gpu-mem 2, scheduled on GPU 3
CASES.append(case(
    "G1b-gpu-job", "job.yaml pod annotated to GPU 3",
    gpus(8, 16384), "GiB", [pod("gpu-job-x7k2p", ["2"], idx="3")],
    [step(req2, [ok("3", "3", "2", "2", "16")], "matched", "gpu-job-x7k2p")]))

# G2 -- allocate.go:182-187
CASES.append(case(
    "G2-no-candidate-8gpu", "no pod at all on an 8-GPU node => poison envs, gRPC OK",
    gpus(8, 16384), "GiB", [], [step(req2, [err("2", "GiB", "2", "16")], "err-no-match")]))

# G3 -- allocate.go:154-180: exactly one GPU => UUID, index 0, no patch
CASES.append(case(
    "G3-single-gpu-fallback", "no candidate and one GPU => NVIDIA_VISIBLE_DEVICES=<UUID>",
    gpus(1, 16384), "GiB", [],
    [step([ids((0, 1), (0, 2), (0, 3))], [ok(uuid(0), "0", "3", "3", "16")], "single-gpu")]))
CASES.append(case(
    "G3b-single-gpu-but-matched", "one GPU and a matching pod: index form, patch issued",
    gpus(1, 16384), "GiB", [pod("p", ["3"], idx="0")],
    [step([ids((0, 1), (0, 2), (0, 3))], [ok("0", "0", "3", "3", "16")], "matched", "p")]))

# G4 -- podutils.go:37-61 + allocate.go:93-111; strconv.Atoi forms (H6)
for tag, idxval, good in [
        ("abc", "abc", None), ("missing", None, None), ("out-of-range", "8", None),
        ("negative", "-1", None), ("space", " 1", None), ("empty", "", None),
        ("underscore", "1_0", None), ("overflow", "9223372036854775808", None),
        ("huge-valid-int", "9223372036854775807", None), ("hex", "0x1", None),
        ("plus", "+3", "3"), ("leading-zeros", "007", "7"), ("minus-zero", "-0", "0")]:
    if good is None:
        st = step(req2, [err("2", "GiB", "2", "16")], "err-bad-index")
    else:
        st = step(req2, [ok(good, good, "2", "2", "16")], "matched", "p")
    CASES.append(case("G4-idx-" + tag, "ALIYUN_COM_GPU_MEM_IDX=%r" % (idxval,),
                      gpus(8, 16384), "GiB", [pod("p", ["2"], idx=idxval)], [st]))

# G5 -- allocate.go:55-57,114-124 (H10): two containers 3 + 1 against limits 3 + 1
req31 = [ids((0, 0), (0, 1), (1, 5)), ids((7, 15))]
CASES.append(case(
    "G5-two-containers", "POD counts all containers, CONTAINER counts its own",
    gpus(8, 16384), "GiB", [pod("multi", ["3", "1"], idx="6")],
    [step(req31, [ok("6", "6", "4", "3", "16"), ok("6", "6", "4", "1", "16")], "matched",
          "multi")]))
CASES.append(case(
    "G5b-limits-sum-mismatch", "pod limits 3+2=5 != request 4 => error for both containers",
    gpus(8, 16384), "GiB", [pod("multi", ["3", "2"], idx="6")],
    [step(req31, [err("4", "GiB", "3", "16"), err("4", "GiB", "1", "16")], "err-no-match")]))
CASES.append(case(
    "G5c-container-without-limit", "containers lacking the resource add nothing; init "
    "containers are not counted (podutils.go:125)",
    gpus(8, 16384), "GiB", [pod("side", ["2", None], idx="1", init_limits=["5"])],
    [step(req2, [ok("1", "1", "2", "2", "16")], "matched", "side")]))

# G6 -- allocate.go:125-127,174-176; buildErrResponse never adds CGPU_DISABLE
CASES.append(case(
    "G6-cgpu-disable-matched", "node label cgpu.disable.isolation=true",
    gpus(8, 16384), "GiB", [pod("p", ["2"], idx="2")],
    [step(req2, [ok("2", "2", "2", "2", "16", cgpu=True)], "matched", "p"),
     step(req2, [err("2", "GiB", "2", "16")], "err-no-match")], cgpu=True))
CASES.append(case(
    "G6b-cgpu-disable-single-gpu", "fallback path also carries CGPU_DISABLE",
    gpus(1, 16384), "GiB", [],
    [step(req2, [ok(uuid(0), "0", "2", "2", "16", cgpu=True)], "single-gpu")], cgpu=True))

# G7 -- podmanager.go:331,342 (H2): n <= 12 => insertion sort with `<=`: ties reversed
CASES.append(case(
    "G7-ties-reversed", "three pods, equal assume time: the LAST listed is claimed first",
    gpus(8, 16384), "GiB",
    [pod("a", ["2"], idx="1", assume="500"), pod("b", ["2"], idx="2", assume="500"),
     pod("c", ["2"], idx="3", assume="500")],
    [step(req2, [ok("3", "3", "2", "2", "16")], "matched", "c"),
     step(req2, [ok("2", "2", "2", "2", "16")], "matched", "b"),
     step(req2, [ok("1", "1", "2", "2", "16")], "matched", "a")]))
CASES.append(case(
    "G7b-ties-mixed", "times [5,3,5,3] => order (3,idx3) (3,idx1) (5,idx2) (5,idx0)",
    gpus(8, 16384), "GiB",
    [pod("p0", ["2"], idx="0", assume="5"), pod("p1", ["2"], idx="1", assume="3"),
     pod("p2", ["2"], idx="2", assume="5"), pod("p3", ["2"], idx="3", assume="3")],
    [step(req2, [ok("3", "3", "2", "2", "16")], "matched", "p3"),
     step(req2, [ok("1", "1", "2", "2", "16")], "matched", "p1"),
     step(req2, [ok("2", "2", "2", "2", "16")], "matched", "p2"),
     step(req2, [ok("0", "0", "2", "2", "16")], "matched", "p0")]))

# G8 -- podutils.go:78-119,64-75
CASES.append(case(
    "G8-assigned-true", "ASSIGNED=true is not a candidate",
    gpus(8, 16384), "GiB", [pod("p", ["2"], assigned="true")],
    [step(req2, [err("2", "GiB", "2", "16")], "err-no-match")]))
CASES.append(case(
    "G8b-assigned-absent", "no ASSIGNED annotation is not a candidate",
    gpus(8, 16384), "GiB", [pod("p", ["2"], assigned=None)],
    [step(req2, [err("2", "GiB", "2", "16")], "err-no-match")]))
CASES.append(case(
    "G8c-assigned-False", "only the exact string \"false\" qualifies",
    gpus(8, 16384), "GiB", [pod("p", ["2"], assigned="False")],
    [step(req2, [err("2", "GiB", "2", "16")], "err-no-match")]))
CASES.append(case(
    "G8d-assume-absent", "no ASSUME_TIME annotation is not a candidate",
    gpus(8, 16384), "GiB", [pod("p", ["2"], assume=None)],
    [step(req2, [err("2", "GiB", "2", "16")], "err-no-match")]))
CASES.append(case(
    "G8e-assume-unparsable", "unparsable ASSUME_TIME counts as 0: sorts before everyone",
    gpus(8, 16384), "GiB",
    [pod("old", ["2"], idx="1", assume="10"), pod("junk", ["2"], idx="2", assume="soon"),
     pod("signed", ["2"], idx="3", assume="+5")],
    # "soon" -> 0 and "+5" -> 0 (ParseUint rejects signs); ties reversed => signed, junk, old
    [step(req2, [ok("3", "3", "2", "2", "16")], "matched", "signed"),
     step(req2, [ok("2", "2", "2", "2", "16")], "matched", "junk"),
     step(req2, [ok("1", "1", "2", "2", "16")], "matched", "old")]))
CASES.append(case(
    "G8f-zero-limit", "a pod asking for 0 gpu-mem is never a candidate",
    gpus(8, 16384), "GiB", [pod("z", ["0"])],
    [step([[]], [err("0", "GiB", "0", "16")], "err-no-match")]))

# G9 -- MiB unit (nvidia.go:31-38, allocate.go:30)
req1024 = [["%s-_-%d" % (uuid(j % 2), 100 + j) for j in range(1024)]]
CASES.append(case(
    "G9-mib-unit", "MiB: DEV = raw total, error string says MiB",
    gpus(2, 16384), "MiB", [pod("m", ["1024"], idx="1")],
    [step(req1024, [ok("1", "1", "1024", "1024", "16384")], "matched", "m"),
     step(req1024, [err("1024", "MiB", "1024", "16384")], "err-no-match")]))
CASES.append(case(
    "G9b-gib-floor", "GiB: 16500 MiB / 1024 floors to 16; every GPU inherits GPU 0 (H7)",
    [[uuid(0), 16500], [uuid(1), 81920]], "GiB", [pod("p", ["2"], idx="1")],
    [step(req2, [ok("1", "1", "2", "2", "16")], "matched", "p")]))

# G10 -- allocate.go:79-111 (A12): no fall-through to the next candidate
CASES.append(case(
    "G10-bad-older-blocks-younger", "older pod of the same size has a bad IDX",
    gpus(8, 16384), "GiB",
    [pod("young-valid", ["2"], idx="4", assume="2000"),
     pod("old-bad", ["2"], idx="nope", assume="1000")],
    [step(req2, [err("2", "GiB", "2", "16")], "err-bad-index"),
     step(req2, [err("2", "GiB", "2", "16")], "err-bad-index")]))

# podmanager.go:273-293
CASES.append(case(
    "N1-other-node", "a pod bound to another node is skipped",
    gpus(8, 16384), "GiB", [pod("elsewhere", ["2"], node="node-b"),
                            pod("here", ["2"], idx="5", assume="9999")],
    [step(req2, [ok("5", "5", "2", "2", "16")], "matched", "here")]))
CASES.append(case(
    "N2-uid-dedup", "second pod with a UID already seen is dropped BEFORE the candidate "
    "filter, even when the first one was not a candidate",
    gpus(8, 16384), "GiB",
    [pod("first", ["2"], assigned="true", uid="same"), pod("second", ["2"], uid="same")],
    [step(req2, [err("2", "GiB", "2", "16")], "err-no-match")]))
CASES.append(case(
    "N3-size-picks-pod", "the request size selects among candidates of different sizes",
    gpus(8, 16384), "GiB",
    [pod("four", ["4"], idx="1", assume="1"), pod("two", ["2"], idx="2", assume="2")],
    [step(req2, [ok("2", "2", "2", "2", "16")], "matched", "two")]))

# H5 -- resource.Quantity.Value() forms
for tag, q, total in [("int", 2, "2"), ("milli", "2000m", "2"), ("round-up", "1500m", "2"),
                      ("exp", "2e0", "2"), ("Ki", "1Ki", "1024"), ("decimal", "2.0", "2")]:
    n = int(total)
    rq = [["%s-_-%d" % (uuid(0), j) for j in range(n)]]
    CASES.append(case("H5-quantity-" + tag, "limit written as %r" % (q,),
                      gpus(8, 16384), "MiB" if n > 16 else "GiB", [pod("q", [q], idx="0")],
                      [step(rq, [ok("0", "0", total, total, "16384" if n > 16 else "16")],
                            "matched", "q")]))

# allocate.go:64-67,133-151
CASES.append(case(
    "F1-lister-fails", "getCandidatePods error => buildErrResponse",
    gpus(8, 16384), "GiB", [pod("p", ["2"])],
    [step(req2, [err("2", "GiB", "2", "16")], "err-pods")], fail_list=True))
CASES.append(case(
    "F2-patch-fails", "PATCH error => buildErrResponse, pod stays unassigned",
    gpus(8, 16384), "GiB", [pod("p", ["2"])],
    [step(req2, [err("2", "GiB", "2", "16")], "err-patch")], fail_patch="always"))
CASES.append(case(
    "F3-patch-conflict-retried", "OptimisticLockErrorMsg is retried once (allocate.go:140)",
    gpus(8, 16384), "GiB", [pod("p", ["2"], idx="2")],
    [step(req2, [ok("2", "2", "2", "2", "16")], "matched", "p")], fail_patch="conflict-once"))

# n > 12 with DISTINCT times: any correct sort agrees (H2) -- 20 pods, sizes 1..20 shuffled
order = [7, 2, 19, 11, 0, 14, 5, 17, 9, 3, 12, 16, 1, 8, 18, 4, 10, 15, 6, 13]
big = [pod("t%02d" % k, ["3"], idx=str(k % 8), assume=str(1000 + 37 * order[k]))
       for k in range(20)]
req3 = [ids((0, 1), (0, 2), (0, 3))]
steps = []
for rank in range(20):
    k = order.index(rank)
    steps.append(step(req3, [ok(str(k % 8), str(k % 8), "3", "3", "16")], "matched",
                      "t%02d" % k))
steps.append(step(req3, [err("3", "GiB", "3", "16")], "err-no-match"))
CASES.append(case("S1-twenty-distinct", "20 equal-size pods, distinct times: oldest first",
                  gpus(8, 16384), "GiB", big, steps))


def run_oracle(c):
    g = R.Globals(c["unit"])
    src = R.PodSource(c["pods"], fail_list=c["fail_list"], fail_patch=c["fail_patch"])
    plug = R.NvidiaDevicePlugin(g, [tuple(x) for x in c["gpus"]], c["node"], src,
                                c["cgpu_disable"], clock=lambda: c["clock"])
    out = []
    for st in c["steps"]:
        before = len(src.patches)
        envs = plug.Allocate(st["request"])
        patch = src.patches[before] if len(src.patches) > before else None
        out.append((envs, patch))
    return out


def pattern_kat(pattern=0):
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "libgs_oracle.so"))
    out = []
    for i, seed in [(0, 0), (1, 0), (0, 1), (255, 1234), (4096, 77), ((1 << 32) - 1, 77),
                    (1 << 32, 77), ((1 << 32) + 5, 0xFFFFFFFF), (11250000000, 20261106)]:
        w = (ctypes.c_uint32 * 4)()
        assert lib.gs_oracle_pattern_word_p(ctypes.c_uint64(i), ctypes.c_uint32(seed),
                                            ctypes.c_uint32(pattern), w) == 0
        out.append({"i": i, "seed": seed, "words": list(w)})
    return out


def main():
    for c in CASES:
        got = run_oracle(c)
        for k, (st, (envs, patch)) in enumerate(zip(c["steps"], got)):
            assert envs == st["expect_envs"], (c["name"], k, envs, st["expect_envs"])
            if st["expect_patch"] is None:
                assert patch is None, (c["name"], k, patch)
            else:
                assert patch == (st["expect_patch"]["namespace"], st["expect_patch"]["name"],
                                 st["expect_patch"]["body"].encode()), (c["name"], k, patch)
    with open(os.path.join(HERE, "allocate_cases.json"), "w") as f:
        json.dump({"generated_by": "tests/golden/make_golden.py",
                   "provenance": "hand-derived from the reference's source lines; confirmed "
                                 "by oracle/ref_model.py; NOT output of the reference",
                   "cases": CASES}, f, indent=1)
    with open(os.path.join(HERE, "pattern_kat.json"), "w") as f:
        json.dump({"generated_by": "tests/golden/make_golden.py",
                   "provenance": "oracle/pattern.c (the probe pattern has no reference "
                                 "counterpart); first vector also checked by hand",
                   "vectors": pattern_kat(0), "hash_vectors": pattern_kat(1)}, f, indent=1)
    print("wrote %d cases" % len(CASES))


if __name__ == "__main__":
    main()

"""The C++ baseline restatement agrees with the golden vectors and the Python oracle, so
the number bench.py times as `cpu_baseline` / `--impl reference` is the same algorithm."""
from oracle import ref_cpp
from oracle import ref_model as R


def test_ref_cpp_matches_golden(golden_cases):
    ran = 0
    for c in golden_cases:
        if c["fail_list"] or c["fail_patch"]:
            continue  # lister/patcher failures are caller-side; not modelled in the baseline
        ref = ref_cpp.RefPlugin([g[0] for g in c["gpus"]], c["gpus"][0][1] if c["gpus"] else 0,
                                c["unit"], c["node"], c["cgpu_disable"])
        ref.set_pods_k8s(c["pods"])
        names = [p["metadata"]["name"] for p in c["pods"]]
        for k, st in enumerate(c["steps"]):
            kind, resp, patched = ref.allocate(R.encode_allocate_request(st["request"]), c["clock"])
            assert R.decode_allocate_response(resp) == st["expect_envs"], (c["name"], k)
            assert kind == st["expect_kind"], (c["name"], k)
            if st["expect_patch"]:
                assert names[patched] == st["expect_patch"]["name"]
                assert ref.last_patch() == st["expect_patch"]["body"].encode()
            else:
                assert patched == -1
        ran += 1
    assert ran >= 40


def test_ref_cpp_list_and_watch_matches_oracle():
    uuids = ["GPU-%08x-0000-0000-0000-%012x" % (i, i) for i in range(3)]
    ref = ref_cpp.RefPlugin(uuids, 2500, "MiB", "n")
    g = R.Globals("MiB")
    devs, _ = R.getDevices(g, [(u, 2500) for u in uuids])
    assert ref.units == 2500 and ref.get_devices() == 7500
    assert ref.list_and_watch() == R.encode_list_and_watch(devs)
    ref.mark_unhealthy(2501)
    devs[2501][1] = "Unhealthy"
    assert ref.list_and_watch() == R.encode_list_and_watch(devs)

"""N > 1 path on CPU: two ranks (gloo), one GPU shard each, no data-path collective.

Covers what bench.py does per rank apart from the probe: each rank builds its own shard's
workload, decides it through the C ABI, and only the bookkeeping (max time, summed
decisions) crosses ranks. Checks that shards are disjoint and each rank's answers equal
the oracle's for its own shard."""
import os
import subprocess
import sys
import textwrap

from tests.conftest import ROOT

WORKER = textwrap.dedent("""
    import ctypes as C, json, os, sys, time
    sys.path.insert(0, %r)
    import torch, torch.distributed as dist
    import bench
    from gpushare_b200 import api, _lib as L
    from oracle import ref_model as R
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    uuid = "GPU-%%08x-0000-0000-0000-%%012x" %% (rank, rank)
    pods_k8s, wires, sizes, _units = bench.build_workload(rank, uuid, 4096)      # small shard: 4 GiB at MiB
    table = api.DeviceTable([uuid], 4096, api.MiB)
    plug = api.NvidiaDevicePlugin(table, bench.NODE)
    plug.set_clock(7)
    plug.set_pods(api.PodList([api.Pod.from_k8s(p) for p in pods_k8s]))
    src = R.PodSource(pods_k8s)
    oracle = R.NvidiaDevicePlugin(R.Globals("MiB"), [(uuid, 4096)], bench.NODE, src, clock=lambda: 7)
    dist.barrier()
    t0 = time.perf_counter()
    n = 0
    for k, w in enumerate(wires):
        d = plug.Allocate(w)
        assert d.kind == "matched" and d.pod_index == k, (rank, k, d.kind)
        assert R.decode_allocate_response(d.response) == oracle.Allocate(R.decode_allocate_request(w))
        n += 1
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64)
    cnt = torch.tensor([n], dtype=torch.int64)
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    dist.all_reduce(cnt)
    uids = [None] * world
    dist.all_gather_object(uids, [p["metadata"]["uid"] for p in pods_k8s])
    if rank == 0:
        print(json.dumps({"decisions": int(cnt.item()), "max_s": float(dt.item()), "world": world,
                          "disjoint": len(set(sum(uids, []))) == sum(len(u) for u in uids),
                          "sizes_differ": True}))
    dist.destroy_process_group()
""")


def test_two_ranks_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    # the small shard (4096 units) cannot hold the 1-40 GiB tenants: shrink them
    env["GS_BENCH_SMALL"] = "1"
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                          "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port", "29731",
                          str(script)], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    assert r["world"] == 2 and r["decisions"] == 16 and r["disjoint"] and r["max_s"] > 0


def test_reference_arm_prints_contract_line():
    """--impl reference: the contract line, the matched over-the-socket legs and the N-GPU
    single-queue stream (short steps: GS_BENCH_SMALL)."""
    env = dict(os.environ, GS_BENCH_SMALL="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                          "--steps", "2", "--warmup", "3"], capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    r = json.loads(out.stdout.strip().splitlines()[-1])
    assert r["impl"] == "reference" and r["metric"] == "allocate_decisions_per_s"
    assert r["cpu_baseline"]["kind"] == "port" and r["cpu_baseline"]["cores"] == 1
    assert r["e2e"]["value"] == r["value"] > 0 and r["e2e"]["h2d_bytes_per_step"] == 0
    assert r["config"]["waves_per_step"] == 8 and r["config"]["request_bytes_avg"] > 0
    assert 0 < r["over_socket"]["value"] <= r["value"] * 1.5
    assert r["over_socket"]["grpcio"]["value"] > 0
    sq = r["single_queue"]
    assert sq["n_gpus"] == 2 and sq["fake_ids"] == 2 * 183359 and sq["pods_in_list"] == 128 and sq["value"] > 0


def test_single_queue_node_workload_is_decided_like_the_oracle():
    """The N-GPU single-queue stream bench.py drives (one plugin over the node, tenants of all
    GPUs interleaved): every request picks its own pod and GPU, in the product and in the oracle."""
    import bench
    from gpushare_b200 import api
    from oracle import ref_model as R
    os.environ["GS_BENCH_SMALL"] = "1"
    try:
        import importlib
        importlib.reload(bench)
        uuids = [bench.synthetic_uuid(g) for g in range(3)]
        pods_k8s, stream, sizes = bench.build_node_workload(3, uuids, 4096)
        assert len(pods_k8s) == 3 * 64 and len(stream) == 3 * 8 and len(sizes) == 3
        table = api.DeviceTable(uuids, 4096, api.MiB)
        plug = api.NvidiaDevicePlugin(table, bench.NODE)
        plug.set_clock(7)
        plug.set_pods(api.PodList([api.Pod.from_k8s(p) for p in pods_k8s]))
        oracle = R.NvidiaDevicePlugin(R.Globals("MiB"), [(u, 4096) for u in uuids], bench.NODE,
                                      R.PodSource(pods_k8s), clock=lambda: 7)
        for g, k, pod_index, w in stream:
            d = plug.Allocate(w)
            assert (d.kind, d.pod_index, d.gpu_index) == ("matched", pod_index, g)
            assert R.decode_allocate_response(d.response) == oracle.Allocate(R.decode_allocate_request(w))
    finally:
        del os.environ["GS_BENCH_SMALL"]
        importlib.reload(bench)


def test_our_arm_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        return
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode != 0 and "no CPU fallback" in (out.stderr + out.stdout)

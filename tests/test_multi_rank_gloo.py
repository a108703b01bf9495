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
    pods_k8s, wires, sizes = bench.build_workload(rank, uuid, 4096)      # small shard: 4 GiB at MiB
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
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                          "--steps", "2", "--warmup", "3"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    r = json.loads(out.stdout.strip().splitlines()[-1])
    assert r["impl"] == "reference" and r["metric"] == "allocate_decisions_per_s"
    assert r["cpu_baseline"]["kind"] == "port" and r["cpu_baseline"]["cores"] == 1
    assert r["e2e"]["value"] == r["value"] > 0 and r["e2e"]["h2d_bytes_per_step"] == 0


def test_our_arm_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        return
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode != 0 and "no CPU fallback" in (out.stderr + out.stdout)

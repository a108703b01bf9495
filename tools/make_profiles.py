"""Turns the raw captures in gpurun_out/ into the tracked summaries under profiles/.

    python tools/make_profiles.py r01      # after a gpurun call that produced launches.csv,
                                           # probe_full.ncu-rep, bench_*.json
"""
import collections
import csv
import json
import os
import shutil
import statistics
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT, PROF = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"


def launches():
    src = os.path.join(OUT, "launches.csv")
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    hdr, data = rows[0], rows[1:]
    ki, vi, gi, bi = (hdr.index(x) for x in ("Kernel Name", "Metric Value", "Grid Size", "Block Size"))
    agg = collections.OrderedDict()
    for r in data:
        agg.setdefault(r[ki].replace("<unnamed>::", ""), []).append((float(r[vi].replace(",", "")), r[gi], r[bi]))
    tot = sum(v[0] for vs in agg.values() for v in vs)
    with open(os.path.join(PROF, tag + "_bench_launches.md"), "w") as f:
        f.write("# %s -- ncu launch list of `python bench.py --steps 3 --warmup 3 --no-cpu-baseline`\n\n" % tag)
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv` on 1 x B200, after the "
                "same command exited 0 without ncu. Times are cold-cache and serialised by the profiler: compare "
                "SHARES, not absolutes. Raw CSV beside this file.\n\n")
        f.write("| kernel | launches | total ms | share | median ms | grid x block |\n|---|---:|---:|---:|---:|---|\n")
        for name, vs in agg.items():
            t = [v[0] for v in vs]
            f.write("| `%s` | %d | %.2f | %.1f%% | %.3f | %s x %s |\n" % (
                name, len(t), sum(t) / 1e6, 100 * sum(t) / tot, statistics.median(t) / 1e6, vs[0][1], vs[0][2]))
        f.write("\nOnly this repository's probe kernels launch on the GPU during the bench; the store pass and "
                "the load+compare pass split the GPU time about evenly, as 1 B stored + 1 B loaded per slice byte "
                "predicts.\n")
    shutil.copy(src, os.path.join(PROF, tag + "_bench_launches.csv"))


WANT = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes.sum.per_second", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "l1tex__m_l1tex2xbar_write_bytes_mem_global_op_tma_st.sum", "sm__cycles_elapsed.avg.per_second",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]


def full():
    rep = os.path.join(OUT, "probe_full.ncu-rep")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    lines = ["| %s | %s | %s |" % (w, units[idx[w]], " | ".join(r[idx[w]][:44] for r in data))
             for w in WANT if w in idx]

    def num(r, k):
        return float(r[idx[k]].replace(",", ""))
    wr = [r for r in data if "write" in r[idx["Kernel Name"]]][0]
    vr = [r for r in data if "verify" in r[idx["Kernel Name"]]][0]
    slice_bytes = 4 << 30
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}

    def b(r, k):
        return num(r, k) * scale[units[idx[k]]]
    traffic = b(wr, "dram__bytes_read.sum") + b(wr, "dram__bytes_write.sum") + \
        b(vr, "dram__bytes_read.sum") + b(vr, "dram__bytes_write.sum")
    with open(os.path.join(PROF, tag + "_probe_ncu_full.md"), "w") as f:
        f.write("# %s -- `ncu --set full` of the two probe kernels\n\n" % tag)
        f.write("Command: `ncu --set full --clock-control none --import-source on -k regex:probe_ -s 2 -c 4 "
                "python tools/probe_prof.py` (one 4 GiB slice, both passes; the same command exited 0 without "
                "ncu first). 1 x B200. Numbers under the profiler are never bench values; they explain them.\n\n")
        f.write("| metric | unit | " + " | ".join("launch %d" % (i + 3) for i in range(len(data))) + " |\n")
        f.write("|---|---|" + "---|" * len(data) + "\n" + "\n".join(lines) + "\n\n")
        f.write("DRAM traffic of one store+load pair: %.3f GB for %.3f GB of algorithmic bytes (ratio %.4f): no "
                "re-reads, no read-for-ownership; the small deficit is store data still dirty in the L2 when the "
                "store kernel ends.\n" % (traffic / 1e9, 2 * slice_bytes / 1e9, traffic / (2 * slice_bytes)))
    json.dump({"source": "profiles/%s_probe_ncu_full.md (ncu --set full, one 4 GiB slice, 1 x B200)" % tag,
               "slice_bytes": slice_bytes,
               "write_kernel": {"dram_bytes_read": b(wr, "dram__bytes_read.sum"),
                                "dram_bytes_write": b(wr, "dram__bytes_write.sum")},
               "verify_kernel": {"dram_bytes_read": b(vr, "dram__bytes_read.sum"),
                                 "dram_bytes_write": b(vr, "dram__bytes_write.sum")},
               "traffic_over_algorithmic": round(traffic / (2 * slice_bytes), 4)},
              open(os.path.join(PROF, "probe_traffic.json"), "w"), indent=1)


def benches():
    for src, dst in (("bench_ours.json", "_bench_1gpu.json"), ("bench_ref.json", "_bench_reference_1gpu.json"),
                     ("allocate_sweep.json", "_allocate_sweep.json"), ("harness_8gpu.json", "_harness_8gpu.json")):
        p = os.path.join(OUT, src)
        if os.path.exists(p) and os.path.getsize(p):
            shutil.copy(p, os.path.join(PROF, tag + dst))


if __name__ == "__main__":
    launches()
    full()
    benches()
    print("profiles updated:", sorted(os.listdir(PROF)))

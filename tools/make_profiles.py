"""Turns the raw captures in gpurun_out/ into the tracked summaries under profiles/.

    python tools/make_profiles.py r02 gpurun_out/r2d   # after a gpurun call that produced launches.csv,
                                                       # probe_full.ncu-rep, probe_bulk_verify.ncu-rep,
                                                       # ncu_traffic40.csv
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
PROF = os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
OUT = os.path.join(ROOT, sys.argv[2]) if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out")


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
        f.write("# %s -- ncu launch list of `python bench.py --steps 2 --warmup 3 --no-cpu-baseline`\n\n" % tag)
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv` on 1 x B200, after the "
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


def raw_rows(name):
    rep = os.path.join(OUT, name)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    return {h: i for i, h in enumerate(hdr)}, units, data


def full():
    idx, units, data = raw_rows("probe_full.ncu-rep")
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
        bulk = os.path.join(OUT, "probe_bulk_verify.ncu-rep")
        if os.path.exists(bulk):
            bidx, bunits, bdata = raw_rows("probe_bulk_verify.ncu-rep")
            extra = ["l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum",
                     "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum"]
            f.write("\n## Side by side: the load+compare pass through bulk async loads (verify variant 521)\n\n"
                    "Command: the same with `python tools/probe_prof.py 4 521 1` and `-k regex:probe_verify_bulk -s 1 "
                    "-c 2`: 64 KiB stages, 8 consumer warps + the producer warp, one CTA per SM. The loads are "
                    "issued by the TMA engine (`..._op_tma_ld` = the slice, SASS `UBLKCP`), compared from shared "
                    "memory. It moves the same bytes with MORE warp instructions than the LDG kernel (LDS + "
                    "compare + mbarrier waits) and is slower; `profiles/r02_read_ceiling.txt` shows why no form "
                    "of this kernel can win: with the compare removed altogether the TMA engine reads 7.37-7.38 "
                    "TB/s, which is what the LDG kernel already sustains. The LDG kernel stays the default.\n\n")
            f.write("| metric | unit | " + " | ".join("launch %d" % (i + 2) for i in range(len(bdata))) + " |\n")
            f.write("|---|---|" + "---|" * len(bdata) + "\n")
            for w in WANT + extra:
                if w in bidx:
                    f.write("| %s | %s | %s |\n" % (w, bunits[bidx[w]], " | ".join(r[bidx[w]][:44] for r in bdata)))
    out = {"source": "profiles/%s_probe_ncu_full.md (ncu --set full, one 4 GiB slice, 1 x B200)" % tag,
           "slice_bytes": slice_bytes,
           "write_kernel": {"dram_bytes_read": b(wr, "dram__bytes_read.sum"),
                            "dram_bytes_write": b(wr, "dram__bytes_write.sum")},
           "verify_kernel": {"dram_bytes_read": b(vr, "dram__bytes_read.sum"),
                             "dram_bytes_write": b(vr, "dram__bytes_write.sum")},
           "traffic_over_algorithmic": round(traffic / (2 * slice_bytes), 4)}
    t40 = os.path.join(OUT, "ncu_traffic40.csv")
    if os.path.exists(t40):
        # the same pair at BASELINE config 4's largest tenant (40 GiB): four metrics per launch
        rows = [r for r in csv.reader(l for l in open(t40) if l.startswith('"'))]
        hdr, data = rows[0], rows[1:]
        ki, mi, vi, ii = (hdr.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "ID"))
        per = collections.OrderedDict()
        for r in data:
            per.setdefault(r[ii], {"kernel": r[ki]})[r[mi]] = float(r[vi].replace(",", ""))
        launches = list(per.values())[-2:]          # the last store + load pair
        t = sum(l["dram__bytes_read.sum"] + l["dram__bytes_write.sum"] for l in launches)
        n40 = 40 << 30
        out = {"source": "profiles/%s_probe_traffic_40gib.csv (ncu --metrics dram__bytes_read.sum,"
                         "dram__bytes_write.sum,..., one 40 GiB slice, 1 x B200)" % tag,
               "slice_bytes": n40, "launches": launches,
               "traffic_over_algorithmic": round(t / (2 * n40), 5),
               "at_4_gib": out}
        shutil.copy(t40, os.path.join(PROF, tag + "_probe_traffic_40gib.csv"))
        with open(os.path.join(PROF, tag + "_probe_ncu_full.md"), "a") as f:
            f.write("\n## DRAM traffic at 40 GiB (BASELINE config 4's largest tenant)\n\n`ncu --metrics dram__bytes_read.sum,"
                    "dram__bytes_write.sum,gpu__time_duration.sum,gpu__dram_throughput... -k regex:probe_ -s 2 -c 4 python "
                    "tools/probe_prof.py 40` (raw CSV: `%s_probe_traffic_40gib.csv`): the last store+load pair moved "
                    "%.3f GB through DRAM for %.3f GB of algorithmic bytes (ratio %.5f).\n\n| kernel | DRAM read | DRAM "
                    "written | duration | dram throughput |\n|---|---:|---:|---:|---:|\n" % (tag, t / 1e9, 2 * n40 / 1e9, t / (2 * n40)))
            for l in launches:
                f.write("| `%s` | %.3f GB | %.3f GB | %.3f ms | %.1f %% of peak |\n" % (
                    l["kernel"].replace("void ", "")[:48], l["dram__bytes_read.sum"] / 1e9, l["dram__bytes_write.sum"] / 1e9,
                    l["gpu__time_duration.sum"] / 1e6, l["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]))
    json.dump(out, open(os.path.join(PROF, "probe_traffic.json"), "w"), indent=1)


def benches():
    for src, dst in (("bench_ours.json", "_bench_1gpu.json"), ("bench_ref.json", "_bench_reference_1gpu.json"),
                     ("allocate_sweep.json", "_allocate_sweep.json"), ("harness_8gpu.json", "_harness_8gpu.json")):
        p = os.path.join(OUT, src)
        if os.path.exists(p) and os.path.getsize(p):
            shutil.copy(p, os.path.join(PROF, tag + dst))


if __name__ == "__main__":
    launches()
    full()
    print("profiles updated:", sorted(os.listdir(PROF)))

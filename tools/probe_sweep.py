"""Tuning aid: time every probe kernel variant on one GPU (run under gpurun).

usage: python tools/probe_sweep.py [GiB] [reps]
Prints one line per (pass, variant code, CTAs/SM): best and median GB/s by CUDA events.
"""
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402


def main():
    gib = float(sys.argv[1]) if len(sys.argv) > 1 else 8.0
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 7
    nbytes = int(gib * (1 << 30)) & ~0xFFFFF
    out = []
    L.load(strict=False)
    with Probe(0) as p:
        print("SMs", p.sm_count, "mem", p.mem_info(), flush=True)
        d = p.alloc(nbytes)
        sl = [(d, nbytes, 1234)]
        p.run(sl)  # warm
        forms_w, forms_v = range(4), range(5)
        for code in [100 + 10 * st + sub for st in (0, 1) for sub in range(5)]:
            p.set_variant(write=code)
            ts = []
            for _ in range(reps):
                _, tm = p.run(sl, L.GS_PROBE_WRITE)
                ts.append(tm.write_ms)
            v, _ = p.run(sl, L.GS_PROBE_VERIFY)
            assert v[0].clean, (code, v)
            rec = dict(kind="write", code=code, ctas=1, best_gbps=nbytes / min(ts) / 1e6,
                       med_gbps=nbytes / statistics.median(ts) / 1e6)
            out.append(rec)
            print(json.dumps(rec), flush=True)
        if os.environ.get("SWEEP_BULK_ONLY"):
            forms_w, forms_v = [], []
        for usel in (0, 1, 2):
            for ctas in (0, 2, 4, 8):
                for fw in forms_w:
                    code = usel * 10 + fw
                    p.set_variant(write=code, ctas_per_sm=ctas)
                    ts = []
                    for _ in range(reps):
                        _, tm = p.run(sl, L.GS_PROBE_WRITE)
                        ts.append(tm.write_ms)
                    best, med = min(ts), statistics.median(ts)
                    rec = dict(kind="write", code=code, ctas=ctas, best_gbps=nbytes / best / 1e6,
                               med_gbps=nbytes / med / 1e6)
                    out.append(rec)
                    print(json.dumps(rec), flush=True)
                p.set_variant(write=0)
                p.run(sl, L.GS_PROBE_WRITE)
                for fv in forms_v:
                    code = usel * 10 + fv
                    p.set_variant(verify=code, ctas_per_sm=ctas)
                    ts = []
                    for _ in range(reps):
                        v, tm = p.run(sl, L.GS_PROBE_VERIFY)
                        assert v[0].clean, v
                        ts.append(tm.verify_ms)
                    best, med = min(ts), statistics.median(ts)
                    rec = dict(kind="verify", code=code, ctas=ctas, best_gbps=nbytes / best / 1e6,
                               med_gbps=nbytes / med / 1e6)
                    out.append(rec)
                    print(json.dumps(rec), flush=True)
        p.free(d)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/probe_sweep.json", "w") as f:
        json.dump(out, f, indent=1)
    for kind in ("write", "verify"):
        top = sorted((r for r in out if r["kind"] == kind), key=lambda r: -r["med_gbps"])[:5]
        print("TOP", kind, json.dumps(top))


if __name__ == "__main__":
    main()

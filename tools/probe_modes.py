"""Tuning aid: separate write-only / verify-only / both-pass timings (run under gpurun)."""
import os
import statistics
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

gib = float(sys.argv[1]) if len(sys.argv) > 1 else 8.0
nbytes = int(gib * (1 << 30)) & ~0xFFFFF
mon = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,clocks.mem,power.draw,clocks_event_reasons.active",
                        "--format=csv,noheader", "-lms", "100"], stdout=subprocess.PIPE, text=True)
with Probe(0) as p:
    d = p.alloc(nbytes)
    sl = [(d, nbytes, 5)]
    for _ in range(3):
        p.run(sl)

    def med(passes, n=7):
        w, v = [], []
        for _ in range(n):
            res, tm = p.run(sl, passes)
            assert res[0].clean
            w.append(tm.write_ms)
            v.append(tm.verify_ms)
        f = lambda ts: nbytes / statistics.median(ts) / 1e6 if statistics.median(ts) > 0 else 0
        return f(w), f(v)

    for rate in (0, 5000, 6500, 7000, 7300, 7450, 7600):
        p.set_write_rate(rate)
        print("rate %5d: write-only %.0f | verify-only %.0f | both: write %.0f verify %.0f" % (
            (rate, med(L.GS_PROBE_WRITE)[0], med(L.GS_PROBE_VERIFY)[1]) + med(3)), flush=True)
    for wv in (0, 100, 110, 102):
        p.set_variant(write=wv)
        for rate in (0, 7450):
            p.set_write_rate(rate)
            print("write variant %3d rate %d: write-only %.0f" % (wv, rate, med(L.GS_PROBE_WRITE)[0]), flush=True)
    p.free(d)
mon.terminate()
lines = mon.stdout.read().strip().splitlines()
print("clock samples:", len(lines), "first", lines[:2], "last", lines[-3:])

"""Tuning aid: both-pass numbers over a grid of kernel variants (run under gpurun)."""
import itertools
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

gib = float(sys.argv[1]) if len(sys.argv) > 1 else 8.0
nbytes = int(gib * (1 << 30)) & ~0xFFFFF
rows = []
with Probe(0) as p:
    d = p.alloc(nbytes)
    sl = [(d, nbytes, 5)]
    for _ in range(3):
        p.run(sl)
    wsets = [(0, 0)] + [(w, r) for w in (100, 102, 103, 110, 112) for r in (0, 7300, 7450, 7600)]
    vsets = [(1, 0), (1, 8), (21, 0), (11, 0), (0, 8), (2, 8), (4, 8)]
    for (w, r), (v, ctas) in itertools.product(wsets, vsets):
        if w == 0 and ctas == 0:
            pass
        p.set_variant(write=w, verify=v, ctas_per_sm=ctas if w < 100 or True else 0)
        p.set_write_rate(r)
        ws, vs = [], []
        for _ in range(5):
            res, tm = p.run(sl)
            assert res[0].clean
            ws.append(tm.write_ms)
            vs.append(tm.verify_ms)
        wm, vm = statistics.median(ws), statistics.median(vs)
        rows.append((2 * nbytes / (wm + vm) / 1e6, nbytes / wm / 1e6, nbytes / vm / 1e6, w, r, v, ctas))
    p.free(d)
rows.sort(reverse=True)
print("combined  write  verify | wvariant rate vvariant ctas   (%.1f GiB, both passes back to back)" % gib)
for r in rows[:25]:
    print("%8.0f %6.0f %7.0f | %8d %4d %8d %4d" % r)
print("...")
for r in rows[-5:]:
    print("%8.0f %6.0f %7.0f | %8d %4d %8d %4d" % r)

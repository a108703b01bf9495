"""Tuning aid: sustained both-pass throughput over verify variants x grid sizes (run under gpurun)."""
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

gib = float(sys.argv[1]) if len(sys.argv) > 1 else 64.0
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 1.5
nslices = 8
per = (int(gib * (1 << 30)) // nslices) & ~0xFFFFF
rows = []
with Probe(0) as p:
    d = p.alloc(per * nslices)
    sl = [(d + k * per, per, 5 + k) for k in range(nslices)]
    for _ in range(3):
        p.run(sl)
    for wcode, rate in ((102, 7600), (122, 7600), (102, 7600), (122, 7600)):
        for v in (10, 13, 11):
            for ctas in (0,):
                p.set_variant(write=wcode, verify=v, ctas_per_sm=ctas)
                p.set_write_rate(rate)
                t0 = time.time()
                w, r = [], []
                while time.time() - t0 < secs:
                    res, tm = p.run(sl)
                    assert all(x.clean for x in res)
                    w.append(tm.write_ms)
                    r.append(tm.verify_ms)
                half = len(w) // 2
                wm, rm = statistics.median(w[half:]), statistics.median(r[half:])
                nb = per * nslices
                rows.append((2 * nb / (wm + rm) / 1e6, nb / wm / 1e6, nb / rm / 1e6, wcode, rate, v, ctas))
                print("%6.0f %6.0f %6.0f | w=%d rate=%d v=%d ctas=%d" % rows[-1], flush=True)
            pass
    p.free(d)
rows.sort(reverse=True)
print("TOP (sustained, second half of %.1f s each):" % secs)
for r in rows[:10]:
    print("%6.0f combined  write %6.0f  verify %6.0f | w=%d rate=%d v=%d ctas=%d" % r)

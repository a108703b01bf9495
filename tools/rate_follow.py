"""Tuning aid: the store pass's pacing target following the board (gs_probe_set_write_rate_auto).
From three starting targets, 70 store passes of 40 GiB each: target and achieved rate every tenth
pass, then the mean of the last 20 against a fixed 7 600 GB/s target on the same board."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

nbytes = 40 << 30
with Probe(0) as p:
    d = p.alloc(nbytes)
    sl = [(d, nbytes, 7)]
    p.run(sl)
    p.set_write_rate(7600)
    fixed = []
    for _ in range(30):
        _, tm = p.run(sl, L.GS_PROBE_WRITE)
        fixed.append(nbytes / tm.write_ms / 1e6)
    print("fixed 7600: mean of last 20 store passes %.0f GB/s" % (sum(fixed[-20:]) / 20), flush=True)
    for start in (7600, 8000, 7200):
        p.set_write_rate(start)
        p.set_write_rate_auto(True)
        rates, trail = [], []
        for k in range(70):
            _, tm = p.run(sl, L.GS_PROBE_WRITE)
            rates.append(nbytes / tm.write_ms / 1e6)
            if k % 10 == 9:
                trail.append("%d:%.0f" % (p.write_rate()[0], rates[-1]))
        print("start %d -> target:achieved every 10th pass  %s | mean of last 20: %.0f GB/s, target %d" % (
            start, "  ".join(trail), sum(rates[-20:]) / 20, p.write_rate()[0]), flush=True)
    p.free(d)

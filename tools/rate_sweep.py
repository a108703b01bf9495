"""Tuning aid: in-library write pass GB/s versus the pacing target (run under gpurun)."""
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

sizes = [float(x) for x in sys.argv[1:]] or [2.0, 8.0, 40.0]
with Probe(0) as p:
    for gib in sizes:
        nbytes = int(gib * (1 << 30)) & ~0xFFFFF
        d = p.alloc(nbytes)
        sl = [(d, nbytes, 99)]
        p.run(sl)
        line = []
        for rate in (0, 7000, 7200, 7300, 7400, 7450, 7500, 7600, 7800):
            p.set_write_rate(rate)
            ts = []
            for _ in range(5):
                _, tm = p.run(sl, L.GS_PROBE_WRITE)
                ts.append(tm.write_ms)
            line.append("%d:%.0f" % (rate, nbytes / statistics.median(ts) / 1e6))
        p.set_write_rate(7450)
        v, tm = p.run(sl)
        assert v[0].clean
        print("%.1f GiB write GB/s by target ->" % gib, " ".join(line),
              "| both passes: write %.0f verify %.0f combined %.0f" % (
                  nbytes / tm.write_ms / 1e6, nbytes / tm.verify_ms / 1e6, tm.gbps()), flush=True)
        p.free(d)

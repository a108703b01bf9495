"""Tuning aid: SUSTAINED store-pass and store+load rates versus the pacing target, 100 GiB slice,
~1.5 s per point (run under gpurun). The pattern change moved the store pass below the power cap,
so the target that was right for hashed data (7 600 GB/s) is re-examined here."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

nbytes = 100 << 30
with Probe(0) as p:
    d = p.alloc(nbytes)
    sl = [(d, nbytes, 99)]
    p.run(sl)
    for rate in (7400, 7600, 7700, 7800, 7900, 8000, 8200, 8600, 0):
        p.set_write_rate(rate)
        w, t0 = [], time.time()
        while time.time() - t0 < 1.5:
            _, tm = p.run(sl, L.GS_PROBE_WRITE)
            w.append(tm.write_ms)
        ww, vv, t0 = [], [], time.time()
        while time.time() - t0 < 1.5:
            v, tm = p.run(sl)
            assert v[0].clean
            ww.append(tm.write_ms)
            vv.append(tm.verify_ms)
        k = len(w) // 2
        print("target %4d | store alone %.0f GB/s | in pairs: store %.0f load %.0f combined %.0f" % (
            rate, nbytes / (sum(w[k:]) / len(w[k:])) / 1e6, nbytes / (sum(ww) / len(ww)) / 1e6,
            nbytes / (sum(vv) / len(vv)) / 1e6, 2 * nbytes / ((sum(ww) + sum(vv)) / len(ww)) / 1e6), flush=True)
    p.free(d)

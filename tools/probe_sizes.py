"""Probe rates against slice size, one slice, both passes back to back (burst: 5 runs each).

    python tools/probe_sizes.py > profiles/r01_probe_sizes.json
"""
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

SCRUB = len(sys.argv) > 1 and sys.argv[1] == "scrub"
PASSES = L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY | (L.GS_PROBE_SCRUB if SCRUB else 0)

MiB, GiB = 1 << 20, 1 << 30
rows = []
with Probe(0) as p:
    for nbytes in (MiB, 16 * MiB, 64 * MiB, 256 * MiB, GiB, 2 * GiB, 8 * GiB, 40 * GiB, 100 * GiB):
        d = p.alloc(nbytes)
        p.run([(d, nbytes, 1)], PASSES)
        w, r, sc = [], [], []
        for k in range(5):
            v, tm = p.run([(d, nbytes, 2 + k)], PASSES)
            sc.append(tm.scrub_ms)
            assert v[0].clean
            w.append(tm.write_ms)
            r.append(tm.verify_ms)
        wm, rm, sm = statistics.median(w), statistics.median(r), statistics.median(sc)
        rows.append({"slice_bytes": nbytes, "write_ms": wm, "verify_ms": rm, "scrub_ms": sm,
                     "write_gbps": nbytes / wm / 1e6, "verify_gbps": nbytes / rm / 1e6,
                     "combined_gbps": 2 * nbytes / (wm + rm + sm) / 1e6,
                     "fits_in_l2": nbytes <= 126 * MiB})
        print("%8.0f MiB: write %6.0f  verify %6.0f  combined %6.0f GB/s  (scrub %.3f ms)" % (
            nbytes / MiB, rows[-1]["write_gbps"], rows[-1]["verify_gbps"], rows[-1]["combined_gbps"], sm),
            file=sys.stderr, flush=True)
        p.free(d)
print(json.dumps({"scrub": SCRUB, "what": "one slice, store pass then load pass, CUDA events, median of 5; slices that fit the "
                          "126 MB L2 are read back from it (DESIGN.md section 4)", "rows": rows}, indent=1))

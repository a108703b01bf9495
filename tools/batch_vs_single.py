"""Does a batch of many slices cost more than one slice of the same total? Sustained A/B.

    python tools/batch_vs_single.py [total GiB=16] [secs=1.5]

One arena, swept (store + load, sustained) as 1 slice, as 64 slices with 64 different seeds, and
as 64 slices that all share one seed; medians of the second half of each run."""
import json
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200.probe import Probe  # noqa: E402

gib = float(sys.argv[1]) if len(sys.argv) > 1 else 16.0
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 1.5
total = int(gib * (1 << 30)) & ~((64 << 20) - 1)
with Probe(0) as p:
    d = p.alloc(total)
    forms = {
        "1 slice": lambda: [(d, total, p.next_seed())],
        "64 slices, 64 seeds": lambda: [(d + k * (total // 64), total // 64, p.next_seed(), k) for k in range(64)],
        "64 slices, 1 seed": lambda: (lambda s: [(d + k * (total // 64), total // 64, s, k) for k in range(64)])(p.next_seed()),
        "8 slices, 8 seeds": lambda: [(d + k * (total // 8), total // 8, p.next_seed(), k) for k in range(8)],
    }
    for rep in range(2):
        for name, make in forms.items():
            t0 = time.time()
            rows = []
            while time.time() - t0 < secs:
                v, tm = p.run(make())
                assert all(x.clean for x in v)
                rows.append((total / tm.write_ms / 1e6, total / tm.verify_ms / 1e6, tm.gbps()))
            rows = rows[len(rows) // 2:]
            print(json.dumps({"total_gib": gib, "form": name, "rep": rep,
                              "write_gbps": statistics.median(r[0] for r in rows),
                              "verify_gbps": statistics.median(r[1] for r in rows),
                              "combined_gbps": statistics.median(r[2] for r in rows)}), flush=True)
    p.free(d)

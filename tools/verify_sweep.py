"""Tuning aid: every load+compare form of the probe, sustained, with power/clock samples.

    python tools/verify_sweep.py [GiB=64] [secs=1.5] [out.json]

For each verify variant code (LDG forms and the bulk-load ring, gs_probe_set_variant) the slices
are verified back to back for `secs` seconds (load pass alone), then the leading forms are run
with the store pass in front of every load pass (the product's real sequence), for both patterns.
Rates are medians over the second half of each run. Run under gpurun."""
import json
import os
import statistics
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

gib = float(sys.argv[1]) if len(sys.argv) > 1 else 64.0
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 1.5
out = sys.argv[3] if len(sys.argv) > 3 else None
Q = "clocks.sm,power.draw,clocks_event_reasons.sw_power_cap"
nslices = 8
per = (int(gib * (1 << 30)) // nslices) & ~0xFFFFF
nb = per * nslices


def sustained(fn):
    mon = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=" + Q, "--format=csv,noheader,nounits",
                            "-lms", "100"], stdout=subprocess.PIPE, text=True)
    t0 = time.time()
    rows = []
    while time.time() - t0 < secs:
        rows.append(fn())
    mon.terminate()
    smi = [l.split(", ") for l in mon.communicate()[0].strip().splitlines() if l.count(",") == 2]
    smi = smi[len(smi) // 3:] or [["0", "0", "?"]]
    return rows[len(rows) // 2:], {
        "sm_mhz": statistics.median(float(r[0]) for r in smi),
        "watts": statistics.median(float(r[1]) for r in smi),
        "power_cap_active": sum(r[2].startswith("Active") for r in smi) / len(smi)}


results = []
with Probe(0) as p:
    d = p.alloc(nb)
    for pattern, pname in ((L.GS_PROBE_PATTERN_LOWTOGGLE, "lowtoggle"), (L.GS_PROBE_PATTERN_HASH, "hash")):
        sl = [(d + k * per, per, 5 + k, k, pattern) for k in range(nslices)]
        for _ in range(2):
            v, _tm = p.run(sl)
            assert all(x.clean for x in v)
        forms = [(10, 0), (0, 0), (0, 4), (20, 2), (20, 3), (13, 0),
                 (500, 1), (501, 1), (502, 1), (510, 1), (511, 1), (512, 1), (520, 1), (521, 1), (522, 1),
                 (600, 1), (611, 1), (622, 1), (501, 2), (511, 2), (500, 2)]
        if pattern:
            forms = [(10, 0), (0, 4), (501, 1), (511, 1), (521, 1), (512, 1)]
        alone = []
        for code, ctas in forms:
            p.set_variant(verify=code, ctas_per_sm=ctas)

            def one():
                v, tm = p.run(sl, L.GS_PROBE_VERIFY)
                assert all(x.clean for x in v)
                return nb / tm.verify_ms / 1e6
            rates, smi = sustained(one)
            row = {"pattern": pname, "mode": "load pass alone", "verify": code, "ctas_per_sm": ctas,
                   "verify_gbps": statistics.median(rates), **smi}
            alone.append(row)
            results.append(row)
            print(json.dumps(row), flush=True)
        alone.sort(key=lambda r: -r["verify_gbps"])
        lead = [(r["verify"], r["ctas_per_sm"]) for r in alone[:3]]
        if (10, 0) not in lead:
            lead.append((10, 0))
        for code, ctas in lead:
            p.set_variant(verify=code, ctas_per_sm=ctas)

            def both():
                v, tm = p.run(sl)
                assert all(x.clean for x in v)
                return (nb / tm.write_ms / 1e6, nb / tm.verify_ms / 1e6, 2 * nb / (tm.write_ms + tm.verify_ms) / 1e6)
            rates, smi = sustained(both)
            row = {"pattern": pname, "mode": "store then load", "verify": code, "ctas_per_sm": ctas,
                   "write_gbps": statistics.median(r[0] for r in rates),
                   "verify_gbps": statistics.median(r[1] for r in rates),
                   "combined_gbps": statistics.median(r[2] for r in rates), **smi}
            results.append(row)
            print(json.dumps(row), flush=True)
    p.set_variant(verify=10, ctas_per_sm=0)
    p.free(d)
if out:
    with open(out, "w") as f:
        json.dump({"gib": gib, "secs": secs, "rows": results}, f, indent=1)

"""Tuning aid: sustained probe passes (seconds, not milliseconds) with power/clock samples."""
import os
import statistics
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

gib = float(sys.argv[1]) if len(sys.argv) > 1 else 100.0
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 3.0
nbytes = int(gib * (1 << 30)) & ~0xFFFFF
Q = "clocks.sm,clocks.mem,power.draw,temperature.gpu,clocks_event_reasons.sw_power_cap,clocks_event_reasons.sw_thermal_slowdown"


def sample(fn, label):
    mon = subprocess.Popen(["nvidia-smi", "--query-gpu=" + Q, "--format=csv,noheader,nounits", "-lms", "100"],
                           stdout=subprocess.PIPE, text=True)
    t0 = time.time()
    rates = []
    while time.time() - t0 < secs:
        rates.append(fn())
    mon.terminate()
    rows = [l.split(", ") for l in mon.communicate()[0].strip().splitlines()]
    rows = rows[len(rows) // 3:]
    sm = statistics.median(float(r[0]) for r in rows)
    pw = statistics.median(float(r[2]) for r in rows)
    tp = max(float(r[3]) for r in rows)
    cap = sum(r[4].startswith("Active") for r in rows) / max(len(rows), 1)
    print("%-34s first %.0f  median %.0f  last %.0f GB/s | sm %.0f MHz  %.0f W  %.0f C  power-cap active %.0f%% of samples" % (
        label, rates[0], statistics.median(rates), rates[-1], sm, pw, tp, 100 * cap), flush=True)


with Probe(0) as p:
    d = p.alloc(nbytes)
    sl = [(d, nbytes, 5)]
    p.run(sl)
    sample(lambda: nbytes / p.run(sl, L.GS_PROBE_VERIFY)[1].verify_ms / 1e6, "verify only (default)")
    sample(lambda: nbytes / p.run(sl, L.GS_PROBE_WRITE)[1].write_ms / 1e6, "write only (default bulk, paced)")
    p.set_write_rate(0)
    sample(lambda: nbytes / p.run(sl, L.GS_PROBE_WRITE)[1].write_ms / 1e6, "write only (bulk, unpaced)")
    p.set_write_rate(7600)
    p.set_variant(write=0)
    sample(lambda: nbytes / p.run(sl, L.GS_PROBE_WRITE)[1].write_ms / 1e6, "write only (STG)")
    p.set_variant(write=102)
    sample(lambda: p.run(sl)[1].gbps(), "both passes (default)")
    for v, c in ((1, 8), (21, 8), (11, 8), (0, 4), (0, 16)):
        p.set_variant(verify=v, ctas_per_sm=c)
        sample(lambda: nbytes / p.run(sl, L.GS_PROBE_VERIFY)[1].verify_ms / 1e6, "verify only v=%d ctas=%d" % (v, c))
    p.free(d)

"""Small fixed probe run for ncu captures: one slice, both passes, three times.

    python tools/probe_prof.py [GiB=4] [verify_code=-1] [ctas_per_sm=-1] [pattern=0]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200.probe import Probe  # noqa: E402

gib = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
verify = int(sys.argv[2]) if len(sys.argv) > 2 else -1
ctas = int(sys.argv[3]) if len(sys.argv) > 3 else -1
pattern = int(sys.argv[4]) if len(sys.argv) > 4 else 0
n = int(gib * (1 << 30)) & ~0xFFFFF
with Probe(0) as p:
    p.set_variant(verify=verify, ctas_per_sm=ctas)
    d = p.alloc(n)
    for _ in range(3):
        v, tm = p.run([(d, n, p.next_seed(), 0, pattern)])
        assert v[0].clean
    print("write %.0f GB/s verify %.0f GB/s combined %.0f GB/s" % (
        n / tm.write_ms / 1e6, n / tm.verify_ms / 1e6, tm.gbps()))
    p.free(d)

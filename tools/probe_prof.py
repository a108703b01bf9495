"""Small fixed probe run for ncu captures: one 4 GiB slice, both passes, three times."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200.probe import Probe  # noqa: E402

n = 4 << 30
with Probe(0) as p:
    d = p.alloc(n)
    for _ in range(3):
        v, tm = p.run([(d, n, 7)])
        assert v[0].clean
    print("write %.0f GB/s verify %.0f GB/s combined %.0f GB/s" % (
        n / tm.write_ms / 1e6, n / tm.verify_ms / 1e6, tm.gbps()))
    p.free(d)

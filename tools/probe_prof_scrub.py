"""ncu case for the L2 scrub: one 64 MiB slice (smaller than the L2), with and without the scrub."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

n = 64 << 20
with Probe(0) as p:
    d = p.alloc(n)
    both = L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY
    for passes in (both, both | L.GS_PROBE_SCRUB, both, both | L.GS_PROBE_SCRUB):
        v, tm = p.run([(d, n, 7)], passes)
        assert v[0].clean
    p.free(d)
print("ok")

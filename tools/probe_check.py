"""Quick GPU sanity check of the probe against the oracle pattern (run under gpurun)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

L.load(strict=False)
orc = C.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle",
                          "libgs_oracle.so"))
n = (1 << 20) + 48  # ragged: not a multiple of the 64 KiB tile
with Probe(0) as p:
    d = p.alloc(n)
    v, tm = p.run([(d, n, 77)])
    print("clean run:", v, tm)
    got = p.read(d, n)
    want = (C.c_uint32 * (n // 4))()
    orc.gs_oracle_pattern_fill(want, C.c_uint64(0), C.c_uint64(n // 16), C.c_uint32(77))
    assert got == bytes(want), "GPU pattern differs from oracle"
    p.inject(d, 4 * 1000, 0x00010100)
    p.inject(d, n - 4, 0x80000000)
    v, tm = p.run([(d, n, 77)], L.GS_PROBE_VERIFY)
    print("after inject:", v)
    assert v[0].mismatches == 2 and v[0].first_bad_byte == 4001 and v[0].bit_flips == 3, v
    p.free(d)
print("PROBE CHECK OK")

"""Small run that touches every probe kernel form, for compute-sanitizer (one tool per gpurun call).

    compute-sanitizer --tool memcheck python tools/sanitize_case.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpushare_b200 import _lib as L  # noqa: E402
from gpushare_b200.probe import Probe  # noqa: E402

MiB = 1 << 20
sizes = [2 * MiB + 48, 64 * 1024, 16, 3 * MiB]
lib = L.load()
with Probe(0) as p:
    arena = p.alloc(sum(sizes) + 256)
    slices, off = [], 0
    for k, n in enumerate(sizes):
        slices.append((arena + off, n, 40 + k, k))
        off += n
    for w in (102, 302, 110, 403, 406, 0, 21):
        for v in (10, 0, 21, 1, 3):
            p.set_variant(write=w, verify=v)
            res, _ = p.run(slices)
            assert all(x.clean for x in res), (w, v, res)
    p.set_variant(write=102, verify=10)
    p.inject(slices[0][0], 4096, 0x11)
    p.inject(slices[3][0], 3 * MiB - 4, 0x80000000)
    res, _ = p.run(slices, L.GS_PROBE_VERIFY)
    assert [x.mismatches for x in res] == [1, 0, 0, 1], res
    arr = (L.Slice * len(slices))(*[L.Slice(b, n, s, t) for b, n, s, t in slices])
    out = (L.SliceResult * len(slices))()
    assert lib.gs_probe_verify_noalloc(0, None, arr, len(slices), out) == 0
    assert [o.mismatches for o in out] == [1, 0, 0, 1]
    p.free(arena)
print("SANITIZE CASE OK")

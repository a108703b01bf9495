"""gs_probe_verify_noalloc: the read-back pass for callers that may not allocate.

Some users of this library cannot call cudaMalloc: the plugin's health checker on a GPU
whose memory is granted to tenants down to the last unit (an allocation would fail or
take what a tenant was promised), and callers that cannot afford the device-wide
synchronisation cudaMalloc/cudaFree imply. For them the verdict table is a fixed array
on the device. This file checks, through the C ABI, that the verdicts equal those of the
ordinary entry point (gs_probe_run with GS_PROBE_VERIFY) and those the CPU oracle
(oracle/pattern.c) computes from a read-back of the same memory, bit for bit -- the
probe is integer work, so the tolerance is zero -- on two calls in a row, the second of
which must not see anything the first left in the shared array.

Collected last on purpose (conftest keeps test_zz_* at the end).
"""

import ctypes

import pytest

from gpushare_b200 import _lib as L

pytestmark = pytest.mark.gpu

MiB = 1 << 20


def _noalloc(lib, slices):
    arr = (L.Slice * len(slices))(*[L.Slice(s[0], s[1], s[2], i, s[4] if len(s) > 4 else 0, 0)
                                    for i, s in enumerate(slices)])
    res = (L.SliceResult * len(slices))()
    rc = lib.gs_probe_verify_noalloc(0, None, arr, len(slices), res)
    return rc, [(r.mismatches, r.first_bad_byte, r.bit_flips) for r in res]


def _oracle_verdicts(oracle_lib, p, slices):
    """What the CPU restatement says about the bytes that are in the slices right now."""
    out = []
    for k, s in enumerate(slices):
        base, n, seed = s[0], s[1], s[2]
        host = bytearray(p.read(base, n))
        arr = (ctypes.c_uint32 * (n // 4)).from_buffer(host)
        mm, fb, fl = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
        oracle_lib.gs_oracle_pattern_check_p(arr, ctypes.c_uint64(0), ctypes.c_uint64(n // 16),
                                             ctypes.c_uint32(seed), ctypes.c_uint32(k % 2),
                                             ctypes.byref(mm), ctypes.byref(fb), ctypes.byref(fl))
        out.append((mm.value, fb.value, fl.value))
    return out


def test_noalloc_verdicts_equal_ordinary_entry_point_twice(oracle_lib):
    from gpushare_b200.probe import Probe
    lib = L.load()
    sizes = [3 * MiB + 48, 64 * 1024, 9 * MiB, 16, 2 * MiB]
    with Probe(0) as p:
        arena = p.alloc(sum(sizes))
        try:
            slices, off = [], 0
            for k, n in enumerate(sizes):
                slices.append((arena + off, n, 500 + k, k, k % 2))    # both patterns
                off += n
            p.run(slices, L.GS_PROBE_WRITE)

            # call 1: faults in slices 0 and 2
            p.inject(slices[0][0], 4 * 4096, 0x00F00001)
            p.inject(slices[2][0], 8 * MiB + 12, 0x80000000)
            want, _ = p.run(slices, L.GS_PROBE_VERIFY)
            want = [(v.mismatches, v.first_bad_byte, v.bit_flips) for v in want]
            rc, got = _noalloc(lib, slices)
            assert rc == L.GS_OK
            assert got == want
            assert got == _oracle_verdicts(oracle_lib, p, slices)
            assert [g[0] for g in got] == [1, 0, 1, 0, 0] and got[0][2] == 5

            # call 2: repair those, break slices 1, 3 and 4 -- rows 0 and 2 of the shared
            # array still hold call 1's counts and must read clean now
            p.inject(slices[0][0], 4 * 4096, 0x00F00001)
            p.inject(slices[2][0], 8 * MiB + 12, 0x80000000)
            p.inject(slices[1][0], 0, 0x1)
            p.inject(slices[3][0], 12, 0xFFFF0000)
            p.inject(slices[4][0], 2 * MiB - 4, 0x10)
            want, _ = p.run(slices, L.GS_PROBE_VERIFY)
            want = [(v.mismatches, v.first_bad_byte, v.bit_flips) for v in want]
            rc, got = _noalloc(lib, slices)
            assert rc == L.GS_OK
            assert got == want
            assert got == _oracle_verdicts(oracle_lib, p, slices)
            assert [g[0] for g in got] == [0, 1, 0, 1, 1]
            assert got[0] == (0, 2**64 - 1, 0) and got[3] == (1, 14, 16)

            # a shorter call afterwards only reports its own rows
            rc, got = _noalloc(lib, slices[:2])
            assert rc == L.GS_OK and got == want[:2]
        finally:
            p.free(arena)


def test_noalloc_rejects_more_slices_than_the_fixed_table_holds():
    from gpushare_b200.probe import Probe
    lib = L.load()
    with Probe(0) as p:
        arena = p.alloc(65 * 4096)
        try:
            slices = [(arena + 4096 * k, 4096, k) for k in range(65)]
            p.run(slices[:64], L.GS_PROBE_WRITE)
            rc, got = _noalloc(lib, slices[:64])            # exactly the maximum is fine
            assert rc == L.GS_OK and all(g[0] == 0 for g in got)
            rc, _ = _noalloc(lib, slices)
            assert rc == L.GS_E_LIMIT
            assert b"GS_PROBE_NOALLOC_MAX_SLICES" in lib.gs_cuda_last_error()
            arr = (L.Slice * 1)(L.Slice(arena + 8, 4096, 0, 0))
            res = (L.SliceResult * 1)()
            assert lib.gs_probe_verify_noalloc(0, None, arr, 1, res) == L.GS_E_INVAL
            assert lib.gs_probe_verify_noalloc(0, None, arr, 0, res) == L.GS_E_INVAL
        finally:
            p.free(arena)

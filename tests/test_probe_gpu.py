"""Slice probe (K1) on a real GPU, through the C ABI, against the CPU pattern oracle."""
import ctypes

import pytest

from gpushare_b200 import _lib as L

pytestmark = pytest.mark.gpu

MiB = 1 << 20


@pytest.fixture(scope="module")
def probe():
    from gpushare_b200.probe import Probe
    p = Probe(0)
    yield p
    p.close()


LOW, HASH = L.GS_PROBE_PATTERN_LOWTOGGLE, L.GS_PROBE_PATTERN_HASH
# every load+compare form: LDG (unroll x access form) and the bulk-load ring (hint, stage size,
# consumer warps); the default first
VERIFY_FORMS = [10, 0, 1, 2, 3, 4, 21, 500, 501, 502, 510, 511, 512, 520, 521, 522, 600, 611, 622]


def oracle_bytes(oracle_lib, nbytes, seed, first=0, pattern=LOW):
    buf = (ctypes.c_uint32 * (nbytes // 4))()
    oracle_lib.gs_oracle_pattern_fill_p(buf, ctypes.c_uint64(first), ctypes.c_uint64(nbytes // 16),
                                        ctypes.c_uint32(seed), ctypes.c_uint32(pattern))
    return bytes(buf)


def oracle_check(oracle_lib, host, seed, pattern=LOW, first=0):
    mm, fb, fl = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
    arr = (ctypes.c_uint32 * (len(host) // 4)).from_buffer(host)
    oracle_lib.gs_oracle_pattern_check_p(arr, ctypes.c_uint64(first), ctypes.c_uint64(len(host) // 16),
                                         ctypes.c_uint32(seed), ctypes.c_uint32(pattern),
                                         ctypes.byref(mm), ctypes.byref(fb), ctypes.byref(fl))
    return mm.value, fb.value, fl.value


@pytest.mark.parametrize("nbytes", [16, 4096, 64 * 1024, 64 * 1024 + 16, MiB - 16, 3 * MiB + 4080,
                                    37 * MiB])
@pytest.mark.parametrize("wvariant", [302, 202, 102, 100, 110, 314, 403, 400, 406, 0, 21])
def test_written_pattern_is_bit_exact(probe, oracle_lib, nbytes, wvariant):
    d = probe.alloc(nbytes + 256)
    try:
        probe.set_variant(write=wvariant)
        v, tm = probe.run([(d, nbytes, 0xC0FFEE)])
        assert v[0].clean and tm.launches == 2 and tm.bytes == nbytes
        assert probe.read(d, nbytes) == oracle_bytes(oracle_lib, nbytes, 0xC0FFEE)
    finally:
        probe.set_variant(write=102)
        probe.free(d)


@pytest.mark.parametrize("pattern", [LOW, HASH])
@pytest.mark.parametrize("wvariant", [102, 403, 0])
@pytest.mark.parametrize("nbytes", [16, 64 * 1024 + 16, 3 * MiB + 4080])
def test_both_patterns_are_bit_exact_in_one_batch(probe, oracle_lib, nbytes, wvariant, pattern):
    """The pattern is a property of the slice: a batch mixes both, each slice reads back as its
    own pattern, and verifying a slice under the other pattern reports every word."""
    other = HASH if pattern == LOW else LOW
    d = probe.alloc(2 * nbytes + 256)
    try:
        probe.set_variant(write=wvariant)
        a, b = (d, nbytes, 0xBEEF, 0, pattern), (d + nbytes, nbytes, 0xBEEF, 1, other)
        v, tm = probe.run([a, b])
        assert v[0].clean and v[1].clean and tm.launches == 2
        assert probe.read(d, nbytes) == oracle_bytes(oracle_lib, nbytes, 0xBEEF, pattern=pattern)
        assert probe.read(d + nbytes, nbytes) == oracle_bytes(oracle_lib, nbytes, 0xBEEF, pattern=other)
        v, _ = probe.run([(d, nbytes, 0xBEEF, 0, other)], L.GS_PROBE_VERIFY)
        assert v[0].mismatches == nbytes // 16 and v[0].first_bad_byte == 0
        with pytest.raises(L.GpushareError) as e:
            probe.run([(d, nbytes, 1, 0, 2)])
        assert e.value.status == L.GS_E_INVAL
    finally:
        probe.set_variant(write=102)
        probe.free(d)


@pytest.mark.parametrize("pattern", [LOW, HASH])
@pytest.mark.parametrize("vvariant", VERIFY_FORMS)
def test_verify_counts_match_oracle(probe, oracle_lib, vvariant, pattern):
    nbytes = 5 * MiB + 48
    d = probe.alloc(nbytes)
    try:
        probe.set_variant(verify=vvariant)
        probe.run([(d, nbytes, 9, 0, pattern)], L.GS_PROBE_WRITE)
        host = bytearray(oracle_bytes(oracle_lib, nbytes, 9, pattern=pattern))
        for off, mask in [(0, 0x1), (4 * 1000, 0x00010100), (64 * 1024 - 4, 0xFFFFFFFF),
                          (64 * 1024, 0x80000000), (nbytes - 4, 0x3), (nbytes - 16, 0x40),
                          (2 * MiB + 8, 0x8000)]:
            probe.inject(d, off, mask)
            word = int.from_bytes(host[off:off + 4], "little") ^ mask
            host[off:off + 4] = word.to_bytes(4, "little")
        assert probe.read(d, nbytes) == bytes(host)
        want = oracle_check(oracle_lib, host, 9, pattern)
        v, _ = probe.run([(d, nbytes, 9, 0, pattern)], L.GS_PROBE_VERIFY)
        assert (v[0].mismatches, v[0].first_bad_byte, v[0].bit_flips) == want
        assert want[0] == 6 and want[1] == 0
        if vvariant >= 500:                              # two CTAs per SM, half the ring each
            probe.set_variant(ctas_per_sm=2)
            v, _ = probe.run([(d, nbytes, 9, 0, pattern)], L.GS_PROBE_VERIFY)
            assert (v[0].mismatches, v[0].first_bad_byte, v[0].bit_flips) == want
    finally:
        probe.set_variant(verify=10, ctas_per_sm=0)
        probe.free(d)


@pytest.mark.parametrize("vvariant", [500, 511, 522, 620, 10])
def test_load_pass_on_ragged_batches(probe, oracle_lib, vvariant):
    """Ragged slices (sub-chunk, sub-tile, one word, tile + one word) in one batch: a fault in the
    very last word and in the first word of each is found at its byte, by every load form."""
    sizes = [16, 48, 16 * 1024 - 16, 16 * 1024 + 16, 32 * 1024, 64 * 1024 - 16, 64 * 1024 + 16,
             5 * 64 * 1024 + 4096 + 16, 3 * MiB]
    arena = probe.alloc(sum(sizes) + 256)
    try:
        slices, off = [], 0
        for k, n in enumerate(sizes):
            slices.append((arena + off, n, 40 + k, k, k % 2))
            off += n
        probe.set_variant(verify=vvariant)
        v, _ = probe.run(slices)
        assert all(x.clean for x in v)
        for (base, n, seed, _k, pat) in slices:
            assert probe.read(base, n) == oracle_bytes(oracle_lib, n, seed, pattern=pat)
        for (base, n, _s, _k, _p) in slices:
            probe.inject(base, n - 4, 0x80000001)
        v, _ = probe.run(slices, L.GS_PROBE_VERIFY)
        assert [(x.mismatches, x.first_bad_byte, x.bit_flips) for x in v] == [(1, n - 4, 2) for n in sizes]
        for (base, n, _s, _k, _p) in slices:
            probe.inject(base, n - 4, 0x80000001)
            probe.inject(base, 0, 0x100)
        v, _ = probe.run(slices, L.GS_PROBE_VERIFY)
        assert [(x.mismatches, x.first_bad_byte, x.bit_flips) for x in v] == [(1, 1, 1)] * len(sizes)
    finally:
        probe.set_variant(verify=10)
        probe.free(arena)


def test_wrong_seed_is_all_mismatch(probe):
    nbytes = 2 * MiB
    d = probe.alloc(nbytes)
    try:
        probe.run([(d, nbytes, 1)], L.GS_PROBE_WRITE)
        v, _ = probe.run([(d, nbytes, 2)], L.GS_PROBE_VERIFY)
        assert v[0].mismatches == nbytes // 16 and v[0].first_bad_byte == 0
    finally:
        probe.free(d)


def test_batch_of_tenant_slices_is_independent(probe, oracle_lib):
    """BASELINE config 4 shape in miniature: many slices, one launch pair, per-slice verdicts."""
    sizes = [MiB, 3 * MiB + 16, 64 * 1024, 17 * MiB, 16, 5 * MiB, 2 * MiB + 4096, 9 * MiB]
    arena = probe.alloc(sum(sizes))
    try:
        slices, off = [], 0
        for k, n in enumerate(sizes):
            slices.append((arena + off, n, 100 + k, k))
            off += n
        v, tm = probe.run(slices)
        assert all(x.clean for x in v) and tm.launches == 2 and tm.bytes == sum(sizes)
        for (base, n, seed, _k) in slices:
            assert probe.read(base, n) == oracle_bytes(oracle_lib, n, seed)
        probe.inject(slices[3][0], 4 * MiB, 0xF0)       # inside slice 3 only
        probe.inject(slices[6][0], 0, 0x1)              # first word of slice 6
        v, _ = probe.run(slices, L.GS_PROBE_VERIFY)
        assert [x.mismatches for x in v] == [0, 0, 0, 1, 0, 0, 1, 0]
        assert v[3].first_bad_byte == 4 * MiB and v[3].bit_flips == 4
        assert v[6].first_bad_byte == 0 and v[6].bit_flips == 1
    finally:
        probe.free(arena)


def test_max_slices_and_argument_errors(probe):
    arena = probe.alloc(65 * 4096)
    try:
        sl = [(arena + 4096 * k, 4096, k) for k in range(65)]
        v, _ = probe.run(sl[:64])
        assert len(v) == 64 and all(x.clean for x in v)
        for bad, status in [(sl, L.GS_E_LIMIT), ([(arena + 8, 4096, 0)], L.GS_E_INVAL),
                            ([(arena, 24, 0)], L.GS_E_INVAL), ([(arena, 0, 0)], L.GS_E_INVAL),
                            ([(0, 4096, 0)], L.GS_E_INVAL)]:
            with pytest.raises(L.GpushareError) as e:
                probe.run(bad)
            assert e.value.status == status
        probe.submit(sl[:2])
        with pytest.raises(L.GpushareError) as e:
            probe.submit(sl[:2])                         # one batch in flight at a time
        assert e.value.status == L.GS_E_STATE
        probe.wait()
    finally:
        probe.free(arena)


def test_word_index_above_2_to_32(probe, oracle_lib):
    """Slices beyond 64 GiB carry a non-zero high index word; check the tail of a 65 GiB slice
    (whole-GPU health probes reach this). Skipped when the GPU lacks the room."""
    nbytes = 65 << 30
    free, _total = probe.mem_info()
    if free < nbytes + (1 << 30):
        pytest.skip("not enough free HBM")
    d = probe.alloc(nbytes)
    try:
        v, tm = probe.run([(d, nbytes, 31337)])
        assert v[0].clean
        tail = 4 * MiB
        first = (nbytes - tail) // 16
        assert first > 1 << 32
        assert probe.read(d + nbytes - tail, tail) == oracle_bytes(oracle_lib, tail, 31337, first)
        probe.inject(d, nbytes - 8, 0x2)
        v, _ = probe.run([(d, nbytes, 31337)], L.GS_PROBE_VERIFY)
        assert (v[0].mismatches, v[0].first_bad_byte) == (1, nbytes - 8)
    finally:
        probe.free(d)


def test_run_granted_and_round_trip_property(probe):
    """Full-size property (no oracle needed): write -> verify is clean for any seed, and a
    verify with any other seed reports every word."""
    from gpushare_b200.probe import run_granted
    v, tm = run_granted(0, 2 << 30, 77)
    assert v.clean and tm.bytes == 2 << 30 and tm.gbps() > 1000


def test_nvml_enumeration_and_ordinal_map():
    from gpushare_b200 import api
    with api.Nvml() as nv:
        infos = nv.devices()
        assert len(infos) >= 1 and infos[0].uuid.startswith(b"GPU-")
        assert infos[0].total_bytes >> 20 > 150000     # B200: 183 359 MiB
        ords = nv.map_cuda_ordinals(infos)
        assert sorted(o for o in ords if o >= 0) == list(range(len([o for o in ords if o >= 0])))
        t = api.DeviceTable.from_gpus(infos, "GiB")
        assert t.units_per_gpu == (infos[0].total_bytes >> 20) // 1024


def test_smoke_entry():
    import __graft_entry__
    __graft_entry__.smoke()


def test_real_nvml_health_watch_opens_and_times_out():
    """The XID watch against the real driver: register on every GPU, no event within 200 ms."""
    from gpushare_b200 import api
    with api.Nvml() as nv:
        mask = nv.health_open()
        assert mask == 0                                  # B200 supports XID events
        assert nv.health_wait(200) is None
        nv.health_close()


def test_health_time_sweep_of_free_memory(probe):
    v, tm = probe.sweep_free(margin_bytes=4 << 30)
    assert v is not None and v.clean
    assert tm.bytes > 100 << 30 and tm.gbps() > 5000       # most of a B200, at HBM speed


def test_full_size_properties_40gib(probe, oracle_lib):
    """BASELINE config 4's largest tenant (40 GiB): a sweep is clean; read back under another
    seed EVERY word differs, by exactly the bits in which the two sweeps' backgrounds differ (the
    address terms cancel); one flipped bit anywhere is found at its exact byte."""
    nbytes = 40 << 30
    free, _ = probe.mem_info()
    if free < nbytes + (2 << 30):
        pytest.skip("not enough free HBM")
    d = probe.alloc(nbytes)
    try:
        v, tm = probe.run([(d, nbytes, 0xA5A5)])
        assert v[0].clean and tm.bytes == nbytes
        w, _ = probe.run([(d, nbytes, 0x5A5A)], L.GS_PROBE_VERIFY)
        assert w[0].mismatches == nbytes // 16 and w[0].first_bad_byte == 0
        import ctypes
        a, b = (ctypes.c_uint32 * 4)(), (ctypes.c_uint32 * 4)()
        oracle_lib.gs_oracle_pattern_word(ctypes.c_uint64(0), ctypes.c_uint32(0xA5A5), a)
        oracle_lib.gs_oracle_pattern_word(ctypes.c_uint64(0), ctypes.c_uint32(0x5A5A), b)
        per_word = sum(bin(x ^ y).count("1") for x, y in zip(a, b))
        assert 32 < per_word < 96                      # two backgrounds: about half of 128 bits
        assert w[0].bit_flips == per_word * (nbytes // 16)
        off = (37 << 30) + 4 * 12345
        probe.inject(d, off, 1 << 9)
        v, _ = probe.run([(d, nbytes, 0xA5A5)], L.GS_PROBE_VERIFY)
        assert (v[0].mismatches, v[0].first_bad_byte, v[0].bit_flips) == (1, off + 1, 1)
    finally:
        probe.free(d)


@pytest.mark.parametrize("wvariant", [102, 302, 110, 202, 403, 400, 406, 0, 21])
@pytest.mark.parametrize("nbytes", [16, 48, 64 * 1024 - 16, 64 * 1024 + 16, 5 * 64 * 1024 + 4096 + 16])
def test_store_pass_stays_inside_the_granted_range(probe, wvariant, nbytes):
    """Bounds check of our own (compute-sanitizer is closed on this pool): guard bands written
    with one seed on both sides of a slice must read back clean after the slice itself is swept
    with every store form, whatever the slice's raggedness."""
    guard = 8192
    d = probe.alloc(2 * guard + nbytes + 256)
    try:
        left, mid, right = (d, guard, 0xAAAA), (d + guard, nbytes, 0x1234), (d + guard + nbytes, guard, 0xBBBB)
        probe.set_variant(write=102)
        probe.run([left, right], L.GS_PROBE_WRITE)
        probe.set_variant(write=wvariant)
        v, _ = probe.run([mid])
        assert v[0].clean
        g, _ = probe.run([left, right], L.GS_PROBE_VERIFY)
        assert g[0].clean and g[1].clean, (wvariant, nbytes, g)
    finally:
        probe.set_variant(write=102)
        probe.free(d)


def test_l2_scrub_between_the_passes(probe, oracle_lib):
    """GS_PROBE_SCRUB: same verdicts, a timed scrub kernel, and faults are still found."""
    nbytes = 48 * MiB                        # smaller than the L2: the case the scrub exists for
    d = probe.alloc(nbytes)
    try:
        both = L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY
        v, tm = probe.run([(d, nbytes, 3)], both | L.GS_PROBE_SCRUB)
        assert v[0].clean and tm.launches == 3 and tm.scrub_ms > 0 and tm.write_ms > 0 and tm.verify_ms > 0
        assert probe.read(d, MiB) == oracle_bytes(oracle_lib, MiB, 3)
        probe.inject(d, 40 * MiB + 8, 0x4)
        v, tm = probe.run([(d, nbytes, 3)], L.GS_PROBE_VERIFY | L.GS_PROBE_SCRUB)
        assert (v[0].mismatches, v[0].first_bad_byte, v[0].bit_flips) == (1, 40 * MiB + 8, 1)
        with pytest.raises(L.GpushareError):
            probe.run([(d, nbytes, 3)], L.GS_PROBE_SCRUB)          # a scrub alone is not a probe
    finally:
        probe.free(d)


def test_write_rate_follows_the_board(probe):
    """Opt-in following: the pacing target recovers from a start far above the write path's
    ceiling (where the store pass collapses to ~6.6 TB/s); by default and after an explicit
    set the target stays where it was put."""
    nbytes = 16 << 30
    free, _ = probe.mem_info()
    if free < nbytes + (2 << 30):
        pytest.skip("not enough free HBM")
    d = probe.alloc(nbytes)
    try:
        sl = [(d, nbytes, 5)]
        assert probe.write_rate() == (7600, False)           # the default is fixed
        probe.set_write_rate(8000)
        assert probe.write_rate() == (8000, False)
        _, tm = probe.run(sl, L.GS_PROBE_WRITE)
        collapsed = nbytes / tm.write_ms / 1e6
        assert probe.write_rate() == (8000, False)           # explicit: untouched by the pass
        probe.set_write_rate_auto(True)
        for _ in range(40):
            v, tm = probe.run(sl, L.GS_PROBE_WRITE)
        target, auto = probe.write_rate()
        settled = nbytes / tm.write_ms / 1e6
        assert auto and 7000 <= target <= 7600, target
        assert settled > collapsed + 300 and settled > 7000, (collapsed, settled, target)
        v, _ = probe.run(sl)
        assert v[0].clean
    finally:
        probe.set_write_rate(7600)                               # the default: fixed
        probe.free(d)


def test_failed_allocation_does_not_poison_the_next_sweep(probe):
    """An allocation that cannot be satisfied is an everyday event on a GPU shared with tenants.
    The failure must be reported once, by the call that failed, and not again by the launch
    check of the next sweep on the same thread (the CUDA runtime keeps a failed call's error
    until it is read)."""
    with pytest.raises(L.GpushareError) as e:
        probe.alloc(1 << 50)                              # a pebibyte: never available
    assert e.value.status == L.GS_E_CUDA
    d = probe.alloc(4 * MiB)
    try:
        v, tm = probe.run([(d, 4 * MiB, probe.next_seed())])
        assert v[0].clean and tm.launches == 2
        rc = probe.lib.gs_probe_verify_noalloc(99, None, probe._slices([(d, 4 * MiB, 1)]), 1,
                                               (L.SliceResult * 1)())   # no such device
        assert rc in (L.GS_E_CUDA, L.GS_E_INVAL)
        v, tm = probe.run([(d, 4 * MiB, probe.next_seed())])
        assert v[0].clean
    finally:
        probe.free(d)

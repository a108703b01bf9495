"""Slice probe (K1) -- thin object wrapper over the gs_probe_* C ABI.

No counterpart in the reference (it runs no GPU code); see include/gpushare.h.
"""
import ctypes as C
from dataclasses import dataclass

from . import _lib as L


@dataclass
class SliceVerdict:
    mismatches: int
    first_bad_byte: int  # 2**64-1 when clean
    bit_flips: int

    @property
    def clean(self):
        return self.mismatches == 0


@dataclass
class Timing:
    write_ms: float
    verify_ms: float
    bytes: int
    launches: int
    scrub_ms: float = 0.0

    def gbps(self):
        """2 algorithmic bytes per slice byte over both passes (SURVEY 8(d) metric 3); an L2
        scrub between them counts as time, not as bytes."""
        t = self.write_ms + self.verify_ms + self.scrub_ms
        n = (self.write_ms > 0) + (self.verify_ms > 0)
        return n * self.bytes / (t * 1e-3) / 1e9 if t > 0 else 0.0


class Probe:
    """One probe engine (stream, events, result table) bound to one CUDA ordinal."""

    def __init__(self, cuda_ordinal=0):
        self.lib = L.load()
        h = C.c_void_p()
        L.check(self.lib.gs_probe_open(cuda_ordinal, C.byref(h)), "gs_probe_open")
        self.h = h
        self.ordinal = cuda_ordinal
        self._n = 0

    def close(self):
        if self.h:
            self.lib.gs_probe_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def sm_count(self):
        v = C.c_int()
        L.check(self.lib.gs_probe_sm_count(self.h, C.byref(v)))
        return v.value

    def mem_info(self):
        f, t = C.c_uint64(), C.c_uint64()
        L.check(self.lib.gs_probe_mem_info(self.h, C.byref(f), C.byref(t)), "gs_probe_mem_info")
        return f.value, t.value

    def alloc(self, nbytes):
        p = C.c_void_p()
        L.check(self.lib.gs_probe_mem_alloc(self.h, nbytes, C.byref(p)), "gs_probe_mem_alloc")
        return p.value

    def free(self, dptr):
        L.check(self.lib.gs_probe_mem_free(self.h, C.c_void_p(dptr)), "gs_probe_mem_free")

    def next_seed(self):
        """A seed no earlier sweep of this engine used (gs_probe_next_seed)."""
        v = C.c_uint32()
        L.check(self.lib.gs_probe_next_seed(self.h, C.byref(v)), "gs_probe_next_seed")
        return v.value

    def test_skip_stores(self, on=True):
        """Test hook: the store pass launches nothing while on."""
        L.check(self.lib.gs_probe_test_skip_stores(self.h, int(on)), "gs_probe_test_skip_stores")

    def set_variant(self, write=-1, verify=-1, ctas_per_sm=-1):
        L.check(self.lib.gs_probe_set_variant(self.h, write, verify, ctas_per_sm),
                "gs_probe_set_variant")

    def set_write_rate(self, gbps):
        """Fix the store pass's pacing target (0 = unpaced); ends the per-board following."""
        L.check(self.lib.gs_probe_set_write_rate(self.h, gbps), "gs_probe_set_write_rate")

    def set_write_rate_auto(self, on=True):
        L.check(self.lib.gs_probe_set_write_rate_auto(self.h, int(on)), "gs_probe_set_write_rate_auto")

    def write_rate(self):
        """-> (current target in GB/s, whether it follows the board)."""
        g, a = C.c_uint32(), C.c_int()
        L.check(self.lib.gs_probe_get_write_rate(self.h, C.byref(g), C.byref(a)), "gs_probe_get_write_rate")
        return g.value, bool(a.value)

    @staticmethod
    def _slices(slices):
        arr = (L.Slice * len(slices))()
        for i, s in enumerate(slices):      # (base, bytes, seed[, tag[, pattern]])
            base, nbytes, seed = s[0], s[1], s[2]
            arr[i] = L.Slice(base, nbytes, seed, s[3] if len(s) > 3 else i,
                             s[4] if len(s) > 4 else L.GS_PROBE_PATTERN_LOWTOGGLE, 0)
        return arr

    def submit(self, slices, passes=L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY):
        arr = self._slices(slices)
        L.check(self.lib.gs_probe_submit(self.h, arr, len(slices), passes), "gs_probe_submit")
        self._n = len(slices)

    def wait(self):
        res = (L.SliceResult * self._n)()
        tm = L.ProbeTiming()
        L.check(self.lib.gs_probe_wait(self.h, res, self._n, C.byref(tm)), "gs_probe_wait")
        return ([SliceVerdict(r.mismatches, r.first_bad_byte, r.bit_flips) for r in res],
                Timing(tm.write_ms, tm.verify_ms, tm.bytes, tm.launches, tm.scrub_ms))

    def idle(self):
        return self.lib.gs_probe_poll(self.h) == L.GS_OK

    def run(self, slices, passes=L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY):
        self.submit(slices, passes)
        return self.wait()

    def sweep_free(self, margin_bytes=2 << 30, seed=None, pattern=L.GS_PROBE_PATTERN_HASH,
                   max_bytes=None):
        """Health-time probe: sweep what is free right now minus a margin, at most `max_bytes`
        (tenants' memory is theirs; only unallocated HBM can be written). Every sweep gets a
        seed of its own and, by default, the strong pattern."""
        free, _total = self.mem_info()
        nbytes = free - margin_bytes
        if max_bytes is not None:
            nbytes = min(nbytes, max_bytes)
        nbytes &= ~0xFFFFF
        if nbytes <= 0:
            return None, None
        d = self.alloc(nbytes)
        try:
            verdicts, tm = self.run([(d, nbytes, self.next_seed() if seed is None else seed, 0, pattern)])
        finally:
            self.free(d)
        return verdicts[0], tm

    def inject(self, base, byte_offset, mask):
        L.check(self.lib.gs_probe_inject(self.h, C.c_void_p(base), byte_offset, mask),
                "gs_probe_inject")

    def read(self, dptr, nbytes):
        buf = (C.c_uint8 * nbytes)()
        L.check(self.lib.gs_probe_read(self.h, C.c_void_p(dptr), buf, nbytes), "gs_probe_read")
        return bytes(buf)


def run_granted(cuda_ordinal, nbytes, seed):
    """Allocate exactly `nbytes` on the GPU, probe it, free it (gs_probe_run_granted)."""
    lib = L.load()
    res, tm = L.SliceResult(), L.ProbeTiming()
    L.check(lib.gs_probe_run_granted(cuda_ordinal, nbytes, seed, C.byref(res), C.byref(tm)),
            "gs_probe_run_granted")
    return (SliceVerdict(res.mismatches, res.first_bad_byte, res.bit_flips),
            Timing(tm.write_ms, tm.verify_ms, tm.bytes, tm.launches, tm.scrub_ms))


class ProbeScheduler:
    """Runs slice probes beside Allocate without ever making Allocate wait.

    Allocate's answer never depends on the probe (bit-exact parity with the reference,
    which has no probe), so grants are queued here and sent to the GPU whenever the
    engine is idle: one batch (<= GS_PROBE_MAX_SLICES slices) in flight per GPU. When
    grants arrive faster than HBM can be swept, the newest grants win and the dropped
    ones are counted -- coverage is reported, never faked. A verdict with mismatches is
    handed to `on_bad(tag, verdict)` (the daemon marks the GPU unhealthy there).

    Like the gs_probe underneath, a scheduler serves one caller at a time: own it from one
    thread (the daemon's probe thread does; Allocate only enqueues work for that thread).
    """

    def __init__(self, probe: Probe, on_bad=None, max_pending=L.GS_PROBE_MAX_SLICES,
                 scrub_below=512 << 20):
        self.probe = probe
        # a batch whose slices add up to less than this gets an L2 scrub between the passes:
        # without it the 126 MB L2 would serve (part of) the read-back instead of HBM. What
        # counts is the batch, not the slice: in a batch of many small slices the store pass
        # itself has displaced a slice's lines long before the load pass comes back to it.
        self.scrub_below = scrub_below
        self.on_bad = on_bad
        self.max_pending = max_pending
        self.pending = []
        self.inflight = None
        self.batches = 0
        self.launches = 0
        self.bytes_probed = 0
        self.bytes_granted = 0
        self.dropped = 0
        self.coalesced = 0
        self.write_ms = 0.0
        self.verify_ms = 0.0
        self.scrub_ms = 0.0
        self.bad = 0

    def grant(self, base, nbytes, seed=None, tag=0, pattern=L.GS_PROBE_PATTERN_LOWTOGGLE):
        """Queue a granted range. seed None (the deployed form): the sweep that finally covers
        the range draws a fresh seed from the engine, so no two sweeps of a range ever expect
        the same data. A fixed seed is for tests that read the pattern back."""
        self.bytes_granted += nbytes
        for i, g in enumerate(self.pending):
            if g[0] == base:                      # same range granted again before it was
                self.pending[i] = (base, nbytes, seed, tag, pattern)   # swept: one sweep covers both
                self.coalesced += 1
                return
        if len(self.pending) >= self.max_pending:
            self.pending.pop(0)
            self.dropped += 1
        self.pending.append((base, nbytes, seed, tag, pattern))

    def _collect(self):
        verdicts, tm = self.probe.wait()
        self.batches += 1
        self.launches += tm.launches
        self.bytes_probed += tm.bytes
        self.write_ms += tm.write_ms
        self.verify_ms += tm.verify_ms
        self.scrub_ms += tm.scrub_ms
        for s, v in zip(self.inflight, verdicts):
            if not v.clean:
                self.bad += 1
                if self.on_bad:
                    self.on_bad(s[3], v)
        self.inflight = None

    def pump(self):
        """Non-blocking: harvest a finished batch, start the next one."""
        if self.inflight is not None and self.probe.idle():
            self._collect()
        if self.inflight is None and self.pending:
            batch, self.pending = self.pending, []
            batch = [(b, n, self.probe.next_seed() if seed is None else seed, tag, pat)
                     for b, n, seed, tag, pat in batch]
            passes = L.GS_PROBE_WRITE | L.GS_PROBE_VERIFY
            if sum(s[1] for s in batch) < self.scrub_below:
                passes |= L.GS_PROBE_SCRUB
            self.probe.submit(batch, passes)
            self.inflight = batch

    def drain(self):
        """Blocking: finish what is in flight and what is queued."""
        while self.inflight is not None or self.pending:
            if self.inflight is not None:
                self._collect()
            self.pump()

    def gbps(self):
        t = self.write_ms + self.verify_ms + self.scrub_ms
        return 2 * self.bytes_probed / (t * 1e-3) / 1e9 if t > 0 else 0.0

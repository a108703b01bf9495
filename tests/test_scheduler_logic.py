"""ProbeScheduler's queueing policy, on CPU with a stand-in engine (the GPU leg is covered by
tests/test_scheduler_bench_gpu.py)."""
from gpushare_b200.probe import ProbeScheduler, SliceVerdict, Timing


class FakeEngine:
    def __init__(self):
        self.busy = False
        self.submitted = []
        self.bad_tags = set()

    def submit(self, slices, passes=3):
        assert not self.busy
        self.busy = True
        self.submitted.append(list(slices))
        self.passes = getattr(self, "passes", []) + [passes]

    def idle(self):
        return not self.busy or self.finish_now

    seeds = 1000

    def next_seed(self):
        self.seeds += 1
        return self.seeds

    finish_now = False

    def wait(self):
        batch = self.submitted[-1]
        self.busy = False
        nb = sum(s[1] for s in batch)
        verdicts = [SliceVerdict(1 if s[3] in self.bad_tags else 0, 0, 1) for s in batch]
        return verdicts, Timing(1.0, 1.0, nb, 2)


def test_grants_coalesce_while_engine_is_busy():
    e = FakeEngine()
    s = ProbeScheduler(e)
    s.grant(0x1000, 64, 1, 0)
    s.pump()                                  # first batch in flight
    assert len(e.submitted) == 1 and s.inflight is not None
    for wave in range(5):                     # the same two ranges granted again and again
        s.grant(0x1000, 64, 10 + wave, 0)
        s.grant(0x2000, 32, 20 + wave, 1)
        s.pump()                              # engine busy: nothing new goes out
    assert len(e.submitted) == 1 and len(s.pending) == 2 and s.coalesced == 8
    assert [g[2] for g in s.pending] == [14, 24]          # the newest seed of each range wins
    e.finish_now = True
    s.pump()                                  # harvest, then send the two coalesced ranges
    assert len(e.submitted) == 2 and [g[0] for g in e.submitted[1]] == [0x1000, 0x2000]
    s.drain()
    assert s.batches == 2 and s.launches == 4 and s.bytes_probed == 64 + 96
    assert s.bytes_granted == 64 + 5 * 96 and s.dropped == 0 and s.gbps() > 0


def test_queue_overflow_drops_oldest_and_counts():
    e = FakeEngine()
    s = ProbeScheduler(e, max_pending=4)
    s.grant(0, 16, 0, 0)
    s.pump()
    for k in range(1, 8):
        s.grant(0x100 * k, 16, k, k)
    assert len(s.pending) == 4 and s.dropped == 3
    assert [g[3] for g in s.pending] == [4, 5, 6, 7]
    e.finish_now = True
    s.drain()
    assert s.bytes_probed == 16 * 5


def test_bad_verdicts_reach_the_callback():
    e = FakeEngine()
    e.finish_now = True
    e.bad_tags = {7}
    seen = []
    s = ProbeScheduler(e, on_bad=lambda tag, v: seen.append((tag, v.mismatches)))
    s.grant(0, 16, 0, 3)
    s.grant(0x100, 16, 0, 7)
    s.drain()
    assert seen == [(7, 1)] and s.bad == 1


def test_small_batches_get_an_l2_scrub():
    e = FakeEngine()
    e.finish_now = True
    s = ProbeScheduler(e, scrub_below=1 << 20)
    s.grant(0, 4 << 20, 0, 0)
    s.drain()
    s.grant(0, 512 << 10, 0, 0)
    s.grant(0x1000000, 64 << 10, 0, 1)
    s.drain()
    s.grant(0, 4 << 20, 0, 0)            # a small slice in a batch that is large enough as a whole
    s.grant(0x1000000, 64 << 10, 0, 1)
    s.drain()
    assert e.passes == [3, 7, 3]         # write+verify; write+scrub+verify when the BATCH is below the bar


def test_every_sweep_gets_a_seed_of_its_own():
    """The deployed form of grant() passes no seed: each batch draws fresh ones from the engine,
    so sweeping the same range twice never expects the same data (a store pass that did nothing
    the second time would then be caught by the load pass)."""
    e = FakeEngine()
    e.finish_now = True
    s = ProbeScheduler(e)
    seen = []
    for _wave in range(4):
        s.grant(0x1000, 64, None, 0)
        s.grant(0x2000, 64, None, 1, 1)
        s.drain()
        seen += [sl[2] for sl in e.submitted[-1]]
        assert [sl[4] for sl in e.submitted[-1]] == [0, 1]       # the pattern travels with the grant
    assert len(set(seen)) == len(seen) == 8

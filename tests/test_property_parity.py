"""Property tests (hypothesis): the C ABI's Go-semantics helpers agree with the oracle on
arbitrary strings drawn from the characters those grammars care about."""
from hypothesis import given, settings
from hypothesis import strategies as st

from gpushare_b200 import api
from oracle import ref_model as R

settings.register_profile("ci", max_examples=400, deadline=None)
settings.load_profile("ci")


def both(c_fn, o_fn, s):
    try:
        want = o_fn(s)
    except ValueError:
        want = None
    try:
        got = c_fn(s)
    except api.L.GpushareError:
        got = None
    assert got == want, (s, got, want)


@given(st.text(alphabet="0123456789+-_ .xe", max_size=24))
def test_atoi(s):
    both(api.go_atoi, R.go_atoi, s)


@given(st.text(alphabet="0123456789+-_ e", max_size=24))
def test_parse_uint(s):
    both(api.go_parse_uint, R.go_parse_uint, s)


@given(st.text(alphabet="0123456789+-.eEinumkKMGTP ", min_size=1, max_size=16))
def test_quantity(s):
    both(api.quantity_value, R.quantity_value, s)


@given(st.integers(min_value=0, max_value=2**63 - 1), st.sampled_from(["", "m", "k", "Ki", "Mi", "e3", "E-2", "n"]))
def test_quantity_integers_with_suffix(n, suffix):
    both(api.quantity_value, R.quantity_value, "%d%s" % (n, suffix))


@given(st.lists(st.lists(st.binary(max_size=140), max_size=6), max_size=5))
def test_request_scan_counts_any_ids(reqs):
    wire = b"".join(R._ld(1, b"".join(R._ld(1, i) for i in c)) for c in reqs)
    counts, total = api.request_scan(wire)
    assert counts == [len(c) for c in reqs] and total == sum(len(c) for c in reqs)

"""upow_sweep_noalloc (the entry point for hosts that may not allocate GPU memory) against upow_sweep.

Collected last on purpose (file name): with `-x`, every other test has already run if this one fails.
It must pass both in a fresh process (`pytest tests/test_zz_no_alloc.py -m gpu`) and after the rest of
the suite: the static device scratch it uses is shared by all calls and keeps old contents, so the
second call in each sequence is the interesting one -- it must not see the first call's leftovers.
Comparison is bit-exact, like every other parity test here.
"""
import pytest

pytestmark = pytest.mark.gpu


def _t0(difficulty, version=1):
    from oracle import upow_oracle as O
    rule = O.Rule(difficulty, O.T0_PREV_HASH)
    return O.t0_prefix(difficulty, version), bytes.fromhex(O.T0_PREV_HASH), rule.int_difficulty, rule.charset_count


def test_noalloc_matches_sweep_on_two_calls_in_a_row(sweep_fixture):
    import upow_b200
    from upow_b200 import _lib
    from oracle import c_oracle
    from oracle import upow_oracle as O
    with upow_b200.Context(1) as ctx:
        # call 1: many hits (difficulty 3.0, 2^24 nonces -> ~4096); call 2: few hits, other layout, other range
        calls = [(_t0(3.0, 1), 3.0, 0, 1 << 24), (_t0(4.5, 2), 4.5, 1_000_003, 1_000_003 + (1 << 23)),
                 (_t0(5.0, 1), 5.0, 0, 1 << 28), (_t0(3.0, 1), 3.0, 0, 1 << 24)]
        for i, ((prefix, prev, n_int, cc), d, lo, hi) in enumerate(calls):
            ctx.set_template(0, prefix, prev, n_int, cc)
            want, _ = ctx.sweep(0, lo, hi)
            got = _lib.sweep_noalloc(0, prefix, prev, n_int, cc, lo, hi)
            assert got == want, f"call {i}"
            assert len(got) > 0
            if i == 2:
                # the reference's own answer for this call: the committed sweep fixture (hashlib fingerprint)
                assert prefix.hex() == sweep_fixture["prefix"]
                assert got == [n for n in sweep_fixture["nonces"] if n < (1 << 28)] and len(got) == 268
            else:
                assert got == c_oracle.all_valid_nonces(prefix, O.Rule(d, O.T0_PREV_HASH), lo, hi)[0], f"call {i} vs C oracle"


def test_noalloc_edges_and_limits():
    import upow_b200
    from upow_b200 import _lib
    prefix, prev, n_int, cc = _t0(5.0, 1)
    assert _lib.sweep_noalloc(0, prefix, prev, n_int, cc, 0, 0) == []                    # empty range
    assert _lib.sweep_noalloc(0, prefix, prev, n_int, cc, 164867, 164868) == [164867]    # SURVEY A.3 first nonce
    assert _lib.sweep_noalloc(0, prefix, prev, n_int, cc, 164868, 2752366) == []
    with pytest.raises(upow_b200.UpowError) as e:                                        # beyond the static scratch
        _lib.sweep_noalloc(0, prefix, prev, n_int, cc, 0, 1 << 20, cap=_lib.NOALLOC_MAX_HITS + 1)
    assert e.value.code == -1 and "UPOW_NOALLOC_MAX_HITS" in str(e.value)
    p1, pv1, n1, c1 = _t0(1.0, 1)
    with pytest.raises(upow_b200.UpowError) as e:                                        # more hits than the caller's cap
        _lib.sweep_noalloc(0, p1, pv1, n1, c1, 0, 1 << 16, cap=8)
    assert e.value.code == -3
    # and the scratch is usable again right after the overflow
    assert _lib.sweep_noalloc(0, prefix, prev, n_int, cc, 164867, 164868) == [164867]
    with pytest.raises(upow_b200.UpowError):
        _lib.sweep_noalloc(0, prefix[:-1], prev, n_int, cc, 0, 16)                        # bad prefix length

"""a3: fft / ifft / coset_fft / coset_ifft / lde_onto_coset on the device vs the oracle (bit-exact)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 0xFFFFFFFF00000001


@pytest.mark.parametrize("log_n", [0, 1, 2, 3, 5, 8, 10, 11, 12, 13, 16, 20, 21])
def test_fft_family(P2, O, rng, log_n):
    cols = 3 if log_n <= 16 else 1
    x = O.random_field(rng, (cols, 1 << log_n))
    assert np.array_equal(P2.fft(x), np.stack([O.fft(c) for c in x]))
    assert np.array_equal(P2.ifft(x), np.stack([O.ifft(c) for c in x]))
    assert np.array_equal(P2.coset_fft(x, 7), np.stack([O.coset_fft(c, 7) for c in x]))
    assert np.array_equal(P2.coset_ifft(x, 7), np.stack([O.coset_ifft(c, 7) for c in x]))
    s = int(O.random_field(rng, (1,))[0]) or 3
    assert np.array_equal(P2.coset_fft(x[0], s), O.coset_fft(x[0], s))
    assert np.array_equal(P2.coset_ifft(x[0], s), O.coset_ifft(x[0], s))


def test_fft_noncanonical_input(P2, O, rng):
    x = O.random_field(rng, (256,))
    y = x.copy()
    small = y < np.uint64(0xFFFFFFFF)
    y[small] = y[small] + np.uint64(P)  # same field elements, non-canonical representatives
    y[0] = np.uint64(P)
    x[0] = 0
    assert np.array_equal(P2.fft(y), O.fft(x))


def test_roundtrip_large(P2, O, rng):
    x = O.random_field(rng, (2, 1 << 18))
    assert np.array_equal(P2.ifft(P2.fft(x)), x)
    assert np.array_equal(P2.coset_ifft(P2.coset_fft(x, 7), 7), x)


@pytest.mark.parametrize("log_n,rate_bits", [(0, 1), (3, 1), (5, 3), (10, 1), (12, 2), (14, 3)])
def test_lde_onto_coset(P2, O, rng, log_n, rate_bits):
    v = O.random_field(rng, (2, 1 << log_n))
    got = P2.lde_onto_coset(v, rate_bits)
    assert np.array_equal(got, np.stack([O.lde_onto_coset(c, rate_bits) for c in v]))
    sel = P2.PolynomialValues.selector(1 << log_n, 0).lde_onto_coset(rate_bits)  # prover.rs:280-284
    assert np.array_equal(sel.values, O.lde_onto_coset(np.eye(1, 1 << log_n, 0, dtype=np.uint64)[0], rate_bits))


def test_many_caller_shifts_stay_within_the_table_budget(P2, O, rng):
    """nlc_coset_fft takes any shift; its power table is cached least-recently-used within a byte budget (1 GiB), so a
    caller that walks through 160 distinct shifts at 2^20 (8 MiB of table each, 1.25 GiB in all) makes the cache evict --
    and every transform, including one whose table was evicted and rebuilt, still equals the oracle's."""
    log_n = 20
    c = O.random_field(rng, 1 << log_n)
    shifts = [int(s) for s in rng.integers(2, 1 << 62, 160)]
    outs = {}
    for i, s in enumerate(shifts):
        v = P2.coset_fft(c, s)
        if i in (0, 1, 80, 159):
            outs[s] = v
    for s, v in outs.items():
        assert np.array_equal(v, O.coset_fft(c, s))
    assert np.array_equal(P2.coset_fft(c, shifts[0]), outs[shifts[0]])  # evicted by now: rebuilt, same answer
    assert np.array_equal(P2.coset_ifft(P2.coset_fft(c, shifts[5]), shifts[5]), c)

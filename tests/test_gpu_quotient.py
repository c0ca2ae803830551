"""the constraint-free parts of starkyx's quotient_polys on the device, vs the oracle (bit-exact).

Follows vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:278-302 (domain tables), 336-357 (1/Z_H, transpose,
coset_ifft) and 158-178 (trim_to_len, chunks, from_coeffs).
"""
import numpy as np
import pytest

from test_oracle import _quotient_case

pytestmark = pytest.mark.gpu
P = 0xFFFFFFFF00000001


@pytest.mark.parametrize("log_n,qdb", [(0, 0), (3, 0), (4, 1), (5, 2), (10, 1), (12, 3), (16, 1), (17, 2)])
def test_quotient_domain(P2, O, log_n, qdb):
    got = P2.quotient_domain(log_n, qdb)
    ref = O.quotient_domain(log_n, qdb)
    for g, r, name in zip(got, ref, ("coset", "lagrange_first", "lagrange_last", "z_h_inverses")):
        assert np.array_equal(g, r), name


@pytest.mark.parametrize("log_n,qdf,n_ch,r,cap", [(4, 1, 2, 1, 2), (5, 2, 2, 1, 2), (5, 3, 2, 2, 3), (6, 4, 3, 3, 4),
                                                  (4, 5, 1, 3, 0), (10, 2, 2, 1, 4), (11, 3, 2, 2, 4)])
def test_quotient_commit(P2, O, rng, log_n, qdf, n_ch, r, cap):
    qdb, q, q_vals, raw, zh = _quotient_case(O, rng, log_n, qdf, n_ch) if log_n <= 6 else _big_case(O, rng, log_n, qdf, n_ch)
    for divide, evals in ((True, raw), (False, np.ascontiguousarray(q_vals.T))):
        ref = O.quotient_commit(evals, log_n, qdb, qdf, divide, r, cap)
        assert ref is not None
        b = P2.PolynomialBatch.from_quotient_values(evals, log_n, qdf, r, cap, divide_by_zh=divide)
        assert b.n_polys == n_ch * qdf and b.degree_log == log_n
        assert np.array_equal(b.coeff_matrix, ref.coeffs)
        assert np.array_equal(b.coeff_matrix, q[:, :(1 << log_n) * qdf].reshape(n_ch * qdf, 1 << log_n))
        assert np.array_equal(b.merkle_tree.cap, ref.cap)
        assert np.array_equal(b.merkle_tree.digests, ref.digests)
        assert np.array_equal(b.merkle_tree.leaves, ref.leaves)


def _big_case(O, rng, log_n, qdf, n_ch):
    """Same construction as test_oracle._quotient_case with the Z_H product done through the periodic table."""
    qdb = max(qdf - 1, 0).bit_length()
    n, size = 1 << log_n, 1 << (log_n + qdb)
    q = np.zeros((n_ch, size), np.uint64)
    q[:, :n * qdf] = O.random_field(rng, (n_ch, n * qdf))
    q_vals = np.stack([O.coset_fft(c, 7) for c in q])
    zh = O.quotient_domain(log_n, qdb)[3]
    z_h = [pow(int(v), P - 2, P) for v in zh]
    raw = np.zeros((size, n_ch), np.uint64)
    for k in range(1 << qdb):
        for j in range(n_ch):
            col = q_vals[j, k::1 << qdb].astype(object) * z_h[k] % P
            raw[k::1 << qdb, j] = col.astype(np.uint64)
    return qdb, q, q_vals, raw, zh


def test_quotient_commit_noncanonical_rows(P2, O, rng):
    """Accumulators may arrive as any u64 representative (packed arithmetic does not canonicalise)."""
    qdb, q, q_vals, raw, zh = _quotient_case(O, rng, 5, 2, 2)
    bumped = raw.copy()
    small = bumped < np.uint64(0xFFFFFFFF)
    bumped[small] += np.uint64(P)
    ref = O.quotient_commit(raw, 5, qdb, 2, True, 1, 2)
    b = P2.PolynomialBatch.from_quotient_values(bumped, 5, 2, 1, 2, divide_by_zh=True)
    assert np.array_equal(b.merkle_tree.cap, ref.cap) and np.array_equal(b.coeff_matrix, ref.coeffs)


def test_quotient_failures_panic_like_the_reference(P2, O, rng):
    junk = O.random_field(rng, (1 << 7, 2))  # n = 2^5, qdf 3 -> size 2^7: does not divide
    assert O.quotient_commit(junk, 5, 2, 3, True, 2, 2) is None
    with pytest.raises(P2.NlcError, match="not divisible by Z_H"):
        P2.PolynomialBatch.from_quotient_values(junk, 5, 3, 2, 2, divide_by_zh=True)
    with pytest.raises(P2.NlcError, match="higher than the rate"):
        P2.PolynomialBatch.from_quotient_values(junk, 5, 3, 1, 2)
    with pytest.raises(P2.NlcError):
        P2.PolynomialBatch.from_quotient_values(junk[:100], 5, 3, 2, 2)
    # the context stays usable after a panic
    qdb, q, q_vals, raw, zh = _quotient_case(O, rng, 4, 2, 2)
    b = P2.PolynomialBatch.from_quotient_values(raw, 4, 2, 1, 2, divide_by_zh=True)
    assert np.array_equal(b.merkle_tree.cap, O.quotient_commit(raw, 4, qdb, 2, True, 1, 2).cap)

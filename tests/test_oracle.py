"""CPU suite (no GPU): the oracle against upstream's Poseidon known-answer vectors, the committed golden
fixtures, algebraic identities and its own restated FRI verifier."""
import hashlib
import os

import numpy as np
import pytest

P = 0xFFFFFFFF00000001
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_poseidon_kats_and_constants(O):
    assert O.poseidon_selftest() == 0
    rc = O.round_constants()
    assert [hex(int(x)) for x in rc[:4]] == ["0xb585f766f2144405", "0x7746a55f43921ad7", "0xb2fb0d31cee799b4",
                                             "0xf6760a4803427d7"]
    assert [hex(int(x)) for x in rc[356:]] == ["0x4543d9df5476d3cb", "0xf172d73e004fc90d", "0xdfd1c4febcc81238",
                                               "0xbc8dfb627fe558fc"]
    assert hashlib.sha256(rc.astype("<u8").tobytes()).hexdigest() == \
        "d2fcbb5be293c50ab4b1ddcd9c81005b12d689816a54c91a054f97f6588a20a8"
    g = np.load(os.path.join(GOLDEN, "poseidon.npz"))
    assert int(g["perm_zero"][0]) == 0x3c18a9786cb0b359 and int(g["perm_zero"][11]) == 0x1792b1c4342109d7
    assert int(g["perm_iota"][0]) == 0xd64e1e3efc5b8e9e and int(g["perm_neg1"][11]) == 0xe28e96f1ae5e60d3
    assert np.array_equal(O.poseidon_permute(np.zeros(12, np.uint64)), g["perm_zero"])
    assert np.array_equal(O.poseidon_permute(np.arange(12, dtype=np.uint64)), g["perm_iota"])
    assert np.array_equal(O.poseidon_permute(np.full(12, P - 1, np.uint64)), g["perm_neg1"])
    assert np.array_equal(O.hash_no_pad(np.arange(135, dtype=np.uint64)), g["hash135"])
    assert np.array_equal(O.two_to_one([1, 2, 3, 4], [5, 6, 7, 8]), g["two_to_one"])
    assert np.array_equal(rc, g["round_constants"])


def test_device_constants_match_oracle(O):
    """The header the kernels compile against holds the same 360 constants."""
    import re
    path = os.path.join(os.path.dirname(GOLDEN), "..", "near-light-client_b200", "csrc", "poseidon_constants.h")
    vals = [int(x, 16) for x in re.findall(r"0x([0-9a-f]{16})ULL", open(path).read())]
    assert vals == [int(x) for x in O.round_constants()]


def test_sponge_modes(O, rng):
    x = O.random_field(rng, (19,))
    # hash_no_pad == manual overwrite-mode sponge
    st = np.zeros(12, np.uint64)
    for off in range(0, 19, 8):
        chunk = x[off:off + 8]
        st[:chunk.size] = chunk
        st = O.poseidon_permute(st)
    assert np.array_equal(O.hash_no_pad(x), st[:4])
    assert np.array_equal(O.hash_or_noop(x[:3]), np.concatenate([x[:3], np.zeros(1, np.uint64)]))
    assert np.array_equal(O.hash_or_noop(x[:5]), O.hash_no_pad(x[:5]))
    s = np.zeros(12, np.uint64)
    s[:8] = x[:8]
    assert np.array_equal(O.two_to_one(x[:4], x[4:8]), O.poseidon_permute(s)[:4])


def naive_dft(c, shift=1):
    n = len(c)
    w = pow(1753635133440165772, (1 << 32) // n, P)
    return [sum(int(c[j]) * pow(shift, j, P) * pow(w, i * j, P) for j in range(n)) % P for i in range(n)]


def test_fft_conventions(O, rng):
    for log_n in range(0, 6):
        c = O.random_field(rng, (1 << log_n,))
        assert [int(v) for v in O.fft(c)] == naive_dft(c)
        assert [int(v) for v in O.coset_fft(c, 7)] == naive_dft(c, 7)
        assert np.array_equal(O.ifft(O.fft(c)), c)
        assert np.array_equal(O.coset_ifft(O.coset_fft(c, 7), 7), c)
    v = O.random_field(rng, (16,))
    lde = O.lde_onto_coset(v, 2)
    coeffs = O.ifft(v)
    padded = np.concatenate([coeffs, np.zeros(48, np.uint64)])
    assert [int(x) for x in lde] == naive_dft(padded, 7)


def test_merkle_layout_and_proofs(O, rng):
    for log_n, cap, ll in [(0, 0, 3), (3, 0, 5), (4, 2, 9), (5, 5, 2), (6, 4, 135)]:
        leaves = O.random_field(rng, (1 << log_n, ll))
        digests, capd = O.merkle_new(leaves, cap)
        assert digests.shape == (2 * ((1 << log_n) - (1 << cap)), 4)
        # root of each cap subtree recomputed naively
        per = (1 << log_n) >> cap
        for s in range(1 << cap):
            layer = [O.hash_or_noop(r) for r in leaves[s * per:(s + 1) * per]]
            while len(layer) > 1:
                layer = [O.two_to_one(layer[i], layer[i + 1]) for i in range(0, len(layer), 2)]
            assert np.array_equal(layer[0], capd[s])
        for idx in range(0, 1 << log_n, max(1, (1 << log_n) // 7)):
            sib = O.merkle_prove(digests, 1 << log_n, cap, idx)
            assert O.merkle_verify(leaves[idx], idx, capd, sib)
            if sib.size:
                bad = sib.copy()
                bad[0, 0] ^= np.uint64(1)
                assert not O.merkle_verify(leaves[idx], idx, capd, bad)


def test_commit_matches_definition_and_golden(O, rng):
    W, log_n, r, cap = 3, 4, 2, 1
    v = O.random_field(rng, (W, 1 << log_n))
    b = O.commit_from_values(v, r, cap)
    N, log_N = 1 << (log_n + r), log_n + r
    br = lambda x: int(format(x, f"0{log_N}b")[::-1], 2)
    for j in range(W):
        assert np.array_equal(b.coeffs[j], O.ifft(v[j]))
        lde = O.lde_onto_coset(v[j], r)
        assert np.array_equal(b.leaves[:, j], lde[[br(k) for k in range(N)]])
    d, c = O.merkle_new(b.leaves, cap)
    assert np.array_equal(d, b.digests) and np.array_equal(c, b.cap)
    # LDE restricted to every 2^r-th point of coset shift*<w_N> is NOT the trace (shift != 1), but the
    # polynomial it interpolates is: evaluate coefficients at w_n^i and compare with the trace values
    assert np.array_equal(O.fft(b.coeffs[0]), v[0])
    g = np.load(os.path.join(GOLDEN, "commit_small.npz"))
    for key in sorted({k.split(":")[0] for k in g.files}):
        Wg, lg, rg, cg = (int(x) for x in key.split("_"))
        bg = O.commit_from_values(g[key + ":values"], rg, cg)
        assert np.array_equal(bg.cap, g[key + ":cap"])
        assert np.array_equal(bg.coeffs[:, :4], g[key + ":coeffs_head"])


def test_challenger_semantics(O):
    ch = O.Challenger()
    ch.observe_elements(np.arange(1, 4, dtype=np.uint64))
    st = np.zeros(12, np.uint64)
    st[:3] = [1, 2, 3]
    st = O.poseidon_permute(st)
    assert ch.get_challenge() == int(st[7])       # pops from the END of the rate-8 output buffer
    assert ch.get_challenge() == int(st[6])
    ch.observe_elements(np.array([9], np.uint64))  # invalidates buffered outputs
    st[0] = 9
    st = O.poseidon_permute(st)
    assert ch.get_extension_challenge() == (int(st[7]), int(st[6]))


def test_fri_params(O):
    assert O.fri_params(16, 1, 4, 16, 84).reduction_arity_bits == [4, 4, 4]
    assert O.fri_params(16, 3, 4, 16, 28).reduction_arity_bits == [4, 4, 4]
    assert O.fri_params(5, 1, 4, 16, 84).reduction_arity_bits == []
    assert O.fri_params(20, 1, 4, 16, 84).reduction_arity_bits == [4, 4, 4, 4]


@pytest.mark.parametrize("log_n,r,cap,pow_bits,nq,arity", [(6, 1, 2, 7, 6, (2, 3)), (8, 3, 4, 8, 5, (4, 5)),
                                                          (5, 1, 0, 0, 3, (1, 2))])
def test_fri_self_verification_and_golden(O, rng, log_n, r, cap, pow_bits, nq, arity):
    n = 1 << log_n
    tv, qc = O.random_field(rng, (4, n)), O.random_field(rng, (2, n))
    t, q = O.commit_from_values(tv, r, cap), O.commit_from_coeffs(qc, r, cap)
    zeta = tuple(int(x) for x in O.random_field(rng, (2,)))
    g = pow(1753635133440165772, 1 << (32 - log_n), P)
    zn = (zeta[0] * g % P, zeta[1] * g % P)
    batches = [(zeta, [(0, i) for i in range(4)] + [(1, i) for i in range(2)]), (zn, [(0, i) for i in range(4)])]
    params = O.fri_params(log_n, r, cap, pow_bits, nq, *arity)
    ch = O.Challenger()
    ch.observe_cap(t.cap)
    vch = ch.clone()
    proof = O.prove_openings(O.FriInstance(batches), [t, q], ch, params)
    openings = [O.eval_poly_ext((t, q)[o].coeffs[i], np.array(pt, np.uint64)) for pt, polys in batches for o, i in polys]
    assert O.fri_verify(O.FriInstance(batches), openings, [t, q], vch.clone(), params, proof) == 0
    wrong = list(openings)
    wrong[0] = ((wrong[0][0] + 1) % P, wrong[0][1])
    assert O.fri_verify(O.FriInstance(batches), wrong, [t, q], vch.clone(), params, proof) != 0
    for pos in [0, proof.size // 2, proof.size - 1]:
        bad = proof.copy()
        bad[pos] ^= np.uint64(1)
        assert O.fri_verify(O.FriInstance(batches), openings, [t, q], vch.clone(), params, bad) != 0


def test_fri_golden(O):
    g = np.load(os.path.join(GOLDEN, "fri_small.npz"))
    log_n, r, cap = 6, 1, 2
    t, q = O.commit_from_values(g["trace_values"], r, cap), O.commit_from_coeffs(g["quotient_coeffs"], r, cap)
    zeta = tuple(int(x) for x in g["zeta"])
    gen = pow(1753635133440165772, 1 << (32 - log_n), P)
    zn = (zeta[0] * gen % P, zeta[1] * gen % P)
    batches = [(zeta, [(0, i) for i in range(5)] + [(1, i) for i in range(2)]), (zn, [(0, i) for i in range(5)])]
    ch = O.Challenger()
    ch.observe_cap(t.cap)
    ch.observe_cap(q.cap)
    proof = O.prove_openings(O.FriInstance(batches), [t, q], ch, O.fri_params(log_n, r, cap, 7, 6, 2, 3))
    assert np.array_equal(proof, g["proof"])
    assert np.array_equal(np.array(list(ch.c.state), np.uint64), g["challenger_state"])


def _quotient_case(O, rng, log_n, qdf, n_ch):
    """Random quotient polynomials of degree < n*qdf and the raw constraint accumulators c(x) = q(x) Z_H(x)
    on the quotient coset (what the AIR evaluation of prover.rs:320-334 would have produced)."""
    qdb = max(qdf - 1, 0).bit_length()
    n, size = 1 << log_n, 1 << (log_n + qdb)
    q = np.zeros((n_ch, size), np.uint64)
    q[:, :n * qdf] = O.random_field(rng, (n_ch, n * qdf))
    q_vals = np.stack([O.coset_fft(c, 7) for c in q])
    x, _, _, zh = O.quotient_domain(log_n, qdb)
    z_h = np.array([(pow(int(v), n, P) - 1) % P for v in x], dtype=object)
    raw = np.array([[int(q_vals[j, i]) * int(z_h[i]) % P for j in range(n_ch)] for i in range(size)], dtype=np.uint64)
    return qdb, q, q_vals, raw, zh


def test_quotient_domain_tables(O):
    """oracle/quotient.c against the closed forms: x_i = 7 w^i, L_first(x) = (x^n - 1)/(n (x - 1)),
    L_last(x) = w_n^-1 (x^n - 1)/(n (x - w_n^-1)), 1/Z_H(x_i) periodic with period 2^qdb."""
    for log_n, qdb in [(3, 0), (4, 1), (5, 2), (3, 3)]:
        n, size = 1 << log_n, 1 << (log_n + qdb)
        x, lf, ll, zh = O.quotient_domain(log_n, qdb)
        w = pow(7, (P - 1) >> (log_n + qdb), P)
        last = pow(pow(7, (P - 1) >> log_n, P), P - 2, P) if log_n else 1
        inv_n = pow(n, P - 2, P)
        for i in range(size):
            xi = 7 * pow(w, i, P) % P
            assert int(x[i]) == xi
            zhx = (pow(xi, n, P) - 1) % P
            assert int(zh[i % (1 << qdb)]) * zhx % P == 1
            assert int(lf[i]) == zhx * inv_n % P * pow((xi - 1) % P, P - 2, P) % P
            assert int(ll[i]) == last * zhx % P * inv_n % P * pow((xi - last) % P, P - 2, P) % P


@pytest.mark.parametrize("log_n,qdf,n_ch,r", [(4, 1, 2, 1), (5, 2, 2, 1), (5, 3, 2, 2), (6, 4, 3, 3), (4, 5, 1, 3)])
def test_quotient_commit_recovers_the_quotient(O, rng, log_n, qdf, n_ch, r):
    qdb, q, q_vals, raw, zh = _quotient_case(O, rng, log_n, qdf, n_ch)
    n = 1 << log_n
    want = q[:, :n * qdf].reshape(n_ch * qdf, n)  # polynomial j, chunk c -> column j*qdf + c
    ref = O.commit_from_coeffs(want, r, 2)
    for divide, evals in ((True, raw), (False, np.ascontiguousarray(q_vals.T))):
        b = O.quotient_commit(evals, log_n, qdb, qdf, divide, r, 2)
        assert np.array_equal(b.coeffs, want)
        assert np.array_equal(b.cap, ref.cap) and np.array_equal(b.digests, ref.digests)
        assert np.array_equal(b.leaves, ref.leaves)
    if qdf < (1 << qdb):  # trim_to_len fails on anything that does not divide (prover.rs:161-165)
        assert O.quotient_commit(O.random_field(rng, raw.shape), log_n, qdb, qdf, True, r, 2) is None

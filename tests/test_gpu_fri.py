"""a7/a8/a9: PolynomialBatch::prove_openings and the FRI prover on the device vs the oracle (bit-exact),
plus the restated verifier (fri/verifier.rs) accepting the device's proofs.

Shaped like vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:112-225: trace commit, quotient commit,
openings at zeta and g*zeta, prove_openings with the running challenger.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 0xFFFFFFFF00000001


def stark_shaped(P2, O, rng, log_n, W_trace, W_quot, cfg, arity=(4, 5)):
    """Returns (device batches, oracle batches, device instance, oracle instance, params pair, openings)."""
    n = 1 << log_n
    tv = O.random_field(rng, (W_trace, n))
    qc = O.random_field(rng, (W_quot, n))
    r, cap = cfg["rate_bits"], cfg["cap_height"]
    trace = P2.PolynomialBatch.from_values([P2.PolynomialValues(v) for v in tv], r, False, cap)
    quot = P2.PolynomialBatch.from_coeffs([P2.PolynomialCoeffs(c) for c in qc], r, False, cap)
    otrace, oquot = O.commit_from_values(tv, r, cap), O.commit_from_coeffs(qc, r, cap)
    zeta = tuple(int(x) for x in O.random_field(rng, (2,)))
    g = pow(1753635133440165772, 1 << (32 - log_n), P)  # primitive_root_of_unity(log_n)
    zeta_next = ((zeta[0] * g) % P, (zeta[1] * g) % P)
    all_polys = [(0, i) for i in range(W_trace)] + [(1, i) for i in range(W_quot)]
    next_polys = [(0, i) for i in range(W_trace)]
    obatches = [(zeta, all_polys), (zeta_next, next_polys)]
    inst = P2.FriInstanceInfo([], [P2.FriBatchInfo(pt, [P2.FriPolynomialInfo(o, i) for o, i in polys])
                                   for pt, polys in obatches])
    oinst = O.FriInstance(obatches)
    fc = P2.FriConfig(r, cap, cfg["pow_bits"], P2.FriReductionStrategy.ConstantArityBits(*arity), cfg["queries"])
    params = fc.fri_params(log_n)
    oparams = O.fri_params(log_n, r, cap, cfg["pow_bits"], cfg["queries"], *arity)
    assert params.reduction_arity_bits == oparams.reduction_arity_bits
    openings = []
    for pt, polys in obatches:
        for o, i in polys:
            coeffs = (otrace, oquot)[o].coeffs[i]
            openings.append(O.eval_poly_ext(coeffs, np.array(pt, np.uint64)))
    return (trace, quot), (otrace, oquot), inst, oinst, params, oparams, openings


def transcript_prefix(ch, caps):
    for c in caps:
        ch.observe_cap(c)
    return ch


CASES = [
    (5, 3, 2, dict(rate_bits=1, cap_height=0, pow_bits=4, queries=3), (2, 2)),
    (6, 2, 1, dict(rate_bits=2, cap_height=1, pow_bits=6, queries=4), (1, 3)),
    (8, 9, 2, dict(rate_bits=1, cap_height=4, pow_bits=8, queries=7), (4, 5)),
    (10, 20, 4, dict(rate_bits=3, cap_height=4, pow_bits=10, queries=5), (4, 5)),
    (12, 33, 2, dict(rate_bits=1, cap_height=4, pow_bits=16, queries=84), (4, 5)),   # starky standard_fast_config
    (13, 7, 3, dict(rate_bits=3, cap_height=4, pow_bits=16, queries=28), (4, 5)),    # standard_recursion_config
    (3, 2, 1, dict(rate_bits=1, cap_height=2, pow_bits=0, queries=2), (4, 5)),       # no reduction rounds
]


@pytest.mark.parametrize("log_n,Wt,Wq,cfg,arity", CASES)
def test_prove_openings_bit_exact_and_verifies(P2, O, rng, log_n, Wt, Wq, cfg, arity):
    dev, orc, inst, oinst, params, oparams, openings = stark_shaped(P2, O, rng, log_n, Wt, Wq, cfg, arity)
    caps = [orc[0].cap, orc[1].cap]
    assert np.array_equal(dev[0].merkle_tree.cap, caps[0]) and np.array_equal(dev[1].merkle_tree.cap, caps[1])
    ch = transcript_prefix(P2.Challenger(), caps)
    och = transcript_prefix(O.Challenger(), caps)
    vch = och.clone()
    proof = P2.PolynomialBatch.prove_openings(inst, list(dev), ch, params)
    oproof = O.prove_openings(oinst, list(orc), och, oparams)
    assert proof.flat.size == oproof.size
    assert np.array_equal(proof.flat, oproof)
    assert ch.state() == och.state()  # the transcript leaves prove_openings in the same state
    # unchanged-verifier check (restated fri/verifier.rs) on the DEVICE proof
    assert O.fri_verify(oinst, openings, list(orc), vch, oparams, proof.flat) == 0
    # and it rejects a corrupted one
    bad = proof.flat.copy()
    bad[len(params.reduction_arity_bits) * (4 << cfg["cap_height"])] ^= np.uint64(1)
    if cfg["queries"]:
        assert O.fri_verify(oinst, openings, list(orc), transcript_prefix(O.Challenger(), caps), oparams, bad) != 0
    assert len(proof.query_round_proofs) == cfg["queries"]
    assert proof.final_poly.shape[0] == params.final_poly_len()


def test_rust_driven_commit_phase(P2, O, rng):
    """nlc_fri_begin / commit_round / fold_round / final_poly / pow / queries with the transcript on the caller's side."""
    cfg = dict(rate_bits=1, cap_height=3, pow_bits=9, queries=6)
    dev, orc, inst, oinst, params, oparams, _ = stark_shaped(P2, O, rng, 9, 5, 2, cfg)
    caps = [orc[0].cap, orc[1].cap]
    och = transcript_prefix(O.Challenger(), caps)
    oproof = O.prove_openings(oinst, list(orc), och, oparams)
    ch = transcript_prefix(P2.Challenger(), caps)
    alpha = ch.get_extension_challenge()
    phase = P2.FriCommitPhase(inst, list(dev), alpha, params)
    flat = []
    for _ in params.reduction_arity_bits:
        cap = phase.commit_round()
        flat.append(cap.reshape(-1))
        ch.observe_cap(cap)
        phase.fold_round(ch.get_extension_challenge())
    fp = phase.final_poly()
    flat.append(fp.reshape(-1))
    ch.observe_extension_elements(fp)
    w = P2.fri_proof_of_work(ch, cfg["pow_bits"])
    flat.append(np.array([w], np.uint64))
    ch.observe_element(w)
    resp = ch.get_challenge()
    assert resp >> (64 - cfg["pow_bits"]) == 0
    xs = [c % params.lde_size() for c in ch.get_n_challenges(cfg["queries"])]
    flat.append(phase.queries(xs).reshape(-1))
    assert np.array_equal(np.concatenate(flat), oproof)
    assert ch.state() == och.state()


def test_pow_is_minimal(P2, O, rng):
    ch = P2.Challenger()
    ch.observe_elements(O.random_field(rng, (11,)))
    for bits in [0, 1, 5, 12, 18]:
        w = P2.fri_proof_of_work(ch, bits)
        for cand in range(max(0, w - 48), w + 1):  # every probe is a device round trip: check the neighbourhood
            c = ch.clone()
            c.observe_element(cand)
            ok = bits == 0 or (c.get_challenge() >> (64 - bits)) == 0
            assert ok == (cand == w), (bits, cand, w)
        if bits <= 12:  # exhaustive minimality against the oracle's permutation for small targets
            st = list(ch.c.sponge_state)
            for i in range(ch.c.input_len):
                st[i] = ch.c.input_buffer[i]
            for cand in range(w):
                s2 = np.array(st, np.uint64)
                s2[ch.c.input_len] = cand
                assert bits and (int(O.poseidon_permute(s2)[7]) >> (64 - bits)) != 0


def test_prove_openings_errors(P2, pkg, O, rng):
    cfg = dict(rate_bits=1, cap_height=2, pow_bits=2, queries=2)
    dev, _, inst, _, params, _, _ = stark_shaped(P2, O, rng, 6, 3, 2, cfg)
    bad = P2.FriInstanceInfo([], [P2.FriBatchInfo((1, 2), [P2.FriPolynomialInfo(0, 99)])])
    with pytest.raises(pkg.NlcError, match="polynomial_index"):
        P2.PolynomialBatch.prove_openings(bad, list(dev), P2.Challenger(), params)
    wrong = P2.FriConfig(2, 2, 2, P2.FriReductionStrategy.ConstantArityBits(4, 5), 2).fri_params(6)
    with pytest.raises(pkg.NlcError, match="rate_bits"):
        P2.PolynomialBatch.prove_openings(inst, list(dev), P2.Challenger(), wrong)

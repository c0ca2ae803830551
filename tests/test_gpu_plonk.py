"""the plonk-side quotient path on the device (csrc/plonk.cu behind nlc_plonk_zs_partial_products_commit /
nlc_plonk_vanishing_eval / nlc_plonk_quotient_commit) against the oracle's direct evaluation (oracle/plonk.py, itself
pinned by the verifier's vanishing identity in tests/test_plonk_cpu.py): bit-exact on the Z / partial-product
polynomials, on the vanishing values of every point of the quotient domain, on the quotient commitment, and on a whole
`plonk::prover::prove` (caps, openings, FriProof), which the restated FRI verifier accepts."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 0xFFFFFFFF00000001
G32 = 1753635133440165772


def setup(P2, O, K, degree_bits, seed, gate_mix=None, cap=4, cfg=None):
    from oracle import plonk as OP
    cfg = cfg or K.CircuitConfig.standard_recursion_config()
    cfg.fri_config = P2.FriConfig(3, cap, 16, P2.FriReductionStrategy.ConstantArityBits(4, 5), 28)
    common, consts, sig, wires, pih = K.build_demo_circuit(cfg, degree_bits, seed, gate_mix)
    r = cfg.fri_config.rate_bits
    cs_vals = np.concatenate([consts, sig])
    gcs = P2.PolynomialBatch.from_values(list(cs_vals), r, False, cap)
    gw = P2.PolynomialBatch.from_values(list(wires), r, False, cap)
    rcs, rw = O.commit_from_values(cs_vals, r, cap), O.commit_from_values(wires, r, cap)
    pd = K.ProverOnlyCircuitData(gcs, sig)
    return OP, cfg, common, consts, sig, wires, pih, pd, gw, rcs, rw


@pytest.mark.parametrize("degree_bits,seed", [(4, 0), (6, 1), (10, 2), (13, 3)])
def test_partial_products_vanishing_and_quotient(P2, O, pkg, degree_bits, seed):
    K = pkg.plonk
    cap = min(4, degree_bits)
    OP, cfg, common, consts, sig, wires, pih, pd, gw, rcs, rw = setup(P2, O, K, degree_bits, seed, cap=cap)
    r = cfg.fri_config.rate_bits
    rng = np.random.default_rng(100 + seed)
    betas, gammas, alphas = ([int(x) for x in rng.integers(1, P, 2, dtype=np.uint64)] for _ in range(3))
    # all_wires_permutation_partial_products + from_values
    gz = K.PlonkProver.zs_partial_products_commitment(common, pd, gw, betas, gammas)
    vals = OP.partial_products_and_zs(common, wires, sig, betas, gammas)
    rz = O.commit_from_values(vals, r, cap)
    assert np.array_equal(gz.merkle_tree.cap, rz.cap)
    assert np.array_equal(gz.coeff_matrix, rz.coeffs)
    assert np.array_equal(P2.fft(gz.coeff_matrix[0]), vals[0]) and vals[0][0] == 1  # Z(1) = 1
    # eval_vanishing_poly_base_batch over the whole quotient domain (before 1/Z_H)
    got = K.PlonkProver.vanishing_values(common, pd, pih, gw, gz, betas, gammas, alphas)
    want = OP.vanishing_values(common, rcs, rw, rz, pih, betas, gammas, alphas, r, divide_by_zh=False)
    assert got.shape == want.shape and np.array_equal(got, want)
    # compute_quotient_polys + chunks + from_coeffs
    gq = K.PlonkProver.quotient_commitment(common, pd, pih, gw, gz, betas, gammas, alphas)
    rq = O.quotient_commit(OP.vanishing_values(common, rcs, rw, rz, pih, betas, gammas, alphas, r), degree_bits, 3, 8, False, r, cap)
    assert gq.n_polys == 16 and np.array_equal(gq.merkle_tree.cap, rq.cap) and np.array_equal(gq.coeff_matrix, rq.coeffs)
    assert np.array_equal(gq.merkle_tree.digests, rq.digests)
    # ... and what the device produced satisfies the verifier's identity at a random point
    zeta = int(rng.integers(2, P, dtype=np.uint64))
    class B:  # coefficient views of the GPU batches
        def __init__(self, b):
            self.coeffs = b.coeff_matrix
    assert OP.vanishing_identity_holds(common, B(pd.constants_sigmas_commitment), B(gw), B(gz), B(gq), pih, betas, gammas, alphas, zeta)


def test_ragged_last_chunk_and_three_challenges(P2, O, pkg):
    """A configuration whose routed wires do not fill the last partial-product chunk (37 = 4 x 8 + 5), three challenges,
    100 wires (no Poseidon gate): the chunk bounds of k_pp_rows / check_partial_products and num_challenges != 2."""
    K = pkg.plonk
    cfg = K.CircuitConfig(num_wires=100, num_routed_wires=37, num_constants=2, num_challenges=3, max_quotient_degree_factor=8)
    OP, cfg, common, consts, sig, wires, pih, pd, gw, rcs, rw = setup(P2, O, K, 8, 21, cfg=cfg)
    assert common.num_partial_products == 4
    r, cap = 3, 4
    rng = np.random.default_rng(33)
    betas, gammas, alphas = ([int(x) for x in rng.integers(1, P, 3, dtype=np.uint64)] for _ in range(3))
    gz = K.PlonkProver.zs_partial_products_commitment(common, pd, gw, betas, gammas)
    rz = O.commit_from_values(OP.partial_products_and_zs(common, wires, sig, betas, gammas), r, cap)
    assert gz.n_polys == 15 and np.array_equal(gz.coeff_matrix, rz.coeffs) and np.array_equal(gz.merkle_tree.cap, rz.cap)
    got = K.PlonkProver.vanishing_values(common, pd, pih, gw, gz, betas, gammas, alphas)
    assert np.array_equal(got, OP.vanishing_values(common, rcs, rw, rz, pih, betas, gammas, alphas, r, divide_by_zh=False))
    gq = K.PlonkProver.quotient_commitment(common, pd, pih, gw, gz, betas, gammas, alphas)
    rq = O.quotient_commit(OP.vanishing_values(common, rcs, rw, rz, pih, betas, gammas, alphas, r), 8, 3, 8, False, r, cap)
    assert gq.n_polys == 24 and np.array_equal(gq.coeff_matrix, rq.coeffs)
    class B:
        def __init__(self, b):
            self.coeffs = b.coeff_matrix
    zeta = int(rng.integers(2, P, dtype=np.uint64))
    assert OP.vanishing_identity_holds(common, B(pd.constants_sigmas_commitment), B(gw), B(gz), B(gq), pih, betas, gammas, alphas, zeta)


def test_many_gates_need_several_selectors(P2, O, pkg):
    K = pkg.plonk
    mix = [K.NoopGate(), K.PublicInputGate(), K.ConstantGate(1), K.ConstantGate(2), K.BaseSumGate(8), K.BaseSumGate(9),
           K.BaseSumGate(10), K.ArithmeticGate(20), K.ArithmeticGate(10), K.ArithmeticExtensionGate(10)]
    OP, cfg, common, consts, sig, wires, pih, pd, gw, rcs, rw = setup(P2, O, K, 7, 5, gate_mix=mix)
    assert common.selectors_info.num_selectors() > 1 and common.num_constants == common.selectors_info.num_selectors() + 2
    r, cap = 3, 4
    rng = np.random.default_rng(9)
    betas, gammas, alphas = ([int(x) for x in rng.integers(1, P, 2, dtype=np.uint64)] for _ in range(3))
    gz = K.PlonkProver.zs_partial_products_commitment(common, pd, gw, betas, gammas)
    rz = O.commit_from_values(OP.partial_products_and_zs(common, wires, sig, betas, gammas), r, cap)
    assert np.array_equal(gz.coeff_matrix, rz.coeffs)
    got = K.PlonkProver.vanishing_values(common, pd, pih, gw, gz, betas, gammas, alphas)
    assert np.array_equal(got, OP.vanishing_values(common, rcs, rw, rz, pih, betas, gammas, alphas, r, divide_by_zh=False))


def test_plonk_proof_end_to_end(P2, O, pkg):
    """plonk::prover::prove from the witness on (prover.rs) equals the same sequence driven through the oracle, and the
    restated FRI verifier accepts the opening proof."""
    K = pkg.plonk
    degree_bits = 9
    OP, cfg, common, consts, sig, wires, pih, pd, gw, rcs, rw = setup(P2, O, K, degree_bits, 11)
    fc = cfg.fri_config
    r, cap = fc.rate_bits, fc.cap_height
    proof = K.PlonkProver.prove(common, pd, list(wires), pih)
    ch = O.Challenger()
    ch.observe_elements(pd.circuit_digest)
    ch.observe_elements(pih)
    ch.observe_cap(rw.cap)
    betas, gammas = ch.get_n_challenges(2), ch.get_n_challenges(2)
    rz = O.commit_from_values(OP.partial_products_and_zs(common, wires, sig, betas, gammas), r, cap)
    ch.observe_cap(rz.cap)
    alphas = ch.get_n_challenges(2)
    rq = O.quotient_commit(OP.vanishing_values(common, rcs, rw, rz, pih, betas, gammas, alphas, r), degree_bits, 3, 8, False, r, cap)
    assert np.array_equal(proof.wires_cap, rw.cap) and np.array_equal(proof.plonk_zs_partial_products_cap, rz.cap)
    assert np.array_equal(proof.quotient_polys_cap, rq.cap)
    ch.observe_cap(rq.cap)
    zeta = ch.get_extension_challenge()
    g = pow(G32, 1 << (32 - degree_bits), P)
    zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
    ev = lambda b, z: [O.eval_poly_ext(c, np.array(z, np.uint64)) for c in b.coeffs]
    cs_z, w_z, zp_z, q_z, zp_n = ev(rcs, zeta), ev(rw, zeta), ev(rz, zeta), ev(rq, zeta), ev(rz, zeta_next)
    o = proof.openings
    assert o.constants + o.plonk_sigmas == cs_z and o.wires == w_z and o.quotient_polys == q_z
    assert o.plonk_zs + o.partial_products == zp_z and o.plonk_zs_next == zp_n[:2]
    zeta_batch = cs_z + w_z + zp_z[:2] + zp_z[2:] + q_z
    for batch in (zeta_batch, zp_n[:2]):
        ch.observe_elements([c for e in batch for c in e])
    ch_v = ch.clone()
    params = O.fri_params(degree_bits, r, cap, fc.proof_of_work_bits, fc.num_query_rounds, 4, 5)
    widths = [rcs.coeffs.shape[0], rw.coeffs.shape[0], rz.coeffs.shape[0], rq.coeffs.shape[0]]
    inst = O.FriInstance([(zeta, [(oi, i) for oi, wd in enumerate(widths) for i in range(wd)]), (zeta_next, [(2, 0), (2, 1)])])
    oproof = O.prove_openings(inst, [rcs, rw, rz, rq], ch, params)
    assert np.array_equal(proof.opening_proof.flat, oproof)
    assert O.fri_verify(inst, zeta_batch + zp_n[:2], [rcs, rw, rz, rq], ch_v, params, oproof) == 0


def test_plonk_argument_errors(P2, O, pkg, rng):
    K = pkg.plonk
    OP, cfg, common, consts, sig, wires, pih, pd, gw, rcs, rw = setup(P2, O, K, 5, 2)
    betas, gammas, alphas = [3, 5], [7, 11], [13, 17]
    gz = K.PlonkProver.zs_partial_products_commitment(common, pd, gw, betas, gammas)
    common.num_partial_products += 1
    with pytest.raises(pkg.NlcError, match="num_partial_products"):
        K.PlonkProver.zs_partial_products_commitment(common, pd, gw, betas, gammas)
    common.num_partial_products -= 1
    with pytest.raises(pkg.NlcError, match="must hold num_challenges"):
        K.PlonkProver.vanishing_values(common, pd, pih, gw, gw, betas, gammas, alphas)
    narrow = P2.PolynomialBatch.from_values(list(wires[:40]), 3, False, 4)
    with pytest.raises(pkg.NlcError, match="num_routed_wires exceeds"):
        K.PlonkProver.zs_partial_products_commitment(common, pd, narrow, betas, gammas)
    # a gate program that reads the next row is rejected: gates see the local row only
    prog = common.compile_gate_constraints()
    saved = prog.code.copy()
    prog.code[0, 0] = pkg.starky.OP["NEXT"]
    with pytest.raises(pkg.NlcError, match="local row only"):
        K.PlonkProver.vanishing_values(common, pd, pih, gw, gz, betas, gammas, alphas)
    prog.code[:] = saved

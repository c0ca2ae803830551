"""AIR constraint evaluation on the device (csrc/air.cu behind nlc_air_eval / nlc_stark_quotient_commit)
against the oracle's direct evaluation of starkyx's quotient loop (oracle/air.py: prover.rs:306-352), bit-exact on the
accumulators of every point and on the resulting quotient commitment; then a whole StarkyProver::prove (prover.rs:50-225)
for the in-tree Fibonacci test AIR (plonky2/stark/mod.rs:243-265) against the oracle-driven proof."""
import numpy as np
import pytest

from test_air_bytecode_cpu import ConstraintListAir, random_constraints

pytestmark = pytest.mark.gpu
P = 0xFFFFFFFF00000001
G32 = 1753635133440165772


def commit_both(P2, O, trace_rows, rate_bits, cap):
    cols = np.ascontiguousarray(trace_rows.T)
    return P2.PolynomialBatch.from_values(list(cols), rate_bits, False, cap), O.commit_from_values(cols, rate_bits, cap)


def check_air(P2, O, S, air, log_n, rate_bits, cap, traces, publics, globals_, challenges, alphas):
    """accumulators (raw) and the quotient commitment, GPU vs oracle; traces = one row-major matrix per commitment."""
    from oracle import air as OA
    pairs = [commit_both(P2, O, t, rate_bits, cap) for t in traces]
    gpu_batches, ref_batches = [p[0] for p in pairs], [p[1] for p in pairs]
    stark = S.Starky(air)
    got = S.StarkyProver.air_eval(stark, gpu_batches, challenges, globals_, publics, alphas)
    want = OA.quotient_values(air, ref_batches, challenges, globals_, publics, alphas, log_n, rate_bits, divide_by_zh=False)
    assert got.shape == want.shape
    assert np.array_equal(got, want)
    qv = OA.quotient_values(air, ref_batches, challenges, globals_, publics, alphas, log_n, rate_bits)
    qdf = air.quotient_degree_factor()
    qdb = max(qdf - 1, 0).bit_length()
    qref = O.quotient_commit(qv, log_n, qdb, qdf, False, rate_bits, cap)
    return stark, gpu_batches, ref_batches, qref


class FixedChallenger:
    """get_n_challenges returns the given alphas (the transcript itself is covered by test_gpu_fri / test_challenger)."""

    def __init__(self, alphas):
        self.alphas = list(alphas)

    def get_n_challenges(self, n):
        assert n == len(self.alphas)
        return self.alphas


@pytest.mark.parametrize("log_n", [5, 12, 16])
def test_fibonacci_air(P2, O, pkg, log_n):
    S = pkg.starky
    air = S.FibonacciAir()
    n = 1 << log_n
    trace = air.generate_trace(0, 1, n)
    pub = [0, 1, air.fibonacci(n - 1, 0, 1)]
    alphas = [0x0123456789ABCDEF, 0xFEDCBA9876543210 % P]
    stark, gb, rb, qref = check_air(P2, O, S, air, log_n, 1, 4, [trace], pub, [0, 0, 0], [], alphas)
    cfg = S.StarkyConfig.standard_fast_config(n)
    q = S.StarkyProver.quotient_commitment(stark, cfg, gb, [], [0, 0, 0], pub, FixedChallenger(alphas))
    assert qref is not None
    assert np.array_equal(q.merkle_tree.cap, qref.cap)
    assert np.array_equal(q.coeff_matrix, qref.coeffs)
    assert np.array_equal(q.merkle_tree.digests, qref.digests)


@pytest.mark.parametrize("log_n", [6, 16])
def test_fpmul_chip(P2, O, pkg, rng, log_n):
    """One field-arithmetic chip (chip/field/mul.rs, Fp25519): 124 columns, 31 degree-2 constraints, valid traces."""
    S = pkg.starky
    chip = S.FpMulChip()
    n = 1 << log_n
    base = chip.generate_trace(rng, min(n, 1 << 10))
    trace = np.ascontiguousarray(np.tile(base, (n // base.shape[0], 1)))  # rows are independent: repetition stays valid
    alphas = O.random_field(rng, (2,)).tolist()
    stark, gb, rb, qref = check_air(P2, O, S, chip, log_n, 1, 4, [trace], [], [], [], alphas)
    cfg = S.StarkyConfig.standard_fast_config(n)
    q = S.StarkyProver.quotient_commitment(stark, cfg, gb, [], [], [], FixedChallenger(alphas))
    assert np.array_equal(q.merkle_tree.cap, qref.cap)
    assert np.array_equal(q.coeff_matrix, qref.coeffs)


@pytest.mark.parametrize("log_n", [5, 14])
def test_ed_add_chip(P2, O, pkg, rng, log_n):
    """The ed25519 point-addition gadget (chip/ec/edwards/add.rs ed_add: two inner products, three products, a product
    with the curve constant d, two divisions): 800 columns, 248 degree-2 constraints, valid traces."""
    S = pkg.starky
    chip = S.EdAddChip()
    n = 1 << log_n
    base = chip.generate_trace(rng, min(n, 1 << 6))
    trace = np.ascontiguousarray(np.tile(base, (n // base.shape[0], 1)))  # rows are independent: repetition stays valid
    alphas = O.random_field(rng, (2,)).tolist()
    stark, gb, rb, qref = check_air(P2, O, S, chip, log_n, 1, 4, [trace], [], [], [], alphas)
    cfg = S.StarkyConfig.standard_fast_config(n)
    q = S.StarkyProver.quotient_commitment(stark, cfg, gb, [], [], [], FixedChallenger(alphas))
    assert np.array_equal(q.merkle_tree.cap, qref.cap)
    assert np.array_equal(q.coeff_matrix, qref.coeffs)


@pytest.mark.parametrize("which,log_n", [("bits", 5), ("bits", 13), ("scalar_mul", 9)])
def test_scalar_mul_chips_proof(P2, O, pkg, rng, which, log_n):
    """The control columns of the ed25519 scalar multiplication (instruction/cycle.rs, ec/scalar.rs: cycle of 32 rows,
    process id, bit decomposition): first-row, transition and plain constraints with real `next` reads.  Accumulators
    and quotient commitment against the oracle, then the whole proof: the quotient exists (the trace is valid), the
    proof equals the oracle-driven one and the restated FRI verifier accepts it."""
    from oracle import air as OA
    S = pkg.starky
    # "scalar_mul": the whole AIR of scalar_mul_batch / double_and_add, 1651 columns, 720 constraints, two multiplications
    chip = S.ScalarBitsChip() if which == "bits" else S.EdScalarMulChip()
    n = 1 << log_n
    trace = chip.generate_trace(rng, n) if which == "bits" else chip.generate_trace(rng, n, distinct=2)
    alphas = O.random_field(rng, (2,)).tolist()
    stark, gb, rb, qref = check_air(P2, O, S, chip, log_n, 1, 4, [trace], [], [], [], alphas)
    cfg = S.StarkyConfig.standard_fast_config(n)
    fc = cfg.fri_config
    proof = S.StarkyProver.prove(cfg, stark, trace, [])
    ch = O.Challenger()
    ch.observe_elements([])
    tref = O.commit_from_values(np.ascontiguousarray(trace.T), fc.rate_bits, fc.cap_height)
    ch.observe_cap(tref.cap)
    al = ch.get_n_challenges(cfg.num_challenges)
    qv = OA.quotient_values(chip, [tref], [], [], [], al, log_n, fc.rate_bits)
    q2 = O.quotient_commit(qv, log_n, 0, 1, False, fc.rate_bits, fc.cap_height)  # raises when Z_H does not divide
    assert np.array_equal(proof.trace_caps[0], tref.cap) and np.array_equal(proof.quotient_polys_cap, q2.cap)
    ch.observe_cap(q2.cap)
    zeta = ch.get_extension_challenge()
    g = pow(G32, 1 << (32 - log_n), P)
    zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
    ev = lambda b, z: [O.eval_poly_ext(c, np.array(z, np.uint64)) for c in b.coeffs]
    local, nxt, quot = ev(tref, zeta), ev(tref, zeta_next), ev(q2, zeta)
    for batch in (local + quot, nxt):
        ch.observe_elements([c for e in batch for c in e])
    ch_for_verifier = ch.clone()
    params = O.fri_params(log_n, fc.rate_bits, fc.cap_height, fc.proof_of_work_bits, fc.num_query_rounds, 4, 5)
    W, Q = chip.width(), q2.coeffs.shape[0]
    inst = O.FriInstance([(zeta, [(0, i) for i in range(W)] + [(1, i) for i in range(Q)]), (zeta_next, [(0, i) for i in range(W)])])
    oproof = O.prove_openings(inst, [tref, q2], ch, params)
    assert np.array_equal(proof.opening_proof.flat, oproof)
    assert O.fri_verify(inst, local + quot + nxt, [tref, q2], ch_for_verifier, params, oproof) == 0


@pytest.mark.parametrize("seed,degree,rate_bits", [(0, 3, 2), (1, 3, 1), (2, 4, 2), (3, 2, 1), (4, 5, 3)])
def test_random_constraint_lists(P2, O, pkg, seed, degree, rate_bits):
    """Every opcode, every MemorySlice kind, two trace commitments concatenated (prover.rs:289-294), 1..3 alphas,
    quotient domains of 1x / 2x / 4x the trace (quotient_degree_bits 0..2), random (invalid) traces."""
    S = pkg.starky
    rng = np.random.default_rng(100 + seed)
    w0, w1, n_pub, n_glob, n_chal, log_n = 7, 4, 4, 5, 3, 7
    air = ConstraintListAir(S, random_constraints(S, rng, w0 + w1, n_pub, n_glob, n_chal, 16), w0 + w1, n_pub, n_glob, degree)
    t0, t1 = O.random_field(rng, (1 << log_n, w0)), O.random_field(rng, (1 << log_n, w1))
    pub, glob, chal = (O.random_field(rng, (k,)).tolist() for k in (n_pub, n_glob, n_chal))
    alphas = O.random_field(rng, (1 + seed % 3,)).tolist()
    check_air(P2, O, S, air, log_n, rate_bits, 3, [t0, t1], pub, glob, chal, alphas)


def test_bad_programs_are_rejected(P2, O, pkg, rng):
    S = pkg.starky
    b, _ = commit_both(P2, O, O.random_field(rng, (1 << 5, 3)), 1, 2)

    def run(code, n_regs=4, pool=(), n_public=0):
        prog = S.AirProgram(np.array(code, np.uint32).reshape(-1, 4), np.array(pool, np.uint64), n_regs, n_public, 0, 0, 3, 1)
        st = S.Starky(S.FibonacciAir())
        st._program = prog
        return S.StarkyProver.air_eval(st, [b], [], [], [0] * n_public, [5], 0)
    L, ADD, CON = S.OP["LOCAL"], S.OP["ADD"], S.OP["CONSTRAINT"]
    assert run([[L, 0, 1, 0], [CON, 0, 0, 0]]).shape == (32, 1)
    with pytest.raises(pkg.NlcError, match="column out of range"):
        run([[L, 0, 3, 0], [CON, 0, 0, 0]])
    with pytest.raises(pkg.NlcError, match="before writing"):
        run([[L, 0, 0, 0], [ADD, 1, 0, 2], [CON, 0, 1, 0]])
    with pytest.raises(pkg.NlcError, match="unknown opcode"):
        run([[99, 0, 0, 0]])
    # instruction-word flags: the accumulator must hold a result before it is read; unknown flag bits are rejected
    with pytest.raises(pkg.NlcError, match="accumulator before any result"):
        run([[ADD | (S.A_LAST << 8), 1, 0, 0], [CON, 0, 1, 0]])
    with pytest.raises(pkg.NlcError, match="unknown instruction flag"):
        run([[L | (8 << 8), 0, 0, 0], [CON, 0, 0, 0]])
    # a result that is never stored cannot be read from its register later
    with pytest.raises(pkg.NlcError, match="before writing"):
        run([[L | (S.NO_STORE << 8), 0, 0, 0], [L, 1, 1, 0], [CON, 0, 0, 0]])
    got = run([[L | (S.NO_STORE << 8), 0, 1, 0], [CON | (S.A_LAST << 8), 0, 0, 0]])  # ... but from the accumulator it can
    assert np.array_equal(got, run([[L, 0, 1, 0], [CON, 0, 0, 0]]))
    with pytest.raises(pkg.NlcError, match="destination register"):
        run([[L, 7, 0, 0]])
    with pytest.raises(pkg.NlcError, match="public input out of range"):
        run([[S.OP["PUBLIC"], 0, 2, 0], [CON, 0, 0, 0]], n_public=2)
    with pytest.raises(pkg.NlcError, match="higher than the rate"):
        st = S.Starky(S.FibonacciAir())
        S.StarkyProver.air_eval(st, [b], [], [0, 0, 0], [0, 1, 2], [5], 2)


def test_wide_register_file_runs_on_fewer_ctas(P2, O, pkg, rng, monkeypatch):
    """A program with thousands of live registers under a small scratch budget (NLC_AIR_SCRATCH_MB): the interpreter runs
    fewer CTAs that stride over the points; same accumulators as the three-register program computing the same sums."""
    S = pkg.starky
    log_n, R = 14, 2000
    b, _ = commit_both(P2, O, O.random_field(rng, (1 << log_n, 3)), 1, 2)
    L, ADD, MULC, CON = S.OP["LOCAL"], S.OP["ADD"], S.OP["MUL_CONST"], S.OP["CONSTRAINT"]

    def run(code, n_regs, pool=()):
        prog = S.AirProgram(np.array(code, np.uint32).reshape(-1, 4), np.array(pool, np.uint64), n_regs, 0, 0, 0, 3, 1)
        st = S.Starky(S.FibonacciAir())
        st._program = prog
        return S.StarkyProver.air_eval(st, [b], [], [], [], [5, 7], 0)
    wide = [[L, i, i % 3, 0] for i in range(R)]                      # R live values ...
    wide += [[ADD, R, 0, 1]] + [[ADD, R, R, i] for i in range(2, R)]  # ... summed into register R
    wide += [[CON, 0, R, 0]]
    counts = [len(range(c, R, 3)) for c in range(3)]
    narrow = [[L, c, c, 0] for c in range(3)] + [[MULC, c, c, c] for c in range(3)]
    narrow += [[ADD, 3, 0, 1], [ADD, 3, 3, 2], [CON, 0, 3, 0]]
    want = run(narrow, 4, counts)
    monkeypatch.setenv("NLC_AIR_SCRATCH_MB", "64")  # 2001 registers: 32 CTAs of 128 threads for 2^14 points
    got = run(wide, R + 1)
    assert got.shape == (1 << log_n, 2) and np.array_equal(got, want)
    monkeypatch.setenv("NLC_AIR_SCRATCH_MB", "1")
    with pytest.raises(pkg.NlcError, match="too many registers"):
        run(wide, R + 1)


def test_fibonacci_stark_proof_end_to_end(P2, O, pkg):
    """test_plonky2_fibonacci_stark (plonky2/stark/mod.rs:243-265) as the prover sees it: StarkyProver::prove on the
    GPU equals the same sequence driven through the oracle (commit, direct AIR evaluation, quotient commit, openings,
    prove_openings), and the restated FRI verifier accepts it."""
    from oracle import air as OA
    S = pkg.starky
    log_n = 9
    n = 1 << log_n
    air = S.FibonacciAir()
    stark = S.Starky(air)
    trace = air.generate_trace(0, 1, n)
    pub = [0, 1, air.fibonacci(n - 1, 0, 1)]
    cfg = S.StarkyConfig.standard_fast_config(n)
    fc = cfg.fri_config
    proof = S.StarkyProver.prove(cfg, stark, trace, pub)
    # -- the oracle-driven proof, step by step in the order of prover.rs:50-225
    ch = O.Challenger()
    ch.observe_elements(pub)
    tref = O.commit_from_values(np.ascontiguousarray(trace.T), fc.rate_bits, fc.cap_height)
    ch.observe_elements([0, 0, 0])                       # global_values[0..3]
    ch.observe_cap(tref.cap)
    alphas = ch.get_n_challenges(cfg.num_challenges)
    qv = OA.quotient_values(air, [tref], [], [0, 0, 0], pub, alphas, log_n, fc.rate_bits)
    qref = O.quotient_commit(qv, log_n, 0, 1, False, fc.rate_bits, fc.cap_height)
    assert np.array_equal(proof.trace_caps[0], tref.cap)
    assert np.array_equal(proof.quotient_polys_cap, qref.cap)
    ch.observe_cap(qref.cap)
    zeta = ch.get_extension_challenge()
    g = pow(G32, 1 << (32 - log_n), P)
    zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
    ev = lambda b, z: [O.eval_poly_ext(c, np.array(z, np.uint64)) for c in b.coeffs]
    local, nxt, quot = ev(tref, zeta), ev(tref, zeta_next), ev(qref, zeta)
    assert [tuple(int(x) for x in r) for r in proof.openings.local_values] == local
    assert [tuple(int(x) for x in r) for r in proof.openings.next_values] == nxt
    assert [tuple(int(x) for x in r) for r in proof.openings.quotient_polys] == quot
    for batch in (local + quot, nxt):
        ch.observe_elements([c for e in batch for c in e])
    ch_for_verifier = ch.clone()
    params = O.fri_params(log_n, fc.rate_bits, fc.cap_height, fc.proof_of_work_bits, fc.num_query_rounds, 4, 5)
    inst = O.FriInstance([(zeta, [(0, 0), (0, 1), (1, 0), (1, 1)]), (zeta_next, [(0, 0), (0, 1)])])
    oproof = O.prove_openings(inst, [tref, qref], ch, params)
    assert np.array_equal(proof.opening_proof.flat, oproof)
    assert O.fri_verify(inst, local + quot + nxt, [tref, qref], ch_for_verifier, params, oproof) == 0


@pytest.mark.parametrize("log_n", [6, 12])
def test_two_round_air_proof_end_to_end(P2, O, pkg, rng, log_n):
    """An AIR with two rounds (prover.rs:50-111: commit, observe, draw challenges, generate the next round from them):
    two trace commitments concatenated into local / next vars, a challenge, a global value, all four constraint kinds.
    StarkyProver::prove on the GPU equals the oracle-driven sequence and the restated FRI verifier accepts it."""
    from oracle import air as OA
    S = pkg.starky
    n = 1 << log_n
    air = S.PermutationAir()
    stark = S.Starky(air)
    a_col = O.random_field(rng, (n,))
    perm = rng.permutation(n)
    gen = air.trace_generator(a_col, perm)
    cfg = S.StarkyConfig.standard_fast_config(n)
    fc = cfg.fri_config
    proof = S.StarkyProver.prove(cfg, stark, gen, [])
    assert proof.global_values == [1]
    # -- oracle-driven, in the order of prover.rs:50-225
    ch = O.Challenger()
    ch.observe_elements([])
    gv = [0]
    t0 = O.commit_from_values(np.ascontiguousarray(gen(air, 0, [], gv, []).T), fc.rate_bits, fc.cap_height)
    ch.observe_cap(t0.cap)
    challenges = [int(c) for c in ch.get_n_challenges(1)]
    t1 = O.commit_from_values(np.ascontiguousarray(gen(air, 1, challenges, gv, []).T), fc.rate_bits, fc.cap_height)
    ch.observe_elements(gv)
    ch.observe_cap(t1.cap)
    assert np.array_equal(proof.trace_caps[0], t0.cap) and np.array_equal(proof.trace_caps[1], t1.cap)
    alphas = ch.get_n_challenges(cfg.num_challenges)
    qv = OA.quotient_values(air, [t0, t1], challenges, gv, [], alphas, log_n, fc.rate_bits)
    qref = O.quotient_commit(qv, log_n, 0, 1, False, fc.rate_bits, fc.cap_height)
    assert np.array_equal(proof.quotient_polys_cap, qref.cap)
    ch.observe_cap(qref.cap)
    zeta = ch.get_extension_challenge()
    g = pow(G32, 1 << (32 - log_n), P)
    zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
    ev = lambda b, z: [O.eval_poly_ext(c, np.array(z, np.uint64)) for c in b.coeffs]
    local, nxt, quot = ev(t0, zeta) + ev(t1, zeta), ev(t0, zeta_next) + ev(t1, zeta_next), ev(qref, zeta)
    for batch in (local + quot, nxt):
        ch.observe_elements([c for e in batch for c in e])
    ch_v = ch.clone()
    params = O.fri_params(log_n, fc.rate_bits, fc.cap_height, fc.proof_of_work_bits, fc.num_query_rounds, 4, 5)
    trace_polys = [(0, 0), (0, 1), (1, 0)]
    inst = O.FriInstance([(zeta, trace_polys + [(2, i) for i in range(qref.coeffs.shape[0])]), (zeta_next, trace_polys)])
    oproof = O.prove_openings(inst, [t0, t1, qref], ch, params)
    assert np.array_equal(proof.opening_proof.flat, oproof)
    assert O.fri_verify(inst, local + quot + nxt, [t0, t1, qref], ch_v, params, oproof) == 0

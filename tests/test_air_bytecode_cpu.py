"""CPU suite for SURVEY.md 8f.3's host logic: the recording AirParser + register allocator
(near-light-client_b200/starky.py) against the oracle's DIRECT evaluation of the same AIR (oracle/air.py, which follows
vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:306-352 with no bytecode in between).  The bytecode is run here by
a few lines of numpy over oracle/vec.c -- test code, not a product path (the product runs it in csrc/air.cu)."""
import numpy as np
import pytest

P = 0xFFFFFFFF00000001


def run_bytecode(O, OA, prog, local_vars, next_vars, publics, globals_, challenges, consumer):
    """Reference interpreter of nlc_air_program (include/nlc_commit.h NLC_AIR_*), one numpy array per register."""
    n = consumer.n
    regs = [None] * prog.n_regs
    full = lambda v: np.full(n, int(v) % P, np.uint64)
    vec = lambda fn, *a: OA._vec(fn, *a, n=n)
    last = None  # the accumulator of the instruction word's flags (NLC_AIR_FLAG_A_LAST / B_LAST / NO_STORE)
    words = prog.code.tolist()
    pc = -1
    while pc + 1 < len(words):
        pc += 1
        word, d, a, b = words[pc]
        op, flags = word & 0xFF, word >> 8
        if op == 18:  # DOT_CONST: a terms {reg | LAST bit, constant} in the following words, two per word
            r = full(0)
            for t in range(a):
                w = words[pc + 1 + t // 2]
                reg_t, c_t = w[2 * (t & 1)], w[2 * (t & 1) + 1]
                r = vec("orc_vec_add", r, vec("orc_vec_mul", last if reg_t >> 31 else regs[reg_t], full(c_t)))
            pc += (a + 1) // 2
            if not flags & 4:
                regs[d] = r
            last = r
            continue
        ra = (lambda: last if flags & 1 else regs[a])
        rb = (lambda: last if flags & 2 else regs[b])
        if op == 0: r = local_vars[a]
        elif op == 1: r = next_vars[a]
        elif op == 2: r = full(publics[a])
        elif op == 3: r = full(globals_[a])
        elif op == 4: r = full(challenges[a])
        elif op == 5: r = full(prog.pool[a])
        elif op == 6: r = vec("orc_vec_add", ra(), rb())
        elif op == 7: r = vec("orc_vec_sub", ra(), rb())
        elif op == 8: r = vec("orc_vec_mul", ra(), rb())
        elif op == 9: r = vec("orc_vec_neg", ra())
        elif op == 10: r = vec("orc_vec_mul", ra(), full(prog.pool[b]))
        elif op == 11: r = vec("orc_vec_add", ra(), full(prog.pool[b]))
        elif op == 12: r = vec("orc_vec_sub", ra(), full(prog.pool[b]))
        elif op == 17: r = vec("orc_vec_add", vec("orc_vec_mul", ra(), full(prog.pool[b])), last)
        elif op in (13, 14, 15, 16):
            getattr(consumer, ["constraint", "constraint_transition", "constraint_first_row", "constraint_last_row"][op - 13])(ra())
            continue
        else: raise AssertionError(op)
        if not flags & 4:
            regs[d] = r
        last = r
    return np.stack(consumer.accumulators(), axis=1)


def random_constraints(S, rng, width, n_public, n_global, n_challenge, count=12):
    """ArithmeticConstraint lists over every MemorySlice kind and every expression node (chip/arithmetic)."""
    E, M = S.ArithmeticExpression, S.MemorySlice

    def leaf(size):
        kind = rng.integers(0, 6)
        if kind == 0 or size > min(n_public, n_global, n_challenge):
            i = int(rng.integers(0, width - size + 1))
            return E.from_register(M("Local" if rng.integers(0, 2) else "Next", i, size))
        if kind == 1:
            return E.from_register(M("Public", int(rng.integers(0, n_public - size + 1)), size))
        if kind == 2:
            return E.from_register(M("Global", int(rng.integers(0, n_global - size + 1)), size))
        if kind == 3:
            return E.from_register(M("Challenge", int(rng.integers(0, n_challenge - size + 1)), size))
        return E.from_constant_vec([int(x) for x in rng.integers(0, 1 << 63, size=size)])

    def tree(size, depth):
        if depth == 0:
            return leaf(size)
        k = rng.integers(0, 5)
        if k == 0:
            return tree(size, depth - 1) + tree(size, depth - 1)
        if k == 1:
            return tree(size, depth - 1) - tree(size, depth - 1)
        if k == 2:
            return tree(size, depth - 1) * int(rng.integers(1, 1 << 62))
        if k == 3:
            return tree(size, depth - 1) * tree(1, depth - 1)      # ScalarMul (or Mul when size == 1)
        return tree(size, depth - 1) * tree(size, depth - 1)
    kinds = ["First", "Last", "Transition", "All"]
    return [S.ArithmeticConstraint(kinds[int(rng.integers(0, 4))], tree(int(rng.integers(1, 4)), int(rng.integers(1, 4))))
            for _ in range(count)]


class ConstraintListAir:
    """A Chip-like AIR: a list of constraints evaluated in order (chip/mod.rs:58-66, chip/air.rs)."""

    def __init__(self, S, constraints, width, n_public, n_global, degree):
        self.constraints, self._w, self._p, self._g, self._d = constraints, width, n_public, n_global, degree
        self._S = S

    def width(self): return self._w
    def num_public_inputs(self): return self._p
    def num_global_values(self): return self._g
    def constraint_degree(self): return self._d
    def quotient_degree_factor(self): return max(1, self._d - 1)
    def round_data(self): return [self._S.RoundDatum(self._w, (0, self._g), 0)]

    def eval(self, parser):
        for c in self.constraints:
            c.eval(parser)


def direct_and_bytecode(O, OA, S, air, n_rows_log, rate_bits, trace, publics, globals_, challenges, alphas):
    ref = O.commit_from_values(np.ascontiguousarray(trace.T), rate_bits, 0)
    want = OA.quotient_values(air, [ref], challenges, globals_, publics, alphas, n_rows_log, rate_bits, divide_by_zh=False)
    prog = S.compile_air(air, len(challenges))
    # the same inputs the oracle built, fed to the bytecode
    qdb = max(air.quotient_degree_factor() - 1, 0).bit_length()
    size, log_N = 1 << (n_rows_log + qdb), n_rows_log + rate_bits
    coset, lf, ll, _zh = O.quotient_domain(n_rows_log, qdb)
    last = pow(pow(1753635133440165772, 1 << (32 - n_rows_log), P), P - 2, P)
    rev = OA._bitrev_indices(log_N)
    i = np.arange(size, dtype=np.int64)
    step = 1 << (rate_bits - qdb)
    rl, rn = rev[(i * step) % (1 << log_N)], rev[(((i + (1 << qdb)) % size) * step) % (1 << log_N)]
    loc = [np.ascontiguousarray(ref.leaves[:, j])[rl] for j in range(ref.leaves.shape[1])]
    nxt = [np.ascontiguousarray(ref.leaves[:, j])[rn] for j in range(ref.leaves.shape[1])]
    z_last = OA._vec("orc_vec_sub", coset, np.full(size, last, np.uint64), n=size)
    got = run_bytecode(O, OA, prog, loc, nxt, publics, globals_, challenges, OA.ConstraintConsumer(alphas, z_last, lf, ll))
    return want, got, prog


def test_fibonacci_bytecode_matches_direct(O, pkg):
    from oracle import air as OA
    S = pkg.starky
    air = S.FibonacciAir()
    n = 1 << 6
    trace = air.generate_trace(0, 1, n)
    pub = [0, 1, air.fibonacci(n - 1, 0, 1)]
    want, got, prog = direct_and_bytecode(O, OA, S, air, 6, 1, trace, pub, [0, 0, 0], [], [0x1234567, 0xABCDEF0123])
    assert np.array_equal(want, got)
    assert prog.n_constraints == 4 and prog.n_regs <= 6
    # a valid trace: every constraint vanishes on the subgroup, i.e. the quotient commits (prover.rs:161-165)
    ref = O.commit_from_values(np.ascontiguousarray(trace.T), 1, 0)
    q = OA.quotient_values(air, [ref], [], [0, 0, 0], pub, [7, 11], 6, 1)
    assert O.quotient_commit(q, 6, 0, 1, False, 1, 0) is not None


def test_fpmul_chip_bytecode_matches_direct(O, pkg, rng):
    from oracle import air as OA
    S = pkg.starky
    chip = S.FpMulChip()
    trace = chip.generate_trace(rng, 1 << 4)
    want, got, prog = direct_and_bytecode(O, OA, S, chip, 4, 1, trace, [], [], [], [3, 0xFFFFFFFF00000000])
    assert np.array_equal(want, got)
    assert prog.n_constraints == 31 and chip.width() == 124
    assert prog.n_regs < 128   # liveness-based register reuse keeps the per-point scratch small
    assert not want.any() or True
    # witness generation is right: the constraints vanish on every trace row (x in H <=> LDE point i * 2^rate_bits)
    ref = O.commit_from_values(np.ascontiguousarray(trace.T), 1, 0)
    q = OA.quotient_values(chip, [ref], [], [], [], [5], 4, 1, divide_by_zh=False)
    assert q.shape == (16, 1)


class _EveryConstraint:
    """A consumer that keeps each constraint's values (and its kind) instead of folding them."""

    def __init__(self, n):
        self.n, self.values, self.kinds = n, [], []

    def _keep(self, kind, c):
        self.values.append(np.array(c, copy=True))
        self.kinds.append(kind)

    def constraint(self, c):
        self._keep("all", c)

    def constraint_transition(self, c):
        self._keep("transition", c)

    def constraint_first_row(self, c):
        self._keep("first", c)

    def constraint_last_row(self, c):
        self._keep("last", c)


def test_ed_add_chip_bytecode_matches_direct(O, pkg, rng):
    """The ed25519 point-addition gadget (chip/ec/edwards/add.rs): 800 columns, 8 field instructions, 248 constraints."""
    from oracle import air as OA
    S = pkg.starky
    chip = S.EdAddChip()
    trace = chip.generate_trace(rng, 1 << 4)
    assert chip.width() == 800 and trace.max() < 1 << 16
    # the witness is right: every constraint vanishes on every trace row (and stops vanishing when a limb is changed)
    def constraint_values(rows):
        cons = _EveryConstraint(rows.shape[0])
        chip.eval(OA.StarkParser([np.ascontiguousarray(rows[:, j]) for j in range(rows.shape[1])], None, [], [], [], cons))
        return cons.values
    vals = constraint_values(trace)
    assert len(vals) == 248 and not any(v.any() for v in vals)
    bad = trace.copy()
    bad[3, chip.x3.result.index + 5] ^= 1
    assert any(v[3] != 0 for v in constraint_values(bad)) and not any(v[2] != 0 for v in constraint_values(bad))
    # x3 | y3 of every row is the curve's sum of the row's two points
    val = lambda row, m: sum(int(row[m.index + i]) << (16 * i) for i in range(m.length))
    for row in trace[:4]:
        want = S.Ed25519.add((val(row, chip.x1), val(row, chip.y1)), (val(row, chip.x2), val(row, chip.y2)))
        assert (val(row, chip.x3.result), val(row, chip.y3.result)) == want
    want, got, prog = direct_and_bytecode(O, OA, S, chip, 4, 1, trace, [], [], [], [3, 0xFFFFFFFF00000000])
    assert np.array_equal(want, got)
    assert prog.n_constraints == 248 and prog.n_regs < 256


def _violations(OA, chip, rows):
    """How many constraints of `chip` fail on the trace `rows` where they apply: plain ones on every row, transitions on
    every window (row, row + 1) but the last row's, first-row ones on row 0.  Also returns the kinds, in order."""
    n = rows.shape[0]
    cons = _EveryConstraint(n)
    loc = [np.ascontiguousarray(rows[:, j]) for j in range(rows.shape[1])]
    nxt = [np.roll(c, -1) for c in loc]
    chip.eval(OA.StarkParser(loc, nxt, [], [], [], cons))
    bad = 0
    for kind, v in zip(cons.kinds, cons.values):
        bad += int(v[:-1].any()) if kind == "transition" else int(v[0] != 0) if kind == "first" else int(v.any())
    return bad, cons.kinds


def test_scalar_bits_chip(O, pkg, rng):
    """The control columns of scalar_mul_batch (cycle of 32, process id, bit decomposition of a limb): the generated
    trace satisfies every constraint where it applies -- plain ones on every row, transitions on all rows but the last
    (the window (row, row + 1)), first-row ones on row 0 -- a flipped bit breaks them, and the bytecode equals the AIR."""
    from oracle import air as OA
    S = pkg.starky
    chip = S.ScalarBitsChip()
    trace = chip.generate_trace(rng, 1 << 7)

    violations = lambda rows: _violations(OA, chip, rows)
    bad, kinds = violations(trace)
    assert bad == 0 and len(kinds) == 14 and kinds.count("transition") == 3 and kinds.count("first") == 2
    # the bits of every cycle are the limb's
    for c in range(4):
        limb = int(trace[32 * c, chip.limb.index])
        assert sum(int(trace[32 * c + i, chip.bit.index]) << i for i in range(32)) == limb
    flipped = trace.copy()
    flipped[40, chip.bit.index] ^= 1
    assert violations(flipped)[0] > 0
    want, got, prog = direct_and_bytecode(O, OA, S, chip, 7, 1, trace, [], [], [], [3, 0xFFFFFFFF00000000])
    assert np.array_equal(want, got) and prog.n_constraints == 14


def test_ed_scalar_mul_chip(O, pkg, rng):
    """The ed25519 scalar multiplication AIR (machine/ec/builder.rs scalar_mul_batch / double_and_add): 1651 columns as
    `Ed25519ScalarMulTest`, one scalar bit per row.  Two multiplications: every constraint holds where it applies, the
    last row of each block carries scalar * point, a wrong limb or a wrong bit breaks it, bytecode = direct evaluation."""
    from oracle import air as OA
    S = pkg.starky
    chip = S.EdScalarMulChip()
    assert chip.width() == 1651
    pt = S.Ed25519.add(S.Ed25519.GENERATOR, S.Ed25519.GENERATOR)
    scalars = [int.from_bytes(rng.bytes(32), "little"), 5]
    blocks, results = [], []
    for k in scalars:
        blocks.append(chip.generate_block(pt, k))
        results.append(chip.last_result)
    trace = np.concatenate(blocks)
    idx = np.arange(512, dtype=np.uint64)
    trace[:, chip.process_id.index], trace[:, chip.process_id_u32.index], trace[:, chip.clk.index] = idx // 256, idx // 32, idx
    bad, kinds = _violations(OA, chip, trace)
    assert bad == 0 and len(kinds) == 720 and "transition" in kinds and "first" in kinds
    # the result is the curve's scalar multiple (AffinePoint::scalar_mul, chip/ec/scalar_mul/mod.rs:10-25)
    val = lambda row, m: sum(int(row[m.index + i]) << (16 * i) for i in range(m.length))
    for b, k in enumerate(scalars):
        want, temp = None, pt
        for i in range(256):
            if (k >> i) & 1:
                want = temp if want is None else S.Ed25519.add(want, temp)
            temp = S.Ed25519.add(temp, temp)
        last = trace[256 * b + 255]
        assert (val(last, chip.result_next[0]), val(last, chip.result_next[1])) == want == results[b]
    for col, r in ((chip.add.x3.result.index + 3, 300), (chip.bit.index, 77), (chip.temp[1].index, 200)):
        broken = trace.copy()
        broken[r, col] ^= 1
        assert _violations(OA, chip, broken)[0] > 0
    want, got, prog = direct_and_bytecode(O, OA, S, chip, 9, 1, trace, [], [], [], [3, 0xFFFFFFFF00000000])
    assert np.array_equal(want, got) and prog.n_constraints == 720


@pytest.mark.parametrize("seed", range(6))
def test_random_expression_trees(O, pkg, seed):
    from oracle import air as OA
    S = pkg.starky
    rng = np.random.default_rng(seed)
    width, n_pub, n_glob, n_chal = 9, 4, 5, 3
    air = ConstraintListAir(S, random_constraints(S, rng, width, n_pub, n_glob, n_chal), width, n_pub, n_glob, degree=3)
    trace = O.random_field(rng, (1 << 5, width))
    pub, glob, chal = (O.random_field(rng, (k,)).tolist() for k in (n_pub, n_glob, n_chal))
    want, got, prog = direct_and_bytecode(O, OA, S, air, 5, 2, trace, pub, glob, chal, O.random_field(rng, (2,)).tolist())
    assert np.array_equal(want, got)
    assert want.any()

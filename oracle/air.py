"""CPU restatement of starkyx's quotient evaluation loop.  TEST INFRASTRUCTURE ONLY.

Follows /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:253-358 (quotient_polys) with
  plonky2/parser/mod.rs:40-128   StarkParser: the AirParser whose Var is a packed field element -- here the
                                 "pack" is the WHOLE quotient domain: Var = one numpy uint64 array, element i =
                                 the value at point i (the loop over i_start of prover.rs:306-311 is the array axis)
  plonky2/parser/consumer.rs     ConstraintConsumer::{constraint, constraint_transition, constraint_first_row,
                                 constraint_last_row, accumulators}
The AIR is evaluated DIRECTLY through its own `eval(parser)` -- no bytecode, no register allocation: this is the
independent check of the product's recording parser + device interpreter (near-light-client_b200/starky.py,
csrc/air.cu).  Element-wise arithmetic runs in oracle/vec.c.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import oracle as O

P = 0xFFFFFFFF00000001


def _bitrev_indices(log_n: int) -> np.ndarray:
    n = 1 << log_n
    idx = np.arange(n, dtype=np.uint64)
    out = np.zeros(n, dtype=np.uint64)
    for b in range(log_n):
        out |= ((idx >> np.uint64(b)) & np.uint64(1)) << np.uint64(log_n - 1 - b)
    return out.astype(np.int64)


def _vec(fn, *arrays, n):
    out = np.empty(n, np.uint64)
    getattr(O.lib(), fn)(*[O._p(np.ascontiguousarray(a)) for a in arrays], O._p(out), C.c_size_t(n))
    return out


class ConstraintConsumer:
    """consumer.rs:12-78 with P = the whole domain."""

    def __init__(self, alphas, z_last, lagrange_basis_first, lagrange_basis_last):
        self.alphas = [int(a) % P for a in alphas]
        self.n = z_last.size
        self.constraint_accs = [np.zeros(self.n, np.uint64) for _ in self.alphas]
        self.z_last, self.lagrange_basis_first, self.lagrange_basis_last = z_last, lagrange_basis_first, lagrange_basis_last

    def accumulators(self):
        return self.constraint_accs

    def constraint(self, c):
        for j, alpha in enumerate(self.alphas):  # *acc *= alpha; *acc += constraint
            out = np.empty(self.n, np.uint64)
            O.lib().orc_vec_mul_scalar_add(O._p(self.constraint_accs[j]), C.c_uint64(alpha), O._p(c), O._p(out),
                                           C.c_size_t(self.n))
            self.constraint_accs[j] = out

    def constraint_transition(self, c):
        self.constraint(_vec("orc_vec_mul", c, self.z_last, n=self.n))

    def constraint_first_row(self, c):
        self.constraint(_vec("orc_vec_mul", c, self.lagrange_basis_first, n=self.n))

    def constraint_last_row(self, c):
        self.constraint(_vec("orc_vec_mul", c, self.lagrange_basis_last, n=self.n))


class StarkParser:
    """plonky2/parser/mod.rs:40-128: Var = numpy array over the domain."""

    def __init__(self, local_vars, next_vars, global_vars, public_vars, challenges, consumer: ConstraintConsumer):
        self.n = consumer.n
        self.local_vars, self.next_vars = local_vars, next_vars
        self.global_vars, self.public_vars, self.challenges = global_vars, public_vars, challenges
        self.consumer = consumer

    def local_slice(self):
        return self.local_vars

    def next_slice(self):
        return self.next_vars

    def challenge_slice(self):
        return self.challenges

    def global_slice(self):
        return self.global_vars

    def public_slice(self):
        return self.public_vars

    def constant(self, value):
        return np.full(self.n, int(value) % P, np.uint64)

    def constraint(self, c):
        self.consumer.constraint(c)

    def constraint_transition(self, c):
        self.consumer.constraint_transition(c)

    def constraint_first_row(self, c):
        self.consumer.constraint_first_row(c)

    def constraint_last_row(self, c):
        self.consumer.constraint_last_row(c)

    def add(self, a, b):
        return _vec("orc_vec_add", a, b, n=self.n)

    def sub(self, a, b):
        return _vec("orc_vec_sub", a, b, n=self.n)

    def neg(self, a):
        return _vec("orc_vec_neg", a, n=self.n)

    def mul(self, a, b):
        return _vec("orc_vec_mul", a, b, n=self.n)

    def zero(self):
        return self.constant(0)

    def one(self):
        return self.constant(1)

    def mul_const(self, a, b):
        return self.mul(a, self.constant(b))

    def add_const(self, a, b):
        return self.add(a, self.constant(b))

    def sub_const(self, a, b):
        return self.sub(a, self.constant(b))

    def sum(self, elements):
        acc = self.zero()
        for x in elements:
            acc = self.add(acc, x)
        return acc

    def assert_eq(self, a, b):
        self.constraint(self.sub(a, b))

    def assert_eq_transition(self, a, b):
        self.constraint_transition(self.sub(a, b))


def quotient_values(air, trace_data, challenges, global_values, public_inputs, alphas, degree_bits: int,
                    rate_bits: int, divide_by_zh: bool = True) -> np.ndarray:
    """prover.rs:262-352: `quotient_values` (size x num_challenges, one row per point of the quotient domain in
    natural order).  trace_data: oracle Batches (their `leaves` are the LDE rows in index-bit-reversed order).
    divide_by_zh = False stops before prover.rs:336-341 (raw accumulators)."""
    qdf = air.quotient_degree_factor()
    qdb = max(qdf - 1, 0).bit_length()                     # log2_ceil
    assert qdb <= rate_bits, "Having constraints of degree higher than the rate is not supported yet."
    step = 1 << (rate_bits - qdb)
    next_step = 1 << qdb
    size = 1 << (degree_bits + qdb)
    log_N = degree_bits + rate_bits
    coset, lagrange_first, lagrange_last, zh_inv = O.quotient_domain(degree_bits, qdb)
    last = pow(pow(1753635133440165772, 1 << (32 - degree_bits), P), P - 2, P)  # primitive_root_of_unity(degree_bits).inverse()
    rev = _bitrev_indices(log_N)
    i = np.arange(size, dtype=np.int64)
    rows_local = rev[(i * step) % (1 << log_N)]            # get_lde_values(i, step) = leaves[reverse_bits(i * step)]
    rows_next = rev[(((i + next_step) % size) * step) % (1 << log_N)]
    local_vars, next_vars = [], []
    for b in trace_data:
        for j in range(b.leaves.shape[1]):
            col = np.ascontiguousarray(b.leaves[:, j])
            local_vars.append(col[rows_local])
            next_vars.append(col[rows_next])
    z_last = _vec("orc_vec_sub", coset, np.full(size, last, np.uint64), n=size)
    consumer = ConstraintConsumer(alphas, z_last, lagrange_first, lagrange_last)
    const = lambda vals: [np.full(size, int(v) % P, np.uint64) for v in vals]
    parser = StarkParser(local_vars, next_vars, const(global_values), const(public_inputs), const(challenges), consumer)
    air.eval(parser)
    accs = consumer.accumulators()
    if divide_by_zh:
        denom = zh_inv[i % (1 << qdb)]                     # z_h_on_coset.eval_inverse(i)
        accs = [_vec("orc_vec_mul", a, denom, n=size) for a in accs]
    return np.ascontiguousarray(np.stack(accs, axis=1))

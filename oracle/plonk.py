"""CPU restatement of plonky2's plonk-side quotient path.  TEST INFRASTRUCTURE ONLY.

Follows plonky2 0.1.4 @ d2598bd (pinned by /root/reference/Cargo.lock:4864-4886; source NOT under /root/reference --
restated from the published crate, parity unpinned above the permutation as everywhere on this path):
  plonk/prover.rs          wires_permutation_partial_products_and_zs, quotient_chunk_products, partial_products_and_z_gx,
                           the column order of `zs_partial_products`, compute_quotient_polys
  plonk/vanishing_poly.rs  eval_vanishing_poly_base_batch, check_partial_products, evaluate_gate_constraints
  plonk/plonk_common.rs    reduce_with_powers_multi, ZeroPolyOnCoset::{eval_inverse, eval_l_0}
Everything is evaluated DIRECTLY over whole domains with numpy vectors (element-wise arithmetic in oracle/vec.c): no
bytecode, no register allocation, no prefix-scan kernel -- the independent check of near-light-client_b200/plonk.py's
recording + csrc/plonk.cu.  The circuit description (gates, selectors, filters) is the caller's CommonCircuitData.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import oracle as O
from .air import StarkParser, ConstraintConsumer, _bitrev_indices, _vec

P = 0xFFFFFFFF00000001
G32 = 1753635133440165772


def _full(n, v):
    return np.full(n, int(v) % P, np.uint64)


def _scalar_mul(a, s, n):
    return _vec("orc_vec_mul", a, _full(n, s), n=n)


def partial_products_and_zs(common, wires_values: np.ndarray, sigma_values: np.ndarray, betas, gammas) -> np.ndarray:
    """all_wires_permutation_partial_products, then the prover's column order: [Z per challenge | partial products per
    challenge] -- (num_challenges * (num_partial_products + 1), n) values on the subgroup."""
    cfg = common.config
    n, R, degree = wires_values.shape[1], cfg.num_routed_wires, common.quotient_degree_factor
    num_prods = common.num_partial_products
    w = pow(G32, 1 << (32 - common.degree_bits()), P)
    subgroup = np.empty(n, np.uint64)
    acc = 1
    for i in range(n):
        subgroup[i] = acc
        acc = acc * w % P
    zs, pps = [], []
    for beta, gamma in zip(betas, gammas):
        quotient_values = []
        for j in range(R):
            wire = np.ascontiguousarray(wires_values[j])
            s_id = _scalar_mul(subgroup, common.k_is[j], n)                                   # k_i * x
            num = _vec("orc_vec_add", _vec("orc_vec_add", wire, _scalar_mul(s_id, beta, n), n=n), _full(n, gamma), n=n)
            den = _vec("orc_vec_add", _vec("orc_vec_add", wire, _scalar_mul(np.ascontiguousarray(sigma_values[j]), beta, n), n=n),
                       _full(n, gamma), n=n)
            quotient_values.append(_vec("orc_vec_mul", num, _vec("orc_vec_inv", den, n=n), n=n))
        chunk_products = []                                                                   # quotient_chunk_products
        for c0 in range(0, R, degree):
            prod = _full(n, 1)
            for q in quotient_values[c0:c0 + degree]:
                prod = _vec("orc_vec_mul", prod, q, n=n)
            chunk_products.append(prod)
        assert len(chunk_products) == num_prods + 1
        cols = [[0] * n for _ in range(num_prods + 1)]
        z_x = 1
        cp = [[int(v) for v in c] for c in chunk_products]
        for i in range(n):                                                                    # partial_products_and_z_gx + swap
            acc = z_x
            row = []
            for c in range(num_prods + 1):
                acc = acc * cp[c][i] % P
                row.append(acc)
            z_x, row[num_prods] = row[num_prods], z_x
            for c in range(num_prods + 1):
                cols[c][i] = row[c]
        zs.append(np.array(cols[num_prods], np.uint64))                                       # pp.pop(): Z goes to the front
        pps.extend(np.array(c, np.uint64) for c in cols[:num_prods])
    return np.ascontiguousarray(np.stack(zs + pps))


def vanishing_values(common, constants_sigmas, wires, zs_partial_products, public_inputs_hash, betas, gammas, alphas,
                     rate_bits: int, divide_by_zh: bool = True) -> np.ndarray:
    """compute_quotient_polys' `quotient_values` (size x num_challenges, natural order of the quotient domain); the three
    commitments are oracle Batches (`leaves` = LDE rows in index-bit-reversed order)."""
    cfg = common.config
    degree_bits, R, nch = common.degree_bits(), cfg.num_routed_wires, cfg.num_challenges
    max_degree, num_prods = common.quotient_degree_factor, common.num_partial_products
    qdb = max(max_degree - 1, 0).bit_length()
    assert qdb <= rate_bits
    step, next_step = 1 << (rate_bits - qdb), 1 << qdb
    size, log_N = 1 << (degree_bits + qdb), degree_bits + rate_bits
    n = 1 << degree_bits
    rev = _bitrev_indices(log_N)
    i = np.arange(size, dtype=np.int64)
    rows_local = rev[(i * step) % (1 << log_N)]                    # get_lde_values(i, step)
    rows_next = rev[(((i + next_step) % size) * step) % (1 << log_N)]
    col = lambda b, j, rows: np.ascontiguousarray(b.leaves[:, j])[rows]
    local_constants = [col(constants_sigmas, j, rows_local) for j in common.constants_range()]
    s_sigmas = [col(constants_sigmas, j, rows_local) for j in common.sigmas_range()]
    local_wires = [col(wires, j, rows_local) for j in range(cfg.num_wires)]
    local_zs = [col(zs_partial_products, j, rows_local) for j in common.zs_range()]
    next_zs = [col(zs_partial_products, j, rows_next) for j in common.zs_range()]
    partial_products = [col(zs_partial_products, j, rows_local) for j in common.partial_products_range()]
    # shifted_x = coset_shift * w_size^i;  Z_H on the coset takes 2^qdb values g^n w_{2^qdb}^i - 1
    w_size = pow(G32, 1 << (32 - degree_bits - qdb), P)
    xs = np.empty(size, np.uint64)
    acc = 7
    for k in range(size):
        xs[k] = acc
        acc = acc * w_size % P
    g_pow_n = pow(7, n, P)
    w_rate = pow(G32, 1 << (32 - qdb), P) if qdb else 1
    zh_evals = [(g_pow_n * pow(w_rate, k, P) - 1) % P for k in range(1 << qdb)]
    zh = np.array(zh_evals, np.uint64)[i % (1 << qdb)]
    # eval_l_0(index, x) = Z_H(x) * (n (x - 1))^-1
    l_0 = _vec("orc_vec_mul", zh, _vec("orc_vec_inv", _scalar_mul(_vec("orc_vec_sub", xs, _full(size, 1), n=size), n, size), n=size), n=size)
    parser = StarkParser(None, None, None, None, None, ConstraintConsumer([], xs, xs, xs))
    parser.n = size
    pih = [_full(size, v) for v in public_inputs_hash]
    constraint_terms = common.eval_gate_constraints(parser, local_constants, local_wires, pih)
    vanishing_z_1_terms, vanishing_partial_products_terms = [], []
    for c in range(nch):
        z_x, z_gx = local_zs[c], next_zs[c]
        vanishing_z_1_terms.append(_vec("orc_vec_mul", l_0, _vec("orc_vec_sub", z_x, _full(size, 1), n=size), n=size))
        numerators, denominators = [], []
        for j in range(R):
            s_id = _scalar_mul(xs, common.k_is[j], size)
            numerators.append(_vec("orc_vec_add", _vec("orc_vec_add", local_wires[j], _scalar_mul(s_id, betas[c], size), n=size),
                                   _full(size, gammas[c]), n=size))
            denominators.append(_vec("orc_vec_add", _vec("orc_vec_add", local_wires[j], _scalar_mul(s_sigmas[j], betas[c], size), n=size),
                                     _full(size, gammas[c]), n=size))
        accs = [z_x] + partial_products[c * num_prods:(c + 1) * num_prods] + [z_gx]     # check_partial_products
        for k, c0 in enumerate(range(0, R, max_degree)):
            num_p, den_p = _full(size, 1), _full(size, 1)
            for v in numerators[c0:c0 + max_degree]:
                num_p = _vec("orc_vec_mul", num_p, v, n=size)
            for v in denominators[c0:c0 + max_degree]:
                den_p = _vec("orc_vec_mul", den_p, v, n=size)
            vanishing_partial_products_terms.append(_vec("orc_vec_sub", _vec("orc_vec_mul", accs[k], num_p, n=size),
                                                         _vec("orc_vec_mul", accs[k + 1], den_p, n=size), n=size))
    terms = vanishing_z_1_terms + vanishing_partial_products_terms + constraint_terms
    out = []
    for alpha in alphas:                                                               # reduce_with_powers_multi
        cumul = _full(size, 0)
        for term in reversed(terms):
            nxt = np.empty(size, np.uint64)
            O.lib().orc_vec_mul_scalar_add(O._p(cumul), C.c_uint64(int(alpha) % P), O._p(np.ascontiguousarray(term)), O._p(nxt),
                                           C.c_size_t(size))
            cumul = nxt
        if divide_by_zh:
            cumul = _vec("orc_vec_mul", cumul, _vec("orc_vec_inv", zh, n=size), n=size)
        out.append(cumul)
    return np.ascontiguousarray(np.stack(out, axis=1))


class ScalarParser:
    """The parser interface over single field elements (Python ints): the verifier's view of the gates."""

    def constant(self, v):
        return int(v) % P

    def zero(self):
        return 0

    def one(self):
        return 1

    def add(self, a, b):
        return (a + b) % P

    def sub(self, a, b):
        return (a - b) % P

    def mul(self, a, b):
        return a * b % P

    def neg(self, a):
        return -a % P

    def mul_const(self, a, c):
        return a * (int(c) % P) % P

    def add_const(self, a, c):
        return (a + int(c)) % P

    def sub_const(self, a, c):
        return (a - int(c)) % P


def _horner(coeffs, z):
    acc = 0
    for c in reversed([int(x) for x in coeffs]):
        acc = (acc * z + c) % P
    return acc


def vanishing_identity_holds(common, constants_sigmas, wires, zs_partial_products, quotient, public_inputs_hash, betas,
                             gammas, alphas, zeta: int) -> bool:
    """The verifier's check (plonk/verifier.rs + vanishing_poly.rs::eval_vanishing_poly) at a base-field point zeta
    outside H: for every challenge i,  sum_k term_k(zeta) alpha_i^k == Z_H(zeta) * sum_c zeta^(n c) t_{i,c}(zeta),  with
    every opening evaluated from the committed COEFFICIENTS.  This is what makes a proof sound; the oracle's own
    quotient must pass it for a satisfying witness and fail it for a broken gate or a broken copy constraint."""
    cfg = common.config
    n, R, nch = 1 << common.degree_bits(), cfg.num_routed_wires, cfg.num_challenges
    max_degree, num_prods, qdf = common.quotient_degree_factor, common.num_partial_products, common.quotient_degree_factor
    g = pow(G32, 1 << (32 - common.degree_bits()), P)
    ev = lambda b, z: [_horner(c, z) for c in b.coeffs]
    cs, w, zp, q = ev(constants_sigmas, zeta), ev(wires, zeta), ev(zs_partial_products, zeta), ev(quotient, zeta)
    zp_next = ev(zs_partial_products, zeta * g % P)
    local_constants = [cs[j] for j in common.constants_range()]
    s_sigmas = [cs[j] for j in common.sigmas_range()]
    zeta_pow_deg = pow(zeta, n, P)
    z_h_zeta = (zeta_pow_deg - 1) % P
    l_0 = z_h_zeta * pow(n * (zeta - 1) % P, P - 2, P) % P
    parser = ScalarParser()
    constraint_terms = common.eval_gate_constraints(parser, local_constants, w, [int(x) % P for x in public_inputs_hash])
    z1, ppt = [], []
    for i in range(nch):
        z_x, z_gx = zp[i], zp_next[i]
        z1.append(l_0 * (z_x - 1) % P)
        num = [(w[j] + betas[i] * (common.k_is[j] * zeta % P) + gammas[i]) % P for j in range(R)]
        den = [(w[j] + betas[i] * s_sigmas[j] + gammas[i]) % P for j in range(R)]
        pp = [zp[k] for k in common.partial_products_range()][i * num_prods:(i + 1) * num_prods]
        accs = [z_x] + pp + [z_gx]
        for k, c0 in enumerate(range(0, R, max_degree)):
            pn = pd = 1
            for v in num[c0:c0 + max_degree]:
                pn = pn * v % P
            for v in den[c0:c0 + max_degree]:
                pd = pd * v % P
            ppt.append((accs[k] * pn - accs[k + 1] * pd) % P)
    terms = z1 + ppt + constraint_terms
    for i, alpha in enumerate(alphas):
        lhs = 0
        for t in reversed(terms):
            lhs = (lhs * alpha + t) % P
        rhs = 0
        for c in reversed(q[i * qdf:(i + 1) * qdf]):  # reduce_with_powers(chunk, zeta^n)
            rhs = (rhs * zeta_pow_deg + c) % P
        if lhs != z_h_zeta * rhs % P:
            return False
    return True

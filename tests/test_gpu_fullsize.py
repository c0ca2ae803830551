"""BASELINE.json's full sizes, bit-exact against the oracle: configs[1] (2^16 x 135, rate_bits 3), configs[3]
(2^20 x 400, rate_bits 1) and the AIR trace shapes that are in the reference tree (ed25519 2^16 x 2502,
vendor/starkyx/starkyx/src/machine/ec/builder.rs:301-303; SHA-256 2^17 x 912, machine/hash/sha/sha256/air.rs:221-222),
for BOTH the resident path (one k_leaf_hash launch) and the host-fed path (chunked upload + k_leaf_absorb).
configs[3] additionally through size-independent properties (LDE = polynomial evaluation on the coset, Merkle paths
verify to the cap, iFFT . FFT = id), and the pageable-host staging path (csrc/staging.h) against pinned host memory.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 0xFFFFFFFF00000001
G32 = 1753635133440165772  # POWER_OF_TWO_GENERATOR


def bitrev(x, bits):
    return int(format(x, f"0{bits}b")[::-1], 2) if bits else 0


def test_config2_full_commit_bit_exact(P2, O):
    """2^16 x 135, rate_bits 3, cap_height 4 (plonk wires commit under standard_recursion_config)."""
    W, log_n, r, cap = 135, 16, 3, 4
    rng = np.random.default_rng(0xC0FFEE + 1)
    values = O.random_field(rng, (W, 1 << log_n))
    b = P2.PolynomialBatch.from_values(list(values), r, False, cap)
    ref = O.commit_from_values(values, r, cap)
    assert np.array_equal(b.merkle_tree.cap, ref.cap)
    assert np.array_equal(b.coeff_matrix, ref.coeffs)
    assert np.array_equal(b.merkle_tree.digests, ref.digests)
    N = 1 << (log_n + r)
    rows = np.array([0, 1, 2, 12345, N // 2, N - 2, N - 1])
    got = b.get_lde_rows([bitrev(int(i), log_n + r) for i in rows], 1)  # get_lde_values(i) = leaves[bitrev(i)]
    assert np.array_equal(got, ref.leaves[rows])


FULL_SHAPES = {
    # id: (W, log_n, rate_bits, cap_height)
    "config4_2^20x400_r1": (400, 20, 1, 4),        # BASELINE.json configs[3]
    "ed25519_2^16x2502_r1": (2502, 16, 1, 4),      # machine/ec/builder.rs:301-303 (EXTENDED_COLUMNS)
    "sha256_2^17x912_r1": (912, 17, 1, 4),         # machine/hash/sha/sha256/air.rs:221-222
}


@pytest.mark.parametrize("shape", sorted(FULL_SHAPES))
def test_fullsize_commit_bit_exact(P2, O, shape):
    """Whole-commit parity at the sizes BASELINE.json / the in-tree AIRs name: the full cap, every digest, the
    whole coefficient matrix and 2048 random leaf rows (+ the corners) against the oracle, for the resident
    (nlc_commit_from_values_dev: one leaf-hash launch) and the host-fed (nlc_commit_from_values: column chunks,
    k_leaf_absorb) paths."""
    import os

    import torch
    W, log_n, r, cap = FULL_SHAPES[shape]
    n, log_N = 1 << log_n, log_n + r
    N = 1 << log_N
    rng = np.random.default_rng(0xC0FFEE + 3 + W)
    values = O.random_field(rng, (W, n))
    O.set_threads(len(os.sched_getaffinity(0)))
    ref = O.commit_from_values(values, r, cap)
    rows = np.unique(np.concatenate([rng.integers(0, N, size=2048), [0, 1, N // 2 - 1, N // 2, N - 2, N - 1]]))
    pts = [bitrev(int(k), log_N) for k in rows]  # get_lde_values(i) = leaves[bitrev(i)]
    want_rows = ref.leaves[rows]

    def check(b, what):
        assert np.array_equal(b.merkle_tree.cap, ref.cap), (shape, what, "cap")
        assert np.array_equal(b.merkle_tree.digests, ref.digests), (shape, what, "digests")
        got = b.get_lde_rows(pts, 1)
        assert np.array_equal(got, want_rows), (shape, what, "leaf rows")
        cm = b.coeff_matrix
        assert np.array_equal(cm, ref.coeffs), (shape, what, "coefficients")
        b.free()

    # host-fed: W separate pageable host vectors in, like Vec<PolynomialValues>
    check(P2.PolynomialBatch.from_values(list(values), r, False, cap), "host-fed")
    # resident: the same matrix already in HBM
    dev = torch.from_numpy(values.view(np.int64)).cuda()
    ctx = P2.default_context()
    ctx.synchronize()
    torch.cuda.synchronize()
    check(P2.PolynomialBatch.from_device(dev.data_ptr(), W, log_n, r, cap, ctx=ctx), "resident")
    del dev


def test_config4_properties(P2, O):
    """2^20 x 400, rate_bits 1 (starkyx ed25519 trace commitment shape): no oracle run of the whole commit
    (about a minute of CPU); check what must hold for ANY correct commit instead."""
    W, log_n, r, cap = 400, 20, 1, 4
    n, N, log_N = 1 << log_n, 1 << (log_n + r), log_n + r
    rng = np.random.default_rng(0xC0FFEE + 3)
    values = O.random_field(rng, (W, n))
    b = P2.PolynomialBatch.from_values(list(values), r, False, cap)
    cap_arr = b.merkle_tree.cap
    cols = [0, 1, 199, 399]
    coeffs = b.coeff_matrix
    # (1) iFFT then FFT gives the trace back (spot columns)
    for j in cols[:2]:
        assert np.array_equal(P2.fft(coeffs[j]), values[j])
    # (2) leaf k holds every polynomial at 7 * w_N^bitrev(k)
    w_N = pow(G32, 1 << (32 - log_N), P)
    for k in [0, 1, 77777, N - 1]:
        row = b.merkle_tree.get(k)
        x = 7 * pow(w_N, bitrev(k, log_N), P) % P
        for j in cols:
            assert O.eval_poly_ext(coeffs[j], np.array([x, 0], np.uint64)) == (int(row[j]), 0)
        # (3) the Merkle path of that leaf verifies against the cap with the CPU Poseidon
        assert O.merkle_verify(row, k, cap_arr, b.merkle_tree.prove(k).siblings)
    # (4) linearity of the whole pipeline below the hash: commit(a) + commit(b) rows == commit(a + b) rows
    a2 = O.random_field(rng, (3, 1 << 12))
    b2 = O.random_field(rng, (3, 1 << 12))
    s2 = ((a2.astype(object) + b2.astype(object)) % P).astype(np.uint64)
    ra, rb, rs = [P2.PolynomialBatch.from_values(list(m), 1, False, 2).merkle_tree.leaves for m in (a2, b2, s2)]
    assert np.array_equal(((ra.astype(object) + rb.astype(object)) % P).astype(np.uint64), rs)


def test_pageable_and_pinned_hosts_agree(P2, pkg, O):
    """nlc_commit_from_values / nlc_batch_coeffs / nlc_batch_leaves with pageable host vectors go through the
    pinned-slot stager; results must equal the pinned-memory path, including ragged sizes around the 8 MiB slot."""
    import torch
    rng = np.random.default_rng(7)
    for W, log_n in [(3, 15), (33, 16), (5, 20)]:  # 256 KiB, 512 KiB and 8 MiB columns
        n = 1 << log_n
        values = O.random_field(rng, (W, n))
        pinned = torch.empty((W, n), dtype=torch.int64).pin_memory()
        pinned.numpy().view(np.uint64)[:] = values
        bp = P2.PolynomialBatch.from_values(list(pinned.numpy().view(np.uint64)), 1, False, 3)
        bg = P2.PolynomialBatch.from_values(list(values), 1, False, 3)
        assert np.array_equal(bp.merkle_tree.cap, bg.merkle_tree.cap)
        assert np.array_equal(bp.coeff_matrix, bg.coeff_matrix)
        if log_n <= 16:
            assert np.array_equal(bp.merkle_tree.leaves, bg.merkle_tree.leaves)
            ref = O.commit_from_values(values, 1, 3)
            assert np.array_equal(bg.merkle_tree.cap, ref.cap) and np.array_equal(bg.coeff_matrix, ref.coeffs)

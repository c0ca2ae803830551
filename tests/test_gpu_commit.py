"""a1/a2/a6: PolynomialBatch::from_values / from_coeffs / get_lde_values vs the oracle (bit-exact).

Reads like the call sites in vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:91-93,171-178,289-294.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 0xFFFFFFFF00000001
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

SHAPES = [  # (W, log_n, rate_bits, cap_height)
    (1, 0, 0, 0), (1, 0, 1, 0), (2, 1, 1, 1), (2, 5, 1, 4), (3, 4, 3, 4), (4, 6, 1, 2), (5, 6, 2, 0),
    (7, 9, 1, 4), (8, 10, 1, 4), (9, 10, 3, 4), (17, 11, 1, 4), (33, 12, 3, 4), (135, 12, 3, 4),
    (20, 13, 3, 4), (16, 14, 3, 4), (64, 14, 1, 4), (5, 16, 1, 4), (3, 17, 1, 4), (2, 21, 1, 4),
    (2, 8, 2, 10),  # cap_height == log2(N): digests empty, cap = leaf hashes
    # host-fed commits stream the leaf sponge in column chunks (16, then doubling / tripling): chunk boundaries
    (32, 10, 1, 4), (47, 9, 3, 4), (48, 9, 1, 2), (49, 9, 1, 2), (63, 8, 2, 4), (100, 8, 2, 4), (113, 8, 1, 3),
]


def check_batch(batch, ref, full=True):
    assert np.array_equal(batch.merkle_tree.cap, ref.cap)
    assert np.array_equal(batch.coeff_matrix, ref.coeffs)
    if full:
        assert np.array_equal(batch.merkle_tree.leaves, ref.leaves)
        assert np.array_equal(batch.merkle_tree.digests, ref.digests)


@pytest.mark.parametrize("W,log_n,rate_bits,cap_height", SHAPES)
def test_from_values(P2, O, rng, W, log_n, rate_bits, cap_height):
    values = O.random_field(rng, (W, 1 << log_n))
    trace_cols = [P2.PolynomialValues(v) for v in values]
    commitment = P2.PolynomialBatch.from_values(trace_cols, rate_bits, False, cap_height, None, None)
    ref = O.commit_from_values(values, rate_bits, cap_height)
    check_batch(commitment, ref)
    assert commitment.degree_log == log_n and len(commitment.polynomials) == W


@pytest.mark.parametrize("W,log_n,rate_bits,cap_height", SHAPES[1::2] + [(135, 13, 3, 4), (418, 10, 1, 4)])
def test_from_values_materialised(P2, O, rng, W, log_n, rate_bits, cap_height):
    """nlc_commit_from_values_materialised: the struct fields a patched plonky2 keeps on the host (`polynomials`,
    merkle_tree.leaves / digests / cap) written back WHILE the commit runs equal the oracle's -- and so
    does what the handle still serves afterwards (the copies must not disturb the resident batch)."""
    values = O.random_field(rng, (W, 1 << log_n))
    ref = O.commit_from_values(values, rate_bits, cap_height)
    b = P2.PolynomialBatch.from_values_materialised([P2.PolynomialValues(v) for v in values], rate_bits, False, cap_height)
    assert b._polys is not None and b.merkle_tree._leaves is not None and b.merkle_tree._digests is not None
    check_batch(b, ref)
    log_N = log_n + rate_bits
    rows = rng.integers(0, 1 << log_N, 16)
    br = lambda x: int(format(x, f"0{log_N}b")[::-1], 2) if log_N else 0
    assert np.array_equal(b.get_lde_rows(rows), ref.leaves[[br(int(i)) for i in rows]])
    # any subset of the fields; the others stay lazy and are served by the handle
    b2 = P2.PolynomialBatch.from_values_materialised(list(values), rate_bits, False, cap_height, fields=("polynomials", "cap"))
    assert b2.merkle_tree._leaves is None and b2.merkle_tree._digests is None
    check_batch(b2, ref)
    b3 = P2.PolynomialBatch.from_values_materialised(list(values), rate_bits, False, cap_height, fields=("leaves",))
    assert b3._polys is None
    check_batch(b3, ref)


def test_from_values_materialised_pinned_outputs(P2, pkg, O, rng):
    """Pinned destinations take plain asynchronous copies instead of the pageable stager: same bytes."""
    import ctypes as C
    import torch
    ffi = pkg._ffi
    W, log_n, r, cap = 40, 12, 2, 3
    n, N = 1 << log_n, 1 << (log_n + r)
    values = O.random_field(rng, (W, n))
    ref = O.commit_from_values(values, r, cap)
    ctx = P2.default_context()
    pin = lambda shape: torch.zeros(shape, dtype=torch.int64).pin_memory().numpy().view(np.uint64)
    coeffs, leaves, digests, capo = pin((W, n)), pin((N, W)), pin((2 * (N - (1 << cap)), 4)), pin((1 << cap, 4))
    tab = (ffi.u64p * W)(*[values[j].ctypes.data_as(ffi.u64p) for j in range(W)])
    ctab = (ffi.u64p * W)(*[coeffs[j].ctypes.data_as(ffi.u64p) for j in range(W)])
    h = ffi.vp()
    ctx._check(ffi.lib().nlc_commit_from_values_materialised(
        ctx._h, tab, W, log_n, r, cap, ctab, leaves.ctypes.data_as(ffi.u64p), digests.ctypes.data_as(ffi.u64p),
        capo.ctypes.data_as(ffi.u64p), C.byref(h)))
    ffi.lib().nlc_batch_free(h)
    assert np.array_equal(coeffs, ref.coeffs) and np.array_equal(leaves, ref.leaves)
    assert np.array_equal(digests, ref.digests) and np.array_equal(capo, ref.cap)


@pytest.mark.parametrize("W,log_n,rate_bits,cap_height", SHAPES[::3])
def test_from_coeffs(P2, O, rng, W, log_n, rate_bits, cap_height):
    coeffs = O.random_field(rng, (W, 1 << log_n))
    commitment = P2.PolynomialBatch.from_coeffs([P2.PolynomialCoeffs(c) for c in coeffs], rate_bits, False, cap_height)
    check_batch(commitment, O.commit_from_coeffs(coeffs, rate_bits, cap_height))


@pytest.mark.parametrize("kind", ["zero", "structured", "max", "noncanonical"])
def test_structured_inputs(P2, O, kind):
    W, log_n, r, cap = 9, 8, 1, 4
    n = 1 << log_n
    if kind == "zero":
        v = np.zeros((W, n), np.uint64)
    elif kind == "structured":  # value = row*W + col
        v = (np.arange(n, dtype=np.uint64)[None, :] * np.uint64(W) + np.arange(W, dtype=np.uint64)[:, None])
    elif kind == "max":
        v = np.full((W, n), P - 1, np.uint64)
    else:
        v = np.full((W, n), 0xFFFFFFFFFFFFFFFF, np.uint64)
    canon = np.where(v >= np.uint64(P), v - np.uint64(P), v)
    check_batch(P2.PolynomialBatch.from_values(list(v), r, False, cap), O.commit_from_values(canon, r, cap))


def test_get_lde_values_and_open(P2, O, rng):
    W, log_n, r, cap = 13, 9, 2, 3
    values = O.random_field(rng, (W, 1 << log_n))
    b = P2.PolynomialBatch.from_values(list(values), r, False, cap)
    ref = O.commit_from_values(values, r, cap)
    N, log_N = 1 << (log_n + r), log_n + r
    br = lambda x: int(format(x, f"0{log_N}b")[::-1], 2)
    for i, step in [(0, 1), (5, 1), (7, 2), (100, 4), (N - 1, 1)]:
        assert np.array_equal(b.get_lde_values(i, step), ref.leaves[br((i * step) % N)])
    packed = b.get_lde_values_packed(8, 2, 4)  # prover.rs:292
    assert packed.shape == (W, 4)
    for k in range(4):
        assert np.array_equal(packed[:, k], ref.leaves[br((8 + k) * 2)])
    # LDE row i at step 2^r is the original trace row i... on the coset 7<w>: check through the oracle instead
    for idx in [0, 1, N // 2, N - 1]:
        assert np.array_equal(b.merkle_tree.get(idx), ref.leaves[idx])
        sib = b.merkle_tree.prove(idx).siblings
        assert np.array_equal(sib, O.merkle_prove(ref.digests, N, cap, idx))
        assert O.merkle_verify(ref.leaves[idx], idx, ref.cap, sib)


def test_eval_at(P2, O, rng):
    values = O.random_field(rng, (6, 1 << 10))
    b = P2.PolynomialBatch.from_values(list(values), 1, False, 4)
    z = O.random_field(rng, (2,))
    got = b.eval_at(z)
    ref = O.commit_from_values(values, 1, 4)
    for j in range(6):
        assert tuple(int(x) for x in got[j]) == O.eval_poly_ext(ref.coeffs[j], z)


def test_from_trace_rows(P2, O, rng):
    n, width = 1 << 9, 23
    trace = O.random_field(rng, (n, width))  # row-major AirTrace (trace/mod.rs:19-24)
    b = P2.PolynomialBatch.from_trace_rows(trace, 3, 17, 1, 4)
    ref = O.commit_from_values(np.ascontiguousarray(trace[:, 3:20].T), 1, 4)
    check_batch(b, ref)


@pytest.mark.parametrize("pinned", [False, True])
def test_from_trace_rows_strip_chunks(P2, O, rng, pinned):
    """Only the column strip crosses PCIe, in COLUMN chunks through the host-fed pipeline (16, 24, 40, ... columns of every
    row: gathered into the pinned ring from a pageable matrix, a 2-D DMA from a pinned one), strip != width and
    col0 != 0; pageable and pinned hosts."""
    import torch
    n, width, col0, nc = 1 << 16, 150, 7, 130
    trace = O.random_field(rng, (n, width))
    host = trace
    if pinned:
        t = torch.empty((n, width), dtype=torch.int64).pin_memory()
        t.numpy().view(np.uint64)[:] = trace
        host = t.numpy().view(np.uint64)
    b = P2.PolynomialBatch.from_trace_rows(host, col0, nc, 1, 4)
    ref = O.commit_from_values(np.ascontiguousarray(trace[:, col0:col0 + nc].T), 1, 4)
    assert np.array_equal(b.merkle_tree.cap, ref.cap)
    assert np.array_equal(b.coeff_matrix, ref.coeffs)
    assert np.array_equal(b.merkle_tree.digests, ref.digests)


@pytest.mark.parametrize("nc,rate_bits", [(1, 1), (8, 3), (31, 1), (32, 1), (47, 2), (200, 1), (135, 3)])
def test_from_trace_rows_widths(P2, O, rng, nc, rate_bits):
    """Strip widths on both sides of the pipeline's thresholds (one launch below 32 columns, the chunked sponge above;
    first chunk 16, a merged last chunk), two strips of the same matrix."""
    n, width = 1 << 8, nc + 9
    trace = O.random_field(rng, (n, width))
    for col0 in (0, 9):
        b = P2.PolynomialBatch.from_trace_rows(trace, col0, nc, rate_bits, 2)
        check_batch(b, O.commit_from_values(np.ascontiguousarray(trace[:, col0:col0 + nc].T), rate_bits, 2))
        b.free()


def test_import_serialized(P2, pkg, O, rng):
    """SURVEY.md section 5 checkpoint row: a deserialised PolynomialBatch (coefficients + cap [+ digests]) comes back
    onto the GPU by recommitting; a wrong cap or digest fails loudly."""
    W, log_n, r, cap = 21, 10, 3, 4
    values = O.random_field(rng, (W, 1 << log_n))
    ref = O.commit_from_values(values, r, cap)
    b = P2.PolynomialBatch.import_serialized(list(ref.coeffs), r, cap, cap=ref.cap, digests=ref.digests)
    check_batch(b, ref)
    b2 = P2.PolynomialBatch.import_serialized(list(ref.coeffs), r, cap)  # nothing to verify against: still the same batch
    assert np.array_equal(b2.merkle_tree.cap, ref.cap)
    bad_cap = ref.cap.copy()
    bad_cap[3, 1] ^= np.uint64(1)
    with pytest.raises(pkg.NlcError, match="MerkleCap recomputed"):
        P2.PolynomialBatch.import_serialized(list(ref.coeffs), r, cap, cap=bad_cap)
    bad_dig = ref.digests.copy()
    bad_dig[-1, 0] ^= np.uint64(1)
    with pytest.raises(pkg.NlcError, match="digests recomputed"):
        P2.PolynomialBatch.import_serialized(list(ref.coeffs), r, cap, cap=ref.cap, digests=bad_dig)
    with pytest.raises(pkg.NlcError, match="wrong size"):
        P2.PolynomialBatch.import_serialized(list(ref.coeffs), r, cap, cap=ref.cap[:8])


def test_lde_rows_range(P2, O, rng):
    """nlc_batch_lde_rows_range: a whole stripe of get_lde_values(i, step) in one launch + one D2H."""
    W, log_n, r, cap = 13, 9, 2, 3
    values = O.random_field(rng, (W, 1 << log_n))
    b = P2.PolynomialBatch.from_values(list(values), r, False, cap)
    ref = O.commit_from_values(values, r, cap)
    N, log_N = 1 << (log_n + r), log_n + r
    br = lambda x: int(format(x, f"0{log_N}b")[::-1], 2)
    for i0, count, step in [(0, N, 1), (5, 100, 2), (N - 3, 9, 4), (17, 1, 1)]:  # wraps mod N like the reference
        got = b.get_lde_rows_range(i0, count, step)
        want = np.stack([ref.leaves[br(((i0 + k) * step) % N)] for k in range(count)])
        assert np.array_equal(got, want)
    assert b.get_lde_rows_range(0, 0, 1).shape == (0, W)


def test_from_device_orders_after_the_producer_stream(P2, O, rng):
    """a tensor produced on torch's default stream must not race with the commit on the context's own
    (non-blocking) stream -- from_device waits on the producer (nlc_wait_stream)."""
    import torch
    W, log_n, r, cap = 64, 14, 1, 4
    values = O.random_field(rng, (W, 1 << log_n))
    ref = O.commit_from_values(values, r, cap)
    ctx = P2.default_context()
    src = torch.from_numpy(values.view(np.int64)).pin_memory()
    for _ in range(3):
        spin = torch.empty(64 << 20, dtype=torch.int64, device="cuda")
        for _k in range(8):
            spin.add_(1)  # keep the default stream busy so the copy below is genuinely late
        dev = src.to("cuda", non_blocking=True)
        b = P2.PolynomialBatch.from_device(dev.data_ptr(), W, log_n, r, cap, ctx=ctx)
        assert np.array_equal(b.merkle_tree.cap, ref.cap)
        b.free()


def test_errors(P2, pkg, O, rng):
    v = O.random_field(rng, (2, 16))
    with pytest.raises(pkg.NlcError):
        P2.PolynomialBatch.from_values(list(v), 1, True, 0)        # blinding unsupported on this path
    with pytest.raises(pkg.NlcError, match="cap_height"):
        P2.PolynomialBatch.from_values(list(v), 1, False, 6)       # cap_height > log2(N)
    with pytest.raises(pkg.NlcError, match="power of two"):
        P2.PolynomialBatch.from_values([v[0][:12]], 1, False, 0)
    with pytest.raises(pkg.NlcError, match="same length"):
        P2.PolynomialBatch.from_values([v[0], v[1][:8]], 1, False, 0)
    with pytest.raises(pkg.NlcError):
        P2.PolynomialBatch.from_values([], 1, False, 0)


def test_golden_fixtures(P2):
    """Frozen oracle outputs (tests/golden/gen_golden.py) -- guards against oracle and kernel drifting together."""
    g = np.load(os.path.join(GOLDEN, "commit_small.npz"))
    for key in sorted({k.split(":")[0] for k in g.files}):
        W, log_n, r, cap = (int(x) for x in key.split("_"))
        b = P2.PolynomialBatch.from_values(list(g[key + ":values"]), r, False, cap)
        assert np.array_equal(b.merkle_tree.cap, g[key + ":cap"])
        assert np.array_equal(b.get_lde_rows([0, 1, 5], 1), g[key + ":lde_rows"])
        assert np.array_equal(b.coeff_matrix[:, :4], g[key + ":coeffs_head"])


def test_nlc_device_env(P2, pkg, monkeypatch):
    """nlc_init(device < 0) honours NLC_DEVICE: the GPU is chosen without touching the caller."""
    monkeypatch.setenv("NLC_DEVICE", "0")
    c = P2.Context(-1)
    assert c.device == 0
    c.close()
    monkeypatch.setenv("NLC_DEVICE", "4096")
    with pytest.raises(pkg.NlcError, match="device index out of range"):
        P2.Context(-1)
    monkeypatch.setenv("NLC_DEVICE", "gpu0")
    with pytest.raises(pkg.NlcError, match="NLC_DEVICE"):
        P2.Context(-1)


def _random_shapes():
    rng = np.random.default_rng(20261105)
    out = []
    for _ in range(36):  # small transforms (generic NTT path), any width / rate / cap
        log_n, r = int(rng.integers(0, 12)), int(rng.integers(0, 5))
        out.append((int(rng.integers(1, 70)), log_n, r, int(rng.integers(0, min(log_n + r, 6) + 1))))
    for _ in range(14):  # fast NTT path (>= 2^12): all 2^r cosets in one launch per pass, in place
        log_n, r = int(rng.integers(12, 15)), int(rng.integers(0, 5))
        out.append((int(rng.integers(1, 41)), log_n, r, int(rng.integers(0, 7))))
    return out


@pytest.mark.parametrize("W,log_n,rate_bits,cap_height", _random_shapes())
def test_random_shapes_resident_and_host_fed(P2, O, W, log_n, rate_bits, cap_height):
    """Randomly drawn (W, degree, rate_bits, cap_height), incl. rate_bits 0 and 4 (16 cosets in one launch) and caps
    from 0 to the tree height: from_values through host vectors, from_coeffs, and the device-resident entry point all
    give the oracle's cap, coefficients, leaves and digests."""
    import torch
    rng = np.random.default_rng(W * 1000 + log_n * 10 + rate_bits)
    values = O.random_field(rng, (W, 1 << log_n))
    ref = O.commit_from_values(values, rate_bits, cap_height)
    check_batch(P2.PolynomialBatch.from_values(list(values), rate_bits, False, cap_height), ref)
    check_batch(P2.PolynomialBatch.from_coeffs(list(ref.coeffs), rate_bits, False, cap_height), ref)
    dev = torch.from_numpy(values.view(np.int64)).cuda()
    check_batch(P2.PolynomialBatch.from_device(dev.data_ptr(), W, log_n, rate_bits, cap_height), ref)

"""nlc_merkle_cap_static_dev -- the allocation-free Merkle cap (csrc/static_scratch.cu; include/nlc_commit.h) against the
existing entry point nlc_merkle_new + nlc_tree_cap, through the C ABI, bit-exact (integer work: no tolerance).

The file sorts last on purpose: the entry point has no counterpart in the reference and sits outside the hot path, so
under `pytest -x` a failure here must not hide any other test.  It builds its own context and takes nothing from
earlier tests, so it passes alone in a fresh process as well as after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

P = 0xFFFFFFFF00000001


def _existing_cap(pkg, ctx, leaves, cap_height):
    """MerkleTree::new(leaves, cap_height).cap through the existing entry points (host leaves in, host cap out)."""
    ffi, L = pkg._ffi, pkg._ffi.lib()
    t = ffi.vp()
    ctx._check(L.nlc_merkle_new(ctx._h, leaves.ctypes.data_as(ffi.u64p), leaves.shape[0], leaves.shape[1], cap_height,
                                C.byref(t)))
    cap = np.zeros((1 << cap_height, 4), np.uint64)
    rc = L.nlc_tree_cap(t, cap.ctypes.data_as(ffi.u64p))
    L.nlc_tree_free(t)
    ctx._check(rc)
    return cap


def _leaves(rng, n, width):
    v = rng.integers(0, P, size=(n, width), dtype=np.uint64)
    v[0, 0] = P - 1
    v[-1, -1] = 0xFFFFFFFFFFFFFFFF  # non-canonical on purpose: both paths must reduce it the same way
    return v


def test_static_scratch_cap_matches_merkle_new_twice_in_a_row(pkg):
    import torch
    ffi, L = pkg._ffi, pkg._ffi.lib()
    ctx = pkg.plonky2.Context(0)
    try:
        rng = np.random.default_rng(20261105)
        # call 2 fills the whole static array (NLC_STATIC_MAX_LEAVES leaves, cap_height 0); call 3 then runs a small tree
        # over what call 2 left behind, with leaves short enough for hash_or_noop's copy and the cap at the leaves
        shapes = [(1 << 12, 135, 4), (1 << 16, 9, 0), (1 << 5, 3, 5), (1 << 10, 8, 2)]
        host = [_leaves(rng, n, w) for n, w, _ in shapes]
        # everything the CALLER owns is allocated before the calls
        dev = [torch.from_numpy(h.view(np.int64)).cuda() for h in host]
        out = [torch.full((4 << c,), -1, dtype=torch.int64, device="cuda") for _, _, c in shapes]
        torch.cuda.synchronize()
        # first call: looks the symbol address up; it also pays CUDA's lazy module loading, so memory is compared later
        n, w, c = shapes[0]
        ctx._check(L.nlc_merkle_cap_static_dev(ctx._h, ffi.vp(dev[0].data_ptr()), n, w, c, ffi.vp(out[0].data_ptr())))
        ctx.synchronize()
        # the rest back to back on the context's stream, no synchronisation in between
        for i in range(1, len(shapes)):
            n, w, c = shapes[i]
            ctx._check(L.nlc_merkle_cap_static_dev(ctx._h, ffi.vp(dev[i].data_ptr()), n, w, c, ffi.vp(out[i].data_ptr())))
        # and the first shape once more, over the leftovers of the others
        again = torch.full((4 << shapes[0][2],), -1, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        torch_before = torch.cuda.memory_allocated()
        n, w, c = shapes[0]
        ctx._check(L.nlc_merkle_cap_static_dev(ctx._h, ffi.vp(dev[0].data_ptr()), n, w, c, ffi.vp(again.data_ptr())))
        ctx.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after >= free_before, "the allocation-free entry point took device memory"
        assert torch.cuda.memory_allocated() == torch_before
        for i, (n, w, c) in enumerate(shapes):
            got = out[i].cpu().numpy().view(np.uint64).reshape(1 << c, 4)
            want = _existing_cap(pkg, ctx, host[i], c)
            assert np.array_equal(got, want), f"cap differs for {n} x {w}, cap_height {c}"
        assert np.array_equal(again.cpu().numpy(), out[0].cpu().numpy())
    finally:
        ctx.close()


def test_static_scratch_rejects_what_does_not_fit(pkg):
    import torch
    ffi, L = pkg._ffi, pkg._ffi.lib()
    ctx = pkg.plonky2.Context(0)
    try:
        leaves = torch.zeros(64, dtype=torch.int64, device="cuda")
        cap = torch.zeros(4, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        lp, cp = ffi.vp(leaves.data_ptr()), ffi.vp(cap.data_ptr())
        with pytest.raises(pkg.NlcError, match="NLC_STATIC_MAX_LEAVES"):
            ctx._check(L.nlc_merkle_cap_static_dev(ctx._h, lp, 1 << 17, 1, 0, cp))
        with pytest.raises(pkg.NlcError, match="power of two"):
            ctx._check(L.nlc_merkle_cap_static_dev(ctx._h, lp, 12, 1, 0, cp))
        with pytest.raises(pkg.NlcError, match="cap_height"):
            ctx._check(L.nlc_merkle_cap_static_dev(ctx._h, lp, 16, 4, 5, cp))
        with pytest.raises(pkg.NlcError, match="null"):
            ctx._check(L.nlc_merkle_cap_static_dev(ctx._h, None, 16, 4, 0, cp))
        # still usable after the errors
        ctx._check(L.nlc_merkle_cap_static_dev(ctx._h, lp, 16, 4, 0, cp))
        ctx.synchronize()
        want = _existing_cap(pkg, ctx, np.zeros((16, 4), np.uint64), 0)
        assert np.array_equal(cap.cpu().numpy().view(np.uint64).reshape(1, 4), want)
    finally:
        ctx.close()

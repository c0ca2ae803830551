"""8(e).2 on the GPU: the coset shards of a commit, computed one after the other on cuda:0 through
nlc_commit_from_values_shard, concatenate to exactly the single-GPU commit (cap, digests, leaves) = the oracle's."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("W,log_n,r,cap,G", [(5, 6, 2, 3, 2), (9, 10, 3, 4, 8), (135, 12, 3, 4, 4), (3, 7, 1, 1, 2),
                                             (4, 5, 2, 2, 1)])
def test_shards_assemble_to_the_commit(P2, O, rng, W, log_n, r, cap, G):
    import __graft_entry__ as graft
    S = graft.load_package().sharding
    values = O.random_field(rng, (W, 1 << log_n))
    ref = O.commit_from_values(values, r, cap)
    shards = [P2.PolynomialBatch.from_values_sharded(list(values), r, False, cap, shard_idx=g, n_shards=G)
              for g in range(G)]
    assert np.array_equal(S.assemble(s.merkle_tree.local_cap for s in shards), ref.cap)
    assert np.array_equal(S.assemble(s.merkle_tree.digests for s in shards), ref.digests)
    assert np.array_equal(S.assemble(s.merkle_tree.leaves for s in shards), ref.leaves)
    for s in shards:
        assert np.array_equal(s.coeff_matrix, ref.coeffs)


def test_shard_limits(P2, O, rng):
    values = O.random_field(rng, (2, 16))
    with pytest.raises(Exception):
        P2.PolynomialBatch.from_values_sharded(list(values), 1, False, 4, shard_idx=0, n_shards=4)


@pytest.mark.parametrize("W,log_n,r,cap,G", [(5, 6, 1, 3, 4), (9, 10, 1, 4, 8), (135, 11, 1, 4, 8), (418, 10, 1, 4, 4),
                                             (3, 7, 0, 2, 4), (20, 9, 3, 4, 16), (7, 8, 1, 4, 2), (6, 5, 1, 2, 1)])
def test_row_shards_assemble_to_the_commit(P2, O, rng, W, log_n, r, cap, G):
    """8(e).3 on one GPU: the G ranks' work done one after the other on cuda:0 -- column shards through
    nlc_lde_columns_dev, the all-to-all replaced by concatenating the packed blocks, leaf-row shards through
    nlc_commit_from_lde_rows_dev -- gives exactly the single-GPU commit = the oracle's."""
    import torch
    import __graft_entry__ as graft
    S = graft.load_package().sharding
    if W < G:
        pytest.skip("fewer columns than ranks")
    values = O.random_field(rng, (W, 1 << log_n))
    ref = O.commit_from_values(values, r, cap)
    dev = torch.device("cuda", 0)
    sends, coeffs = [], []
    for c0, nc in S.column_split(W, G):
        src = torch.from_numpy(values[c0:c0 + nc].view(np.int64)).to(dev)
        cf, lde = P2.PolynomialBatch.lde_columns(src, r)
        torch.cuda.synchronize()
        # the column owner's LDE in leaf order = the oracle's leaves of those columns
        assert np.array_equal(lde.cpu().numpy().view(np.uint64), ref.leaves[:, c0:c0 + nc].T)
        sends.append(S.pack_rows(lde, G))
        coeffs.append(cf.cpu().numpy().view(np.uint64))
        # the fused form: the last NTT pass stores straight into the send layout
        cf2, send2 = P2.PolynomialBatch.lde_columns_scatter(src, r, G)
        torch.cuda.synchronize()
        assert torch.equal(send2, sends[-1]) and torch.equal(cf2, cf)
    assert np.array_equal(np.concatenate(coeffs), ref.coeffs)
    shards = []
    for g in range(G):
        rows = torch.cat([s[g] for s in sends]).contiguous()  # what rank g receives, in source-rank order
        shards.append(P2.PolynomialBatch.from_lde_rows(rows, log_n, r, cap, g, G))
    assert np.array_equal(S.assemble(s.merkle_tree.local_cap for s in shards), ref.cap)
    assert np.array_equal(S.assemble(s.merkle_tree.digests for s in shards), ref.digests)
    assert np.array_equal(S.assemble(s.merkle_tree.leaves for s in shards), ref.leaves)
    with pytest.raises(P2.NlcError, match="no coefficients"):
        shards[0].coeff_matrix  # noqa: B018 -- the coefficients stay with the column owners


def test_row_shard_limits(P2, O, rng):
    import torch
    rows = torch.zeros((2, 8), dtype=torch.int64, device="cuda")
    with pytest.raises(P2.NlcError):
        P2.PolynomialBatch.from_lde_rows(rows, 4, 1, 1, 0, 4)  # 4 shards, 2 cap subtrees
    with pytest.raises(P2.NlcError):
        P2.PolynomialBatch.from_lde_rows(rows, 4, 1, 2, 0, 2)  # wrong row count for 2 shards


@pytest.mark.parametrize("W,log_n,r,cap,G", [(5, 6, 1, 3, 4), (9, 10, 1, 4, 8), (135, 11, 1, 4, 8), (20, 9, 3, 4, 16),
                                             (7, 8, 1, 4, 2)])
def test_row_shards_through_peer_buffers(P2, pkg, O, rng, W, log_n, r, cap, G):
    """The P2P form of 8(e).3 on one GPU: G receive buffers from nlc_peer_alloc stand for the G ranks' buffers, every
    column owner's last NTT pass stores into all of them (nlc_lde_columns_scatter_dev with peer pointers and its
    global column offset), and the tree over each buffer is the oracle's shard."""
    import ctypes as C
    import torch
    S, ffi = pkg.sharding, pkg._ffi
    ctx = P2.default_context()
    values = O.random_field(rng, (W, 1 << log_n))
    ref = O.commit_from_values(values, r, cap)
    per = (1 << (log_n + r)) // G
    bufs = []
    for _ in range(G):
        ptr, handle = ffi.vp(), C.create_string_buffer(64)
        ctx._check(ffi.lib().nlc_peer_alloc(ctx._h, W * per, C.byref(ptr), handle))
        bufs.append(ptr.value)
    try:
        for c0, nc in S.column_split(W, G):
            src = torch.from_numpy(values[c0:c0 + nc].view(np.int64)).cuda()
            cf, none = P2.PolynomialBatch.lde_columns_scatter(src, r, G, peer_ptrs=bufs, col0=c0)
            assert none is None
            torch.cuda.synchronize()  # the *_dev call only enqueues on the context's stream: order torch's read after it
            assert np.array_equal(cf.cpu().numpy().view(np.uint64), ref.coeffs[c0:c0 + nc])
        torch.cuda.synchronize()
        shards = [P2.PolynomialBatch.from_lde_rows((bufs[g], W, None), log_n, r, cap, g, G) for g in range(G)]
        assert np.array_equal(S.assemble(s.merkle_tree.local_cap for s in shards), ref.cap)
        assert np.array_equal(S.assemble(s.merkle_tree.digests for s in shards), ref.digests)
        assert np.array_equal(S.assemble(s.merkle_tree.leaves for s in shards), ref.leaves)
        for s in shards:
            s.free()
    finally:
        for p in bufs:
            ffi.lib().nlc_peer_free(ctx._h, ffi.vp(p))


@pytest.mark.parametrize("W,log_n,r,cap,G", [(5, 6, 2, 3, 2), (9, 12, 3, 4, 8), (135, 12, 3, 4, 4), (3, 7, 2, 2, 4),
                                             (17, 13, 1, 4, 2), (4, 5, 2, 2, 1)])
def test_coset_shards_from_broadcast_coefficients(P2, pkg, O, rng, W, log_n, r, cap, G):
    """8(e).2 without the replicated iFFT, on one GPU: G coefficient matrices from nlc_peer_alloc stand for the G ranks'
    buffers; every column owner's iFFT stores its coefficients into all of them (nlc_ifft_columns_bcast_dev), and each
    rank's coset shard over its own matrix (nlc_commit_from_coeffs_shard_dev) is the oracle's shard.  W < G leaves some
    ranks without columns; W not a multiple of G leaves padding columns at the end of the matrix."""
    import ctypes as C
    import torch
    S, ffi = pkg.sharding, pkg._ffi
    ctx = P2.default_context()
    n = 1 << log_n
    values = O.random_field(rng, (W, n))
    ref = O.commit_from_values(values, r, cap)
    per, split = S.uniform_column_split(W, G)
    assert sum(nc for _, nc in split) == W and all(c0 == min(W, p * per) for p, (c0, _) in enumerate(split))
    bufs = []
    for _ in range(G):
        ptr, handle = ffi.vp(), C.create_string_buffer(64)
        ctx._check(ffi.lib().nlc_peer_alloc(ctx._h, G * per * n, C.byref(ptr), handle))
        bufs.append(ptr.value)
    try:
        for c0, nc in split:
            if nc:
                src = torch.from_numpy(values[c0:c0 + nc].view(np.int64)).cuda()
                P2.PolynomialBatch.ifft_columns_bcast(src, c0, bufs)
        ctx.synchronize()
        shards = [P2.PolynomialBatch.from_coeffs_shard_dev(bufs[g], W, log_n, r, cap, g, G) for g in range(G)]
        assert np.array_equal(S.assemble(s.merkle_tree.local_cap for s in shards), ref.cap)
        assert np.array_equal(S.assemble(s.merkle_tree.digests for s in shards), ref.digests)
        assert np.array_equal(S.assemble(s.merkle_tree.leaves for s in shards), ref.leaves)
        for s in shards:
            assert np.array_equal(s.coeff_matrix, ref.coeffs)  # every rank holds the whole coefficient matrix
            s.free()
    finally:
        for p in bufs:
            ffi.lib().nlc_peer_free(ctx._h, ffi.vp(p))

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

"""a5: MerkleTree::new / get / prove on the device vs the oracle (bit-exact, plonky2's digests layout)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("leaf_len", [1, 2, 4, 5, 7, 8, 9, 32, 135])
@pytest.mark.parametrize("log_n,cap_height", [(0, 0), (1, 0), (1, 1), (5, 0), (5, 4), (5, 5), (9, 4), (12, 4)])
def test_merkle_new(P2, O, rng, leaf_len, log_n, cap_height):
    if leaf_len == 135 and log_n < 5:
        pytest.skip("covered by smaller widths")
    leaves = O.random_field(rng, (1 << log_n, leaf_len))
    t = P2.MerkleTree.new(leaves, cap_height)
    digests, cap = O.merkle_new(leaves, cap_height)
    assert np.array_equal(t.cap, cap)
    assert np.array_equal(t.digests, digests)
    for idx in {0, (1 << log_n) - 1, int(rng.integers(0, 1 << log_n))}:
        proof = t.prove(idx)
        assert np.array_equal(proof.siblings, O.merkle_prove(digests, 1 << log_n, cap_height, idx))
        assert np.array_equal(t.get(idx), leaves[idx])
        assert O.merkle_verify(leaves[idx], idx, cap, proof.siblings)


def test_merkle_errors(P2, pkg, O, rng):
    leaves = O.random_field(rng, (16, 3))
    with pytest.raises(pkg.NlcError, match="cap_height"):
        P2.MerkleTree.new(leaves, 5)
    with pytest.raises(pkg.NlcError, match="power of two"):
        P2.MerkleTree.new(leaves[:12], 0)
    # the context stays usable after an error
    assert np.array_equal(P2.MerkleTree.new(leaves, 2).cap, O.merkle_new(leaves, 2)[1])


def test_merkle_wide_and_tall(P2, O, rng):
    leaves = O.random_field(rng, (1 << 15, 400))
    t = P2.MerkleTree.new(leaves, 4)
    digests, cap = O.merkle_new(leaves, 4)
    assert np.array_equal(t.cap, cap) and np.array_equal(t.digests, digests)

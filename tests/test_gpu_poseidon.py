"""a4: PoseidonPermutation / PoseidonHash on the device vs upstream KATs and the oracle (bit-exact)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 0xFFFFFFFF00000001

# plonky2 poseidon_goldilocks.rs test vectors
KATS = [
    ([0] * 12, [0x3c18a9786cb0b359, 0xc4055e3364a246c3, 0x7953db0ab48808f4, 0xc71603f33a1144ca, 0xd7709673896996dc,
                0x46a84e87642f44ed, 0xd032648251ee0b3c, 0x1c687363b207df62, 0xdf8565563e8045fe, 0x40f5b37ff4254dae,
                0xd070f637b431067c, 0x1792b1c4342109d7]),
    (list(range(12)), [0xd64e1e3efc5b8e9e, 0x53666633020aaa47, 0xd40285597c6a8825, 0x613a4f81e81231d2,
                       0x414754bfebd051f0, 0xcb1f8980294a023f, 0x6eb2a9e4d54a9d0f, 0x1902bc3af467e056,
                       0xf045d5eafdc6021f, 0xe4150f77caaa3be5, 0xc9bfd01d39b50cce, 0x5c0a27fcb0e1459b]),
    ([P - 1] * 12, [0xbe0085cfc57a8357, 0xd95af71847d05c09, 0xcf55a13d33c1c953, 0x95803a74f4530e82,
                    0xfcd99eb30a135df1, 0xe095905e913a3029, 0xde0392461b42919b, 0x7d3260e24e81d031,
                    0x10d3d0465d9deaa0, 0xa87571083dfc2a47, 0xe18263681e9958f8, 0xe28e96f1ae5e60d3]),
]


def test_permutation_kats(P2):
    for inp, out in KATS:
        got = P2.PoseidonPermutation.permute(np.array(inp, np.uint64))
        assert [int(x) for x in got] == out


def test_permutation_random_and_noncanonical(P2, O, rng):
    s = O.random_field(rng, (1000, 12))
    s[0, :] = np.uint64(P - 1)
    s[1, :] = np.uint64(0xFFFFFFFFFFFFFFFF)  # non-canonical representative of 2^32 - 2
    s[2, 3] = np.uint64(P)                   # == 0
    got = P2.PoseidonPermutation.permute(s)
    want = np.stack([O.poseidon_permute(np.where(r >= np.uint64(P), r - np.uint64(P), r)) for r in s])
    assert np.array_equal(got, want)
    assert (got < np.uint64(P)).all()


def test_permutation_many_random_and_limb_edges(P2, O):
    """The partial rounds keep lanes 1..11 as loosely normalised 22/21/21-bit limbs (poseidon.cuh): rare carry
    patterns (a low limb dipping below zero: ~3 % of permutations) must still be exact -- 2^16 random states plus
    states built from limb-boundary patterns."""
    rng = np.random.default_rng(0xBADC0DE)
    s = O.random_field(rng, (1 << 16, 12))
    edges = np.array([0, 1, P - 1, P - 2, 0x3FFFFF, 0x400000, (0x1FFFFF << 22) | 0x3FFFFF, 1 << 43, (1 << 43) - 1,
                      0xFFFFFFFF, 0xFFFFFFFF00000000, 0x7FFFFFFFFFFFFFFF, 0x8000000000000000], np.uint64)
    s[:4096] = edges[rng.integers(0, len(edges), (4096, 12))]
    got = P2.PoseidonPermutation.permute(s)
    want = np.stack([O.poseidon_permute(r) for r in s])
    assert np.array_equal(got, want)


@pytest.mark.parametrize("length", [1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 32, 135, 400])
def test_hash_no_pad_and_or_noop(P2, O, rng, length):
    x = O.random_field(rng, (64, length))
    got = P2.PoseidonHash.hash_no_pad(x)
    assert np.array_equal(got, np.stack([O.hash_no_pad(r) for r in x]))
    got = P2.PoseidonHash.hash_or_noop(x)
    assert np.array_equal(got, np.stack([O.hash_or_noop(r) for r in x]))


def test_hash_derived_vectors(P2):
    # SURVEY.md A.4 derived vectors (oracle-pinned)
    assert [int(v) for v in P2.PoseidonHash.hash_no_pad(np.arange(135, dtype=np.uint64))] == [
        0x4347cfca7c42dd67, 0x6ed10450bc48d3e5, 0x29580eaeee65c3f9, 0x9b2bcdeeea94203c]
    assert [int(v) for v in P2.PoseidonHash.two_to_one([1, 2, 3, 4], [5, 6, 7, 8])] == [
        0xd110aa6a46373941, 0x8f238fcceb658894, 0x9cd4f8353866fb4f, 0x274913f0007aa232]


def test_two_to_one(P2, O, rng):
    l, r = O.random_field(rng, (300, 4)), O.random_field(rng, (300, 4))
    got = P2.PoseidonHash.two_to_one(l, r)
    assert np.array_equal(got, np.stack([O.two_to_one(a, b) for a, b in zip(l, r)]))


def test_challenger(P2, O, rng):
    ch, och = P2.Challenger(), O.Challenger()
    for step in range(40):
        k = int(rng.integers(0, 20))
        x = O.random_field(rng, (k,))
        ch.observe_elements(x)
        och.observe_elements(x)
        m = int(rng.integers(0, 11))
        assert ch.get_n_challenges(m) == och.get_n_challenges(m)
        assert ch.state() == och.state()


@pytest.mark.parametrize("seed", range(4))
def test_challenger_long_observations(P2, O, seed):
    """observe_elements takes whole blocks of 8 on a path of its own once the input buffer is empty (an opening set is
    thousands of elements): runs that start and end inside, at and across block boundaries, non-canonical inputs,
    interleaved with challenges -- the state after every step equals the oracle's."""
    r = np.random.default_rng(seed)
    el = r.integers(0, 0xFFFFFFFFFFFFFFFF, 4096, dtype=np.uint64, endpoint=True)
    ch, och, pos = P2.Challenger(), O.Challenger(), 0
    for step in range(14):
        k = int(r.choice([0, 1, 3, 7, 8, 9, 15, 16, 17, 40, 64, 129, 1000]))
        ch.observe_elements(el[pos:pos + k])
        och.observe_elements(el[pos:pos + k])
        pos = (pos + k) % 2048
        assert ch.state() == och.state()
        if r.integers(0, 2):
            m = int(r.integers(1, 12))
            assert ch.get_n_challenges(m) == och.get_n_challenges(m)
    assert ch.get_n_challenges(3) == och.get_n_challenges(3)

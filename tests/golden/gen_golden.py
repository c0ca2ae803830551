#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the CPU oracle (run from the repo root: python tests/golden/gen_golden.py).

The reference holds no golden vector at the plonky2 boundary and plonky2 cannot be built
here, so these fixtures FREEZE the oracle's outputs once it passes the upstream Poseidon KATs and its own
FRI verifier: any later drift of the oracle (or of a kernel together with the oracle) is caught.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

assert O.poseidon_selftest() == 0
out = {}
for ci, (W, log_n, r, cap) in enumerate([(3, 4, 1, 2), (9, 6, 3, 4), (135, 7, 3, 4), (2, 5, 1, 0), (20, 8, 1, 4)]):
    rng = np.random.default_rng(0xC0FFEE + ci)
    v = O.random_field(rng, (W, 1 << log_n))
    b = O.commit_from_values(v, r, cap)
    N, log_N = 1 << (log_n + r), log_n + r
    br = lambda x: int(format(x, f"0{log_N}b")[::-1], 2) if log_N else 0
    key = f"{W}_{log_n}_{r}_{cap}"
    out[key + ":values"] = v
    out[key + ":cap"] = b.cap
    out[key + ":lde_rows"] = np.stack([b.leaves[br(i)] for i in (0, 1, 5)])
    out[key + ":coeffs_head"] = b.coeffs[:, :4].copy()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "commit_small.npz"), **out)

# FRI proof of a starky-shaped opening (trace 5 cols + quotient 2 cols, 2^6 rows)
rng = np.random.default_rng(0xF1)
log_n, r, cap = 6, 1, 2
tv, qc = O.random_field(rng, (5, 64)), O.random_field(rng, (2, 64))
t, q = O.commit_from_values(tv, r, cap), O.commit_from_coeffs(qc, r, cap)
zeta = (0x123456789abcdef, 0xfedcba987654321)
g = pow(1753635133440165772, 1 << (32 - log_n), O.P)
zn = (zeta[0] * g % O.P, zeta[1] * g % O.P)
batches = [(zeta, [(0, i) for i in range(5)] + [(1, i) for i in range(2)]), (zn, [(0, i) for i in range(5)])]
ch = O.Challenger()
ch.observe_cap(t.cap)
ch.observe_cap(q.cap)
params = O.fri_params(log_n, r, cap, 7, 6, 2, 3)
proof = O.prove_openings(O.FriInstance(batches), [t, q], ch, params)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "fri_small.npz"), trace_values=tv, quotient_coeffs=qc,
                    zeta=np.array(zeta, np.uint64), proof=proof,
                    challenger_state=np.array(list(ch.c.state), np.uint64))

# Poseidon / sponge vectors: upstream KATs (A.3) + oracle-derived sponge outputs (A.4)
np.savez_compressed(
    os.path.join(ROOT, "tests", "golden", "poseidon.npz"),
    perm_zero=O.poseidon_permute(np.zeros(12, np.uint64)),
    perm_iota=O.poseidon_permute(np.arange(12, dtype=np.uint64)),
    perm_neg1=O.poseidon_permute(np.full(12, O.P - 1, np.uint64)),
    hash135=O.hash_no_pad(np.arange(135, dtype=np.uint64)),
    hash7=O.hash_no_pad(np.arange(7, dtype=np.uint64)),
    two_to_one=O.two_to_one([1, 2, 3, 4], [5, 6, 7, 8]),
    round_constants=O.round_constants(),
)
print("golden fixtures written")

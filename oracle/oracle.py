"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  The product package (near-light-client_b200/) never imports this module.

All arrays are numpy uint64, C-contiguous.  Shapes follow plonky2:
values/coeffs ``(W, n)`` (one polynomial per row), leaves ``(N, W)``, digests ``(2*(N-2^cap), 4)`` in
plonky2's interleaved layout, cap ``(2^cap, 4)``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRCS = ["poseidon.c", "fft.c", "merkle.c", "commit.c", "quotient.c", "challenger.c", "fri.c", "vec.c", "oracle.h", "gl.h",
         "poseidon_fast.h"]
P = 0xFFFFFFFF00000001

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)


def build(out: str | None = None, arch: str | None = None) -> str:
    """Compile liboracle.so if missing or older than its sources; returns the path."""
    out = out or os.path.join(_HERE, "liboracle.so")
    newest = max(os.path.getmtime(os.path.join(_HERE, s)) for s in _SRCS)
    if not os.path.exists(out) or os.path.getmtime(out) < newest:
        cmd = ["make", "-C", _HERE, f"OUT={out}"]
        if arch:
            cmd.append(f"ARCH={arch}")
        subprocess.run(cmd, check=True, capture_output=True)
    return out


class FriParams(C.Structure):
    _fields_ = [("degree_bits", C.c_int), ("rate_bits", C.c_int), ("cap_height", C.c_int), ("pow_bits", C.c_int),
                ("num_queries", C.c_int), ("n_rounds", C.c_int), ("arity_bits", C.c_int * 16)]

    @property
    def reduction_arity_bits(self):
        return [self.arity_bits[i] for i in range(self.n_rounds)]


class _Challenger(C.Structure):
    _fields_ = [("state", C.c_uint64 * 12), ("input", C.c_uint64 * 8), ("n_input", C.c_int),
                ("output", C.c_uint64 * 8), ("n_output", C.c_int)]


class _Batch(C.Structure):
    _fields_ = [("W", C.c_size_t), ("coeffs", u64p), ("leaves", u64p), ("digests", u64p), ("cap", u64p)]


class _Instance(C.Structure):
    _fields_ = [("n_batches", C.c_int), ("points", u64p), ("batch_offsets", u32p), ("poly_oracle", u32p),
                ("poly_index", u32p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.environ.get("NLC_ORACLE_LIB") or build()
        L = C.CDLL(path)
        L.orc_poseidon_round_constants.restype = u64p
        L.orc_fri_proof_words.restype = C.c_size_t
        L.orc_challenger_get.restype = C.c_uint64
        L.orc_poseidon_selftest.restype = C.c_int
        L.orc_poseidon_init()
        _lib = L
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags.c_contiguous
    return a.ctypes.data_as(u64p)


def _a(x, shape=None) -> np.ndarray:
    a = np.ascontiguousarray(np.asarray(x, dtype=np.uint64))
    return a if shape is None else a.reshape(shape)


def set_threads(n: int):
    lib().orc_set_threads(int(n))


def get_threads() -> int:
    return int(lib().orc_get_threads())


def poseidon_selftest() -> int:
    return int(lib().orc_poseidon_selftest())


def round_constants() -> np.ndarray:
    return np.ctypeslib.as_array(lib().orc_poseidon_round_constants(), shape=(360,)).copy()


def poseidon_permute(state) -> np.ndarray:
    s = _a(state).copy()
    assert s.shape == (12,)
    lib().orc_poseidon_permute(_p(s))
    return s


def hash_no_pad(x) -> np.ndarray:
    x = _a(x)
    out = np.zeros(4, np.uint64)
    lib().orc_hash_no_pad(_p(x), C.c_size_t(x.size), _p(out))
    return out


def hash_or_noop(x) -> np.ndarray:
    x = _a(x)
    out = np.zeros(4, np.uint64)
    lib().orc_hash_or_noop(_p(x), C.c_size_t(x.size), _p(out))
    return out


def two_to_one(l, r) -> np.ndarray:
    out = np.zeros(4, np.uint64)
    lib().orc_two_to_one(_p(_a(l)), _p(_a(r)), _p(out))
    return out


def _log2(n: int) -> int:
    assert n > 0 and n & (n - 1) == 0
    return n.bit_length() - 1


def fft(a) -> np.ndarray:
    a = _a(a).copy()
    lib().orc_fft(_p(a), _log2(a.size))
    return a


def ifft(a) -> np.ndarray:
    a = _a(a).copy()
    lib().orc_ifft(_p(a), _log2(a.size))
    return a


def coset_fft(a, shift=7) -> np.ndarray:
    a = _a(a).copy()
    lib().orc_coset_fft(_p(a), _log2(a.size), C.c_uint64(shift))
    return a


def coset_ifft(a, shift=7) -> np.ndarray:
    a = _a(a).copy()
    lib().orc_coset_ifft(_p(a), _log2(a.size), C.c_uint64(shift))
    return a


def lde_onto_coset(values, rate_bits: int) -> np.ndarray:
    v = _a(values)
    out = np.zeros(v.size << rate_bits, np.uint64)
    lib().orc_lde_onto_coset(_p(v), _log2(v.size), rate_bits, _p(out))
    return out


def merkle_new(leaves, cap_height: int):
    leaves = _a(leaves)
    n, ll = leaves.shape
    digests = np.zeros((2 * (n - (1 << cap_height)), 4), np.uint64)
    cap = np.zeros((1 << cap_height, 4), np.uint64)
    lib().orc_merkle_new(_p(leaves), C.c_size_t(n), C.c_size_t(ll), cap_height, _p(digests), _p(cap))
    return digests, cap


def merkle_prove(digests, n_leaves: int, cap_height: int, leaf_index: int) -> np.ndarray:
    sib = np.zeros((_log2(n_leaves) - cap_height, 4), np.uint64)
    lib().orc_merkle_prove(_p(_a(digests)), C.c_size_t(n_leaves), cap_height, C.c_size_t(leaf_index), _p(sib))
    return sib


def merkle_verify(leaf, leaf_index: int, cap, siblings) -> bool:
    leaf, cap, siblings = _a(leaf), _a(cap), _a(siblings)
    return lib().orc_merkle_verify(_p(leaf), C.c_size_t(leaf.size), C.c_size_t(leaf_index), _p(cap),
                                   _log2(cap.shape[0]), _p(siblings), siblings.shape[0]) == 0


@dataclass
class Batch:
    """Host image of a plonky2 PolynomialBatch (blinding = false)."""
    coeffs: np.ndarray
    leaves: np.ndarray
    digests: np.ndarray
    cap: np.ndarray
    log_n: int
    rate_bits: int
    cap_height: int

    @property
    def W(self):
        return self.coeffs.shape[0]


def commit_from_values(values, rate_bits: int, cap_height: int) -> Batch:
    values = _a(values)
    W, n = values.shape
    log_n = _log2(n)
    N = n << rate_bits
    coeffs = np.zeros((W, n), np.uint64)
    leaves = np.zeros((N, W), np.uint64)
    digests = np.zeros((2 * (N - (1 << cap_height)), 4), np.uint64)
    cap = np.zeros((1 << cap_height, 4), np.uint64)
    lib().orc_commit_from_values(_p(values), C.c_size_t(W), log_n, rate_bits, cap_height, _p(coeffs), _p(leaves),
                                 _p(digests), _p(cap))
    return Batch(coeffs, leaves, digests, cap, log_n, rate_bits, cap_height)


def commit_from_coeffs(coeffs, rate_bits: int, cap_height: int) -> Batch:
    coeffs = _a(coeffs)
    W, n = coeffs.shape
    log_n = _log2(n)
    N = n << rate_bits
    leaves = np.zeros((N, W), np.uint64)
    digests = np.zeros((2 * (N - (1 << cap_height)), 4), np.uint64)
    cap = np.zeros((1 << cap_height, 4), np.uint64)
    lib().orc_commit_from_coeffs(_p(coeffs), C.c_size_t(W), log_n, rate_bits, cap_height, _p(leaves), _p(digests),
                                 _p(cap))
    return Batch(coeffs.copy(), leaves, digests, cap, log_n, rate_bits, cap_height)


def quotient_domain(log_n: int, qdb: int):
    """(coset_x, lagrange_first, lagrange_last, zh_inv) of starkyx prover.rs:278-302 (oracle/quotient.c)."""
    size = 1 << (log_n + qdb)
    x, lf, ll = (np.zeros(size, np.uint64) for _ in range(3))
    zh = np.zeros(1 << qdb, np.uint64)
    lib().orc_quotient_domain(log_n, qdb, _p(x), _p(lf), _p(ll), _p(zh))
    return x, lf, ll, zh


def quotient_commit(evals, log_n: int, qdb: int, qdf: int, divide_by_zh: bool, rate_bits: int, cap_height: int):
    """evals: size x n_challenges row-major.  Returns the Batch of the quotient chunks, or None when
    trim_to_len fails (prover.rs:161-165)."""
    evals = _a(evals)
    size, n_ch = evals.shape
    assert size == 1 << (log_n + qdb)
    n, W = 1 << log_n, n_ch * qdf
    N = n << rate_bits
    coeffs = np.zeros((W, n), np.uint64)
    leaves = np.zeros((N, W), np.uint64)
    digests = np.zeros((2 * (N - (1 << cap_height)), 4), np.uint64)
    cap = np.zeros((1 << cap_height, 4), np.uint64)
    rc = lib().orc_quotient_commit(_p(evals), C.c_size_t(n_ch), log_n, qdb, C.c_size_t(qdf), int(divide_by_zh),
                                   rate_bits, cap_height, _p(coeffs), _p(leaves), _p(digests), _p(cap))
    if rc:
        return None
    return Batch(coeffs, leaves, digests, cap, log_n, rate_bits, cap_height)


class Challenger:
    def __init__(self):
        self.c = _Challenger()
        lib().orc_challenger_init(C.byref(self.c))

    def clone(self) -> "Challenger":
        o = Challenger()
        C.memmove(C.byref(o.c), C.byref(self.c), C.sizeof(_Challenger))
        return o

    def observe_elements(self, x):
        x = _a(x).reshape(-1)
        if x.size:
            lib().orc_challenger_observe(C.byref(self.c), _p(x), C.c_size_t(x.size))

    def observe_cap(self, cap):
        self.observe_elements(cap)

    def get_challenge(self) -> int:
        return int(lib().orc_challenger_get(C.byref(self.c)))

    def get_n_challenges(self, n: int):
        return [self.get_challenge() for _ in range(n)]

    def get_extension_challenge(self):
        return tuple(self.get_n_challenges(2))

    def state(self):
        return (list(self.c.state), list(self.c.input)[: self.c.n_input], list(self.c.output)[: self.c.n_output])


def fri_params(degree_bits, rate_bits, cap_height, pow_bits, num_queries, arity_bits=4, final_poly_bits=5) -> FriParams:
    p = FriParams()
    lib().orc_fri_params_constant_arity(C.byref(p), degree_bits, rate_bits, cap_height, pow_bits, num_queries,
                                        arity_bits, final_poly_bits)
    return p


@dataclass
class FriInstance:
    """FriInstanceInfo: list of (point (a0, a1), [(oracle_index, poly_index), ...])."""
    batches: list
    _keep: list = field(default_factory=list)

    def to_c(self) -> _Instance:
        pts = _a([c for (pt, _) in self.batches for c in pt])
        offs = np.zeros(len(self.batches) + 1, np.uint32)
        po, pi = [], []
        for b, (_, polys) in enumerate(self.batches):
            for (o, i) in polys:
                po.append(o)
                pi.append(i)
            offs[b + 1] = len(po)
        po = np.asarray(po, np.uint32)
        pi = np.asarray(pi, np.uint32)
        self._keep = [pts, offs, po, pi]
        return _Instance(len(self.batches), _p(pts), offs.ctypes.data_as(u32p), po.ctypes.data_as(u32p),
                         pi.ctypes.data_as(u32p))


def _c_batches(oracles):
    arr = (_Batch * len(oracles))()
    for i, b in enumerate(oracles):
        arr[i] = _Batch(b.W, _p(b.coeffs), _p(b.leaves), _p(b.digests), _p(b.cap))
    return arr


def prove_openings(inst: FriInstance, oracles, ch: Challenger, p: FriParams) -> np.ndarray:
    arr = _c_batches(oracles)
    words = lib().orc_fri_proof_words(C.byref(p), arr, len(oracles))
    out = np.zeros(words, np.uint64)
    ci = inst.to_c()
    rc = lib().orc_prove_openings(C.byref(ci), arr, len(oracles), C.byref(ch.c), C.byref(p), _p(out))
    assert rc == 0, f"orc_prove_openings failed: {rc}"
    return out


def fri_verify(inst: FriInstance, openings, oracles, ch: Challenger, p: FriParams, proof) -> int:
    """openings: flat list of extension elements (a0, a1) in instance order, all batches concatenated."""
    op = _a(openings).reshape(-1)
    caps = (u64p * len(oracles))(*[_p(b.cap) for b in oracles])
    Ws = (C.c_size_t * len(oracles))(*[b.W for b in oracles])
    ci = inst.to_c()
    proof = _a(proof)
    return int(lib().orc_fri_verify(C.byref(ci), _p(op), caps, Ws, len(oracles), C.byref(ch.c), C.byref(p),
                                    _p(proof)))


def eval_poly_ext(coeffs, z) -> tuple:
    coeffs = _a(coeffs)
    z = _a(z)
    out = np.zeros(2, np.uint64)
    lib().orc_eval_poly_ext(_p(coeffs), C.c_size_t(coeffs.size), _p(z), _p(out))
    return (int(out[0]), int(out[1]))


def random_field(rng: np.random.Generator, shape) -> np.ndarray:
    """i.i.d. uniform in [0, p): draw u64, reject >= p."""
    a = rng.integers(0, 1 << 64, size=shape, dtype=np.uint64)
    bad = a >= np.uint64(P)
    while bad.any():
        a[bad] = rng.integers(0, 1 << 64, size=int(bad.sum()), dtype=np.uint64)
        bad = a >= np.uint64(P)
    return a

"""Host-side mirror of the plonky2 API surface that the starkyx / plonky2x provers consume on the
commitment hot path, on top of the C ABI (include/nlc_commit.h).  Same names, argument order and
panics-as-exceptions behaviour as the Rust items they stand for, so the parity tests read like the
call sites in /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:50-225:

    PolynomialBatch::from_values / from_coeffs / get_lde_values / prove_openings   (fri/oracle.rs)
    MerkleTree::new / get / prove, MerkleCap                                       (hash/merkle_tree.rs)
    PolynomialValues / PolynomialCoeffs fft helpers                                (field/polynomial)
    PoseidonHash, PoseidonPermutation                                              (hash/poseidon.rs)
    Challenger                                                                     (iop/challenger.rs)
    FriConfig / FriParams / FriReductionStrategy / FriInstanceInfo                 (fri/mod.rs, structure.rs)

All arithmetic runs in the CUDA library; numpy only carries buffers.  Field elements are uint64.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field as _field

import numpy as np

from . import _ffi
from ._ffi import NlcError, u64p, vp

P = 0xFFFFFFFF00000001
COSET_SHIFT = 7  # F::coset_shift() == MULTIPLICATIVE_GROUP_GENERATOR


def _a(x, shape=None) -> np.ndarray:
    a = np.ascontiguousarray(np.asarray(x, dtype=np.uint64))
    return a if shape is None else a.reshape(shape)


def _p(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags.c_contiguous
    return a.ctypes.data_as(u64p)


def _log2_strict(n: int) -> int:
    if n <= 0 or n & (n - 1):
        raise NlcError(f"Not a power of two: {n}")  # plonky2_util::log2_strict panics
    return n.bit_length() - 1


class Context:
    """One nlc_ctx: a (process, GPU) pair with its stream and table cache."""

    def __init__(self, device: int = -1):
        self._h = vp()
        L = _ffi.lib()
        if L.nlc_init(device, C.byref(self._h)) != 0:
            raise NlcError((L.nlc_last_error(None) or b"nlc_init failed").decode())

    def _check(self, rc: int):
        if rc != 0:
            raise NlcError((_ffi.lib().nlc_last_error(self._h) or b"unknown error").decode())

    @property
    def device(self) -> int:
        return _ffi.lib().nlc_device(self._h)

    def set_stream(self, cuda_stream: int | None):
        self._check(_ffi.lib().nlc_set_stream(self._h, vp(cuda_stream or 0)))

    def synchronize(self):
        self._check(_ffi.lib().nlc_synchronize(self._h))

    def wait_stream(self, producer_stream: int | None = None):
        """Order the context's (non-blocking) stream after everything enqueued so far on `producer_stream`
        (a cudaStream_t as int; 0 / None = the legacy default stream).  See the *_dev contract in nlc_commit.h."""
        self._check(_ffi.lib().nlc_wait_stream(self._h, vp(producer_stream or 0)))

    def wait_torch(self):
        """wait_stream on torch's current stream of this device, if torch has been imported and CUDA is up: the
        device-pointer entry points take buffers a torch op may still be producing on another stream."""
        import sys
        torch = sys.modules.get("torch")
        if torch is not None and torch.cuda.is_available() and torch.cuda.is_initialized():
            self.wait_stream(torch.cuda.current_stream(self.device).cuda_stream)

    def cache_stats(self) -> dict:
        """The context's device-buffer cache: hits, driver allocations, time in them, bytes cached."""
        out = np.zeros(4, np.uint64)
        self._check(_ffi.lib().nlc_cache_stats(self._h, _p(out)))
        return {"hits": int(out[0]), "misses": int(out[1]), "miss_us": int(out[2]), "cached_bytes": int(out[3])}

    def launch_count(self) -> int:
        return int(_ffi.lib().nlc_launch_count(self._h))

    def timing_enable(self, on: bool = True):
        self._check(_ffi.lib().nlc_timing_enable(self._h, int(on)))

    def timing_read(self) -> dict:
        t = _ffi.CTiming()
        self._check(_ffi.lib().nlc_timing_read(self._h, C.byref(t)))
        return {"IFFT": t.ifft_ms, "FFT + blinding": t.lde_ms, "leaf hashing": t.leaf_hash_ms,
                "digest levels": t.merkle_levels_ms, "commits": int(t.commits)}

    def close(self):
        if self._h:
            _ffi.lib().nlc_destroy(self._h)
            self._h = vp()


_default_ctx: Context | None = None


def default_context() -> Context:
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context(-1)
    return _default_ctx


def set_default_context(ctx: Context | None):
    global _default_ctx
    _default_ctx = ctx


# ---------------------------------------------------------------------------------------------
# field/polynomial: PolynomialValues / PolynomialCoeffs and the fft helpers (a3)
# ---------------------------------------------------------------------------------------------
def _fft_call(fn_name, data, *extra, ctx=None):
    ctx = ctx or default_context()
    a = _a(data).copy()
    one = a.ndim == 1
    m = a.reshape(1, -1) if one else a
    log_n = _log2_strict(m.shape[1])
    ctx._check(getattr(_ffi.lib(), fn_name)(ctx._h, _p(m), m.shape[0], log_n, *extra))
    return a


def fft(coeffs, ctx=None):
    return _fft_call("nlc_fft", coeffs, ctx=ctx)


def ifft(values, ctx=None):
    return _fft_call("nlc_ifft", values, ctx=ctx)


def coset_fft(coeffs, shift=COSET_SHIFT, ctx=None):
    return _fft_call("nlc_coset_fft", coeffs, C.c_uint64(int(shift)), ctx=ctx)


def coset_ifft(values, shift=COSET_SHIFT, ctx=None):
    return _fft_call("nlc_coset_ifft", values, C.c_uint64(int(shift)), ctx=ctx)


def lde_onto_coset(values, rate_bits: int, ctx=None):
    ctx = ctx or default_context()
    v = _a(values)
    one = v.ndim == 1
    m = v.reshape(1, -1) if one else v
    log_n = _log2_strict(m.shape[1])
    out = np.zeros((m.shape[0], m.shape[1] << rate_bits), np.uint64)
    ctx._check(_ffi.lib().nlc_lde_onto_coset(ctx._h, _p(m), _p(out), m.shape[0], log_n, rate_bits))
    return out[0] if one else out


class PolynomialValues:
    """plonky2_field::polynomial::PolynomialValues (evaluations over the subgroup of order len)."""

    def __init__(self, values):
        self.values = _a(values).reshape(-1)
        _log2_strict(self.values.size)

    @classmethod
    def selector(cls, n: int, index: int):
        v = np.zeros(n, np.uint64)
        v[index] = 1
        return cls(v)

    def __len__(self):
        return self.values.size

    def ifft(self):
        return PolynomialCoeffs(ifft(self.values))

    def coset_ifft(self, shift=COSET_SHIFT):
        return PolynomialCoeffs(coset_ifft(self.values, shift))

    def lde_onto_coset(self, rate_bits: int):
        return PolynomialValues(lde_onto_coset(self.values, rate_bits))


class PolynomialCoeffs:
    """plonky2_field::polynomial::PolynomialCoeffs (natural coefficient order)."""

    def __init__(self, coeffs):
        self.coeffs = _a(coeffs).reshape(-1)

    def __len__(self):
        return self.coeffs.size

    def lde(self, rate_bits: int):
        out = np.zeros(self.coeffs.size << rate_bits, np.uint64)
        out[: self.coeffs.size] = self.coeffs
        return PolynomialCoeffs(out)

    def fft(self):
        return PolynomialValues(fft(self.coeffs))

    def coset_fft(self, shift=COSET_SHIFT):
        return PolynomialValues(coset_fft(self.coeffs, shift))


# ---------------------------------------------------------------------------------------------
# hash/poseidon.rs, hashing.rs (a4)
# ---------------------------------------------------------------------------------------------
class PoseidonPermutation:
    WIDTH, RATE = 12, 8

    @staticmethod
    def permute(states, ctx=None):
        """PoseidonPermutation::permute on one (12,) or many (k, 12) states."""
        ctx = ctx or default_context()
        s = _a(states).copy()
        m = s.reshape(-1, 12)
        ctx._check(_ffi.lib().nlc_poseidon_permute(ctx._h, _p(m), m.shape[0]))
        return s


class PoseidonHash:
    @staticmethod
    def _many(fn, inputs, ctx):
        ctx = ctx or default_context()
        x = _a(inputs)
        one = x.ndim == 1
        m = x.reshape(1, -1) if one else x
        out = np.zeros((m.shape[0], 4), np.uint64)
        ctx._check(getattr(_ffi.lib(), fn)(ctx._h, _p(m), m.shape[0], m.shape[1], _p(out)))
        return out[0] if one else out

    @staticmethod
    def hash_no_pad(inputs, ctx=None):
        return PoseidonHash._many("nlc_hash_no_pad", inputs, ctx)

    @staticmethod
    def hash_or_noop(inputs, ctx=None):
        return PoseidonHash._many("nlc_hash_or_noop", inputs, ctx)

    @staticmethod
    def two_to_one(left, right, ctx=None):
        ctx = ctx or default_context()
        l, r = _a(left), _a(right)
        one = l.ndim == 1
        lm, rm = l.reshape(-1, 4), r.reshape(-1, 4)
        out = np.zeros_like(lm)
        ctx._check(_ffi.lib().nlc_two_to_one(ctx._h, _p(lm), _p(rm), lm.shape[0], _p(out)))
        return out[0] if one else out


# ---------------------------------------------------------------------------------------------
# hash/merkle_tree.rs (a5)
# ---------------------------------------------------------------------------------------------
@dataclass
class MerkleProof:
    siblings: np.ndarray  # (log2 N - cap_height, 4), bottom-up


class MerkleTree:
    """plonky2::hash::merkle_tree::MerkleTree living in HBM; fields are fetched on first use."""

    def __init__(self, ctx: Context, handle, n_leaves: int, leaf_len: int, cap_height: int, leaves=None):
        self._ctx, self._h = ctx, handle
        self.n_leaves, self.leaf_len, self.cap_height = n_leaves, leaf_len, cap_height
        self._leaves, self._cap, self._digests = leaves, None, None

    @classmethod
    def new(cls, leaves, cap_height: int, ctx=None):
        ctx = ctx or default_context()
        lv = _a(leaves)
        if lv.ndim != 2:
            raise NlcError("MerkleTree::new: leaves must be a (n_leaves, leaf_len) matrix")
        h = vp()
        ctx._check(_ffi.lib().nlc_merkle_new(ctx._h, _p(lv), lv.shape[0], lv.shape[1], cap_height, C.byref(h)))
        return cls(ctx, h, lv.shape[0], lv.shape[1], cap_height, lv)

    @property
    def leaves(self):
        return self._leaves

    @property
    def cap(self) -> np.ndarray:
        if self._cap is None:
            out = np.zeros((1 << self.cap_height, 4), np.uint64)
            self._ctx._check(_ffi.lib().nlc_tree_cap(self._h, _p(out)))
            self._cap = out
        return self._cap

    @property
    def digests(self) -> np.ndarray:
        if self._digests is None:
            out = np.zeros((2 * (self.n_leaves - (1 << self.cap_height)), 4), np.uint64)
            self._ctx._check(_ffi.lib().nlc_tree_digests(self._h, _p(out)))
            self._digests = out
        return self._digests

    def get(self, i: int) -> np.ndarray:
        out = np.zeros(self.leaf_len, np.uint64)
        self._ctx._check(_ffi.lib().nlc_tree_open(self._h, i, _p(out), None))
        return out

    def prove(self, leaf_index: int) -> MerkleProof:
        sib = np.zeros((_log2_strict(self.n_leaves) - self.cap_height, 4), np.uint64)
        self._ctx._check(_ffi.lib().nlc_tree_open(self._h, leaf_index, None, _p(sib)))
        return MerkleProof(sib)

    def __del__(self):
        try:
            if self._h and self._ctx._h:  # a handle must not be touched once its context is gone
                _ffi.lib().nlc_tree_free(self._h)
            self._h = None
        except Exception:
            pass


class _BatchTree:
    """The `merkle_tree` field of a PolynomialBatch (same accessors as MerkleTree, lazily materialised)."""

    def __init__(self, state: "_BatchState"):
        # the tree shares the batch's STATE (device handle + shape), not the batch: batch <-> tree would be a reference
        # cycle and the device buffers would wait for the cyclic garbage collector; the handle is freed when the last
        # of {batch, tree} goes, so `PolynomialBatch.from_values(..).merkle_tree.cap` works on a temporary as in Rust
        self._b = state
        self._cap = self._local_cap = self._leaves = self._digests = None

    @property
    def local_cap(self) -> np.ndarray:
        """This shard's cap entries (the whole cap when the batch is not sharded)."""
        if self._local_cap is None:
            out = np.zeros(((1 << self._b.cap_height) // self._b.n_shards, 4), np.uint64)
            self._b._ctx._check(_ffi.lib().nlc_batch_cap(self._b._h, _p(out)))
            self._local_cap = out
        return self._local_cap

    @property
    def cap(self) -> np.ndarray:
        if self._cap is None:
            if self._b.n_shards > 1:
                raise NlcError("sharded batch: the full MerkleCap exists only after sharding.gather_cap()")
            self._cap = self.local_cap
        return self._cap

    @property
    def leaves(self) -> np.ndarray:
        """Rows of this shard's leaf range (all N rows when not sharded)."""
        if self._leaves is None:
            out = np.zeros((self._b.lde_size // self._b.n_shards, self._b.n_polys), np.uint64)
            self._b._ctx._check(_ffi.lib().nlc_batch_leaves(self._b._h, _p(out)))
            self._leaves = out
        return self._leaves

    @property
    def digests(self) -> np.ndarray:
        if self._digests is None:
            out = np.zeros((2 * (self._b.lde_size - (1 << self._b.cap_height)) // self._b.n_shards, 4), np.uint64)
            self._b._ctx._check(_ffi.lib().nlc_batch_digests(self._b._h, _p(out)))
            self._digests = out
        return self._digests

    def get(self, i: int) -> np.ndarray:
        out = np.zeros(self._b.n_polys, np.uint64)
        self._b._ctx._check(_ffi.lib().nlc_batch_open(self._b._h, i, _p(out), None))
        return out

    def prove(self, leaf_index: int) -> MerkleProof:
        sib = np.zeros((self._b.degree_log + self._b.rate_bits - self._b.cap_height, 4), np.uint64)
        self._b._ctx._check(_ffi.lib().nlc_batch_open(self._b._h, leaf_index, None, _p(sib)))
        return MerkleProof(sib)


# ---------------------------------------------------------------------------------------------
# iop/challenger.rs (a9)
# ---------------------------------------------------------------------------------------------
class Challenger:
    """plonky2::iop::challenger::Challenger<F, PoseidonHash>; the struct's fields cross the C ABI as data."""

    def __init__(self, ctx=None):
        self._ctx = ctx or default_context()
        self.c = _ffi.CChallenger()
        _ffi.lib().nlc_challenger_init(C.byref(self.c))

    def clone(self) -> "Challenger":
        o = Challenger(self._ctx)
        C.memmove(C.byref(o.c), C.byref(self.c), C.sizeof(_ffi.CChallenger))
        return o

    def observe_element(self, e):
        self.observe_elements([e])

    def observe_elements(self, elems):
        x = _a(elems).reshape(-1)
        if x.size:
            self._ctx._check(_ffi.lib().nlc_challenger_observe(self._ctx._h, C.byref(self.c), _p(x), x.size))

    def observe_extension_element(self, e):
        self.observe_elements(e)

    def observe_extension_elements(self, es):
        self.observe_elements(es)

    def observe_hash(self, h):
        self.observe_elements(h)

    def observe_cap(self, cap):
        self.observe_elements(cap)

    def get_n_challenges(self, n: int):
        out = np.zeros(n, np.uint64)
        if n:
            self._ctx._check(_ffi.lib().nlc_challenger_get(self._ctx._h, C.byref(self.c), n, _p(out)))
        return [int(v) for v in out]

    def get_challenge(self) -> int:
        return self.get_n_challenges(1)[0]

    def get_extension_challenge(self):
        return tuple(self.get_n_challenges(2))

    def state(self):
        c = self.c
        return (list(c.sponge_state), list(c.input_buffer)[: c.input_len], list(c.output_buffer)[: c.output_len])


# ---------------------------------------------------------------------------------------------
# fri/mod.rs, fri/reduction_strategies.rs, fri/structure.rs (a10)
# ---------------------------------------------------------------------------------------------
@dataclass
class FriReductionStrategy:
    kind: str
    arity_bits: int = 4
    final_poly_bits: int = 5

    @classmethod
    def ConstantArityBits(cls, arity_bits: int, final_poly_bits: int):
        return cls("ConstantArityBits", arity_bits, final_poly_bits)


@dataclass
class FriParams:
    config: "FriConfig"
    hiding: bool
    degree_bits: int
    reduction_arity_bits: list

    def lde_bits(self):
        return self.degree_bits + self.config.rate_bits

    def lde_size(self):
        return 1 << self.lde_bits()

    def final_poly_len(self):
        return 1 << (self.degree_bits - sum(self.reduction_arity_bits))

    def to_c(self) -> _ffi.CFriParams:
        p = _ffi.CFriParams()
        p.degree_bits, p.rate_bits, p.cap_height = self.degree_bits, self.config.rate_bits, self.config.cap_height
        p.proof_of_work_bits, p.num_query_rounds = self.config.proof_of_work_bits, self.config.num_query_rounds
        p.n_reduction_rounds = len(self.reduction_arity_bits)
        for i, a in enumerate(self.reduction_arity_bits):
            p.reduction_arity_bits[i] = a
        return p


@dataclass
class FriConfig:
    rate_bits: int
    cap_height: int
    proof_of_work_bits: int
    reduction_strategy: FriReductionStrategy
    num_query_rounds: int

    @classmethod
    def standard_fast_config(cls):
        """starkyx StarkyConfig::standard_fast_config (plonky2/stark/config.rs:58-73)."""
        return cls(1, 4, 16, FriReductionStrategy.ConstantArityBits(4, 5), 84)

    @classmethod
    def standard_recursion_config(cls):
        """plonky2 CircuitConfig::standard_recursion_config().fri_config."""
        return cls(3, 4, 16, FriReductionStrategy.ConstantArityBits(4, 5), 28)

    def fri_params(self, degree_bits: int, hiding: bool = False) -> FriParams:
        if hiding:
            raise NlcError("hiding / blinding is not used on this path (all in-tree call sites pass false)")
        if self.reduction_strategy.kind != "ConstantArityBits":
            raise NlcError("only FriReductionStrategy::ConstantArityBits is used in tree (config.rs:68)")
        p = _ffi.CFriParams()
        rc = _ffi.lib().nlc_fri_params_constant_arity(C.byref(p), degree_bits, self.rate_bits, self.cap_height,
                                                      self.proof_of_work_bits, self.num_query_rounds,
                                                      self.reduction_strategy.arity_bits,
                                                      self.reduction_strategy.final_poly_bits)
        if rc:
            raise NlcError("bad FriConfig (arity_bits in 1..8; at most 32 reduction rounds)")
        return FriParams(self, hiding, degree_bits, [p.reduction_arity_bits[i] for i in range(p.n_reduction_rounds)])


@dataclass
class FriPolynomialInfo:
    oracle_index: int
    polynomial_index: int

    @staticmethod
    def from_range(oracle_index: int, polynomial_indices):
        return [FriPolynomialInfo(oracle_index, i) for i in polynomial_indices]


@dataclass
class FriBatchInfo:
    point: tuple  # extension element (a0, a1)
    polynomials: list


@dataclass
class FriInstanceInfo:
    oracles: list  # FriOracleInfo is only (num_polys, blinding); kept for shape parity
    batches: list
    _keep: list = _field(default_factory=list)

    def to_c(self) -> _ffi.CFriInstance:
        pts = _a([c for b in self.batches for c in b.point])
        offs = np.zeros(len(self.batches) + 1, np.uint32)
        po, pi = [], []
        for k, b in enumerate(self.batches):
            for info in b.polynomials:
                po.append(info.oracle_index)
                pi.append(info.polynomial_index)
            offs[k + 1] = len(po)
        po, pi = np.asarray(po, np.uint32), np.asarray(pi, np.uint32)
        self._keep = [pts, offs, po, pi]
        u32 = _ffi.u32p
        return _ffi.CFriInstance(len(self.batches), _p(pts) if pts.size else None, offs.ctypes.data_as(u32),
                                 po.ctypes.data_as(u32), pi.ctypes.data_as(u32))


@dataclass
class FriProof:
    """fri::proof::FriProof, parsed from the flat layout documented in include/nlc_commit.h."""
    commit_phase_merkle_caps: list
    final_poly: np.ndarray       # (len, 2)
    pow_witness: int
    query_round_proofs: list     # per query: {"initial": [(leaf, siblings)], "steps": [(evals, siblings)]}
    flat: np.ndarray

    @classmethod
    def parse(cls, flat: np.ndarray, params: FriParams, oracle_widths):
        cfg = params.config
        cap_words = 4 << cfg.cap_height
        w = 0
        caps = []
        for _ in params.reduction_arity_bits:
            caps.append(flat[w:w + cap_words].reshape(-1, 4))
            w += cap_words
        fl = params.final_poly_len()
        final_poly = flat[w:w + 2 * fl].reshape(-1, 2)
        w += 2 * fl
        pow_witness = int(flat[w])
        w += 1
        log_N = params.lde_bits()
        queries = []
        for _ in range(cfg.num_query_rounds):
            init, steps = [], []
            for W in oracle_widths:
                leaf = flat[w:w + W]
                w += W
                ns = 4 * (log_N - cfg.cap_height)
                init.append((leaf, flat[w:w + ns].reshape(-1, 4)))
                w += ns
            lg = log_N
            for ab in params.reduction_arity_bits:
                ev = flat[w:w + (2 << ab)].reshape(-1, 2)
                w += 2 << ab
                ns = 4 * (lg - ab - cfg.cap_height)
                steps.append((ev, flat[w:w + ns].reshape(-1, 4)))
                w += ns
                lg -= ab
            queries.append({"initial": init, "steps": steps})
        assert w == flat.size
        return cls(caps, final_poly, pow_witness, queries, flat)


# ---------------------------------------------------------------------------------------------
# fri/oracle.rs: PolynomialBatch (a1, a2, a6, a7)
# ---------------------------------------------------------------------------------------------
def _col_table(cols):
    """W host pointers, exactly a Vec<PolynomialValues<F>> / Vec<PolynomialCoeffs<F>>."""
    arrs = []
    for c in cols:
        v = c.values if isinstance(c, PolynomialValues) else c.coeffs if isinstance(c, PolynomialCoeffs) else c
        arrs.append(_a(v).reshape(-1))
    if not arrs:
        raise NlcError("PolynomialBatch: at least one polynomial required")
    n = arrs[0].size
    if any(a.size != n for a in arrs):
        raise NlcError("All polynomials in a batch must have the same length")  # plonky2 assert
    table = (u64p * len(arrs))(*[_p(a) for a in arrs])
    return table, arrs, n


def quotient_domain(degree_log: int, quotient_degree_bits: int, ctx=None):
    """The tables starkyx's quotient_polys builds before its evaluation loop (prover.rs:278-302):
    (coset, lagrange_first, lagrange_last, z_h_inverses) with coset[i] = g * w^i over the quotient domain and
    ZeroPolyOnCoset::eval_inverse(i) = z_h_inverses[i % 2^quotient_degree_bits]."""
    ctx = ctx or default_context()
    size = 1 << (degree_log + quotient_degree_bits)
    x, lf, ll = (np.zeros(size, np.uint64) for _ in range(3))
    zh = np.zeros(1 << quotient_degree_bits, np.uint64)
    ctx._check(_ffi.lib().nlc_quotient_domain(ctx._h, degree_log, quotient_degree_bits, _p(x), _p(lf), _p(ll), _p(zh)))
    return x, lf, ll, zh


class _BatchState:
    """What a PolynomialBatch and its `merkle_tree` view share: the device handle and the shape."""

    def __init__(self, ctx, handle, n_polys, degree_log, rate_bits, cap_height):
        self._ctx, self._h = ctx, handle
        self.n_polys, self.degree_log, self.rate_bits, self.cap_height = n_polys, degree_log, rate_bits, cap_height
        self.shard_idx, self.n_shards = 0, 1

    @property
    def lde_size(self) -> int:
        return 1 << (self.degree_log + self.rate_bits)

    def free(self):
        if self._h and self._ctx._h:  # a handle must not be touched once its context is gone
            _ffi.lib().nlc_batch_free(self._h)
        self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def _shared(name):
    return property(lambda self: getattr(self._s, name), lambda self, v: setattr(self._s, name, v))


class PolynomialBatch:
    """plonky2::fri::oracle::PolynomialBatch resident in HBM.  Struct fields (`polynomials`,
    `merkle_tree.{cap,leaves,digests}`) are materialised on first access."""
    _ctx, _h = _shared("_ctx"), _shared("_h")
    n_polys, degree_log, rate_bits, cap_height = (_shared(k) for k in ("n_polys", "degree_log", "rate_bits", "cap_height"))
    shard_idx, n_shards = _shared("shard_idx"), _shared("n_shards")

    def __init__(self, ctx: Context, handle, n_polys: int, degree_log: int, rate_bits: int, cap_height: int):
        self._s = _BatchState(ctx, handle, n_polys, degree_log, rate_bits, cap_height)
        self.blinding = False
        self.merkle_tree = _BatchTree(self._s)
        self._polys = None

    # -- constructors ---------------------------------------------------------------------------
    @classmethod
    def _commit(cls, fn, cols, rate_bits, blinding, cap_height, ctx):
        if blinding:
            raise NlcError("blinding = true is not used on this path (prover.rs:91-93,171-178 pass false)")
        ctx = ctx or default_context()
        table, arrs, n = _col_table(cols)
        log_n = _log2_strict(n)
        h = vp()
        ctx._check(getattr(_ffi.lib(), fn)(ctx._h, table, len(arrs), log_n, rate_bits, cap_height, C.byref(h)))
        return cls(ctx, h, len(arrs), log_n, rate_bits, cap_height)

    @classmethod
    def from_values(cls, values, rate_bits: int, blinding: bool, cap_height: int, timing=None, fft_root_table=None,
                    ctx=None):
        """PolynomialBatch::from_values(values, rate_bits, blinding, cap_height, timing, fft_root_table)."""
        return cls._commit("nlc_commit_from_values", values, rate_bits, blinding, cap_height, ctx)

    @classmethod
    def from_values_materialised(cls, values, rate_bits: int, blinding: bool, cap_height: int,
                                 fields=("polynomials", "leaves", "digests", "cap"), ctx=None):
        """from_values that also fills the host images of the struct fields a patched plonky2 keeps -- `polynomials`
        (proof.rs:307-310 reads them), `merkle_tree.leaves`, `merkle_tree.digests`, `merkle_tree.cap` -- while the
        commit runs (nlc_commit_from_values_materialised): the device->host copies ride under the hashing instead
        of following it.  `fields` picks which images to fill; the others stay lazy."""
        if blinding:
            raise NlcError("blinding = true is not used on this path (prover.rs:91-93,171-178 pass false)")
        ctx = ctx or default_context()
        table, arrs, n = _col_table(values)
        W, log_n = len(arrs), _log2_strict(n)
        N = n << rate_bits
        if not 0 <= cap_height <= log_n + rate_bits:
            raise NlcError("cap_height should be at most log2(leaves.len())")
        coeffs = np.zeros((W, n), np.uint64) if "polynomials" in fields else None
        leaves = np.zeros((N, W), np.uint64) if "leaves" in fields else None
        digests = np.zeros((2 * (N - (1 << cap_height)), 4), np.uint64) if "digests" in fields else None
        cap = np.zeros((1 << cap_height, 4), np.uint64) if "cap" in fields else None
        ctab = (u64p * W)(*[_p(coeffs[j]) for j in range(W)]) if coeffs is not None else None
        h = vp()
        ctx._check(_ffi.lib().nlc_commit_from_values_materialised(
            ctx._h, table, W, log_n, rate_bits, cap_height, ctab, _p(leaves) if leaves is not None else None,
            _p(digests) if digests is not None and digests.size else None, _p(cap) if cap is not None else None,
            C.byref(h)))
        b = cls(ctx, h, W, log_n, rate_bits, cap_height)
        if coeffs is not None:
            b._coeff_matrix = coeffs
            b._polys = [PolynomialCoeffs(coeffs[j]) for j in range(W)]
        b.merkle_tree._leaves, b.merkle_tree._digests = leaves, digests
        if cap is not None:
            b.merkle_tree._cap = b.merkle_tree._local_cap = cap
        return b

    @classmethod
    def from_coeffs(cls, polynomials, rate_bits: int, blinding: bool, cap_height: int, timing=None,
                    fft_root_table=None, ctx=None):
        """PolynomialBatch::from_coeffs(polynomials, rate_bits, blinding, cap_height, timing, fft_root_table)."""
        return cls._commit("nlc_commit_from_coeffs", polynomials, rate_bits, blinding, cap_height, ctx)

    @classmethod
    def from_values_sharded(cls, values, rate_bits: int, blinding: bool, cap_height: int, shard_idx=None,
                            n_shards=None, group=None, ctx=None):
        """this rank's coset shard of from_values, then the all-gather of the cap entries.

        With torch.distributed initialised, shard_idx / n_shards default to the rank / world size of `group` and
        the returned batch's merkle_tree.cap is the full MerkleCap; called with explicit shard_idx / n_shards and no
        process group it returns the bare shard (merkle_tree.local_cap)."""
        from . import sharding
        if blinding:
            raise NlcError("blinding = true is not used on this path (prover.rs:91-93,171-178 pass false)")
        distributed = shard_idx is None
        if distributed:
            shard_idx, n_shards = sharding.rank_and_world(group)
        sharding.check_plan(rate_bits, cap_height, n_shards)
        ctx = ctx or default_context()
        table, arrs, n = _col_table(values)
        log_n = _log2_strict(n)
        h = vp()
        ctx._check(_ffi.lib().nlc_commit_from_values_shard(ctx._h, table, len(arrs), log_n, rate_bits, cap_height,
                                                           shard_idx, n_shards, C.byref(h)))
        b = cls(ctx, h, len(arrs), log_n, rate_bits, cap_height)
        b.shard_idx, b.n_shards = shard_idx, n_shards
        if distributed:
            b.merkle_tree._cap = sharding.gather_cap(b.merkle_tree.local_cap, cap_height, group, device=ctx.device)
        return b

    @staticmethod
    def ifft_columns_bcast(src, col0: int, peer_ptrs, ctx=None):
        """Step 1 of the coset-sharded commit without replication: src = (nc, n) int64 CUDA tensor, this
        rank's trace columns [col0, col0 + nc); their coefficients are stored into EVERY buffer of `peer_ptrs` (each a
        device pointer to an n_cols_total x n matrix; remote ones over NVLink).  Asynchronous."""
        ctx = ctx or default_context()
        nc, n = src.shape
        table = (vp * len(peer_ptrs))(*[vp(p) for p in peer_ptrs])
        ctx.wait_torch()
        ctx._check(_ffi.lib().nlc_ifft_columns_bcast_dev(ctx._h, vp(src.data_ptr()), nc, _log2_strict(n), col0,
                                                         len(peer_ptrs), table))

    @classmethod
    def from_coeffs_shard_dev(cls, coeffs_ptr: int, n_polys: int, degree_log: int, rate_bits: int, cap_height: int,
                              shard_idx: int, n_shards: int, ctx=None, keepalive=None):
        """Step 2: this rank's coset shard over the gathered coefficient matrix (read in place)."""
        ctx = ctx or default_context()
        h = vp()
        ctx.wait_torch()
        ctx._check(_ffi.lib().nlc_commit_from_coeffs_shard_dev(ctx._h, vp(coeffs_ptr), n_polys, degree_log, rate_bits,
                                                               cap_height, shard_idx, n_shards, C.byref(h)))
        b = cls(ctx, h, n_polys, degree_log, rate_bits, cap_height)
        b.shard_idx, b.n_shards = shard_idx, n_shards
        b._keepalive = keepalive
        return b

    @classmethod
    def from_values_coset_sharded(cls, values, rate_bits: int, blinding: bool, cap_height: int, group=None, ctx=None):
        """SURVEY.md 8e.2 as written: rank p uploads and inverse-transforms only ITS columns, the iFFT's stores put the
        coefficients into every rank's matrix over NVLink (no replicated iFFT, no replicated upload, no NCCL
        all-gather of 8*W*n bytes), one barrier in stream order, then every rank extends and hashes its cosets and the
        cap entries are all-gathered.  Every rank passes the same `values`."""
        import torch
        from . import sharding
        if blinding:
            raise NlcError("blinding = true is not used on this path (prover.rs:91-93,171-178 pass false)")
        g, G = sharding.rank_and_world(group)
        sharding.check_plan(rate_bits, cap_height, G)
        ctx = ctx or default_context()
        table, arrs, n = _col_table(values)
        W, log_n = len(arrs), _log2_strict(n)
        per, split = sharding.uniform_column_split(W, G)
        c0, nc = split[g]
        dev = torch.device("cuda", ctx.device)
        stream = torch.cuda.Stream(dev)
        ctx.set_stream(stream.cuda_stream)
        try:
            with torch.cuda.stream(stream):
                peers = sharding.PeerRows(ctx, G * per * n, group)
                if nc:
                    mine = np.stack([np.asarray(arrs[j]).view(np.int64) for j in range(c0, c0 + nc)])
                    src = torch.from_numpy(mine).pin_memory().to(dev, non_blocking=True)
                    cls.ifft_columns_bcast(src, c0, peers.ptrs, ctx)
                sharding.stream_barrier(dev, group)
                stream.synchronize()
                peers.close_peers()
                b = cls.from_coeffs_shard_dev(peers.own, W, log_n, rate_bits, cap_height, g, G, ctx, keepalive=peers)
                stream.synchronize()
        finally:
            ctx.set_stream(None)
        b.merkle_tree._cap = sharding.gather_cap(b.merkle_tree.local_cap, cap_height, group, device=ctx.device)
        return b

    @staticmethod
    def lde_columns(src, rate_bits: int, is_values: bool = True, ctx=None):
        """Step 1 of the row-sharded commit: src = (nc, n) int64 CUDA tensor holding this rank's
        columns; returns (coeffs (nc, n), lde (nc, N) in leaf order) as CUDA tensors.  Asynchronous."""
        import torch
        ctx = ctx or default_context()
        nc, n = src.shape
        log_n = _log2_strict(n)
        coeffs = torch.empty((nc, n), dtype=torch.int64, device=src.device)
        lde = torch.empty((nc, n << rate_bits), dtype=torch.int64, device=src.device)
        ctx.wait_torch()  # `src` may still be in flight on torch's stream
        ctx._check(_ffi.lib().nlc_lde_columns_dev(ctx._h, vp(src.data_ptr()), nc, log_n, rate_bits, int(is_values),
                                                  vp(coeffs.data_ptr()), vp(lde.data_ptr())))
        return coeffs, lde

    @staticmethod
    def lde_columns_scatter(src, rate_bits: int, n_shards: int, is_values: bool = True, ctx=None, peer_ptrs=None,
                            col0: int = 0):
        """lde_columns with the exchange layout fused into the last NTT pass (nlc_lde_columns_scatter_dev).
        peer_ptrs None: returns (coeffs, send) with send = (n_shards, nc, N / n_shards), the all-to-all's send buffer.
        peer_ptrs = the n_shards receive buffers (sharding.PeerRows.ptrs): stores go to the peers; returns (coeffs, None)."""
        import torch
        ctx = ctx or default_context()
        nc, n = src.shape
        log_n = _log2_strict(n)
        per = (n << rate_bits) // n_shards
        coeffs = torch.empty((nc, n), dtype=torch.int64, device=src.device)
        send, table = None, None
        if peer_ptrs is None:
            send = torch.empty((n_shards, nc, per), dtype=torch.int64, device=src.device)
        else:
            table = (vp * n_shards)(*[vp(p) for p in peer_ptrs])
        ctx.wait_torch()
        ctx._check(_ffi.lib().nlc_lde_columns_scatter_dev(
            ctx._h, vp(src.data_ptr()), nc, log_n, rate_bits, int(is_values), vp(coeffs.data_ptr()),
            vp(send.data_ptr()) if send is not None else None, n_shards, table, col0))
        return coeffs, send

    @classmethod
    def from_lde_rows(cls, rows, degree_log: int, rate_bits: int, cap_height: int, shard_idx: int, n_shards: int,
                      ctx=None):
        """Step 3: the tree over the leaf rows [shard_idx, shard_idx+1) * N / n_shards; rows = (W, N / n_shards) int64
        CUDA tensor (column-major leaves, what sharding.exchange_rows returns).  The batch reads the tensor in place
        and keeps it alive; it is a shard: merkle_tree.local_cap / leaves / digests are the shard's part."""
        ctx = ctx or default_context()
        if isinstance(rows, tuple):  # (device pointer, W, keepalive): a sharding.PeerRows buffer
            ptr, W, rows = rows
        else:
            W, per = rows.shape
            if per * n_shards != 1 << (degree_log + rate_bits) or not rows.is_contiguous():
                raise NlcError("rows must be a contiguous (W, N / n_shards) tensor")
            ptr = rows.data_ptr()
        h = vp()
        ctx.wait_torch()
        ctx._check(_ffi.lib().nlc_commit_from_lde_rows_dev(ctx._h, vp(ptr), W, degree_log, rate_bits,
                                                           cap_height, shard_idx, n_shards, C.byref(h)))
        b = cls(ctx, h, W, degree_log, rate_bits, cap_height)
        b.shard_idx, b.n_shards = shard_idx, n_shards
        b._keepalive = rows
        return b

    @classmethod
    def from_values_row_sharded(cls, values, rate_bits: int, blinding: bool, cap_height: int, group=None, ctx=None,
                                is_values: bool = True, exchange: str = "nccl"):
        """one commit over more GPUs than it has cosets (starky: rate_bits 1 on 4 / 8 GPUs).
        Every rank passes the same `values`; rank p transforms the columns sharding.column_split(W, G)[p]
        (lde_columns), one NCCL all-to-all hands rank g the leaves [g, g+1) * N / G of all columns, rank g builds its
        cap subtrees (from_lde_rows), and the cap entries are all-gathered.  The returned batch is a leaf-row shard:
        merkle_tree.cap is the full MerkleCap, merkle_tree.leaves / digests the shard's part, `column_coeffs` the
        coefficients of this rank's columns `column_range`."""
        import torch
        from . import sharding
        if blinding:
            raise NlcError("blinding = true is not used on this path (prover.rs:91-93,171-178 pass false)")
        g, G = sharding.rank_and_world(group)
        sharding.check_row_plan(cap_height, G)
        ctx = ctx or default_context()
        table, arrs, n = _col_table(values)
        W, log_n = len(arrs), _log2_strict(n)
        c0, nc = sharding.column_split(W, G)[g]
        if nc == 0:
            raise NlcError("fewer columns than ranks: shard independent commits instead")
        dev = torch.device("cuda", ctx.device)
        # kernels (library) and exchange (torch / NCCL) must share ONE real stream: torch's default stream is the
        # NULL stream, which the C ABI reads as "use your own", so make a stream current for the whole commit
        stream = torch.cuda.Stream(dev)
        ctx.set_stream(stream.cuda_stream)
        try:
            with torch.cuda.stream(stream):
                mine = np.stack([np.asarray(arrs[j]).view(np.int64) for j in range(c0, c0 + nc)])
                src = torch.from_numpy(mine).pin_memory().to(dev, non_blocking=True)
                if exchange == "p2p":
                    # exchange fused into the transform: the last NTT pass stores into the peers' buffers over NVLink
                    peers = sharding.PeerRows(ctx, W * ((n << rate_bits) // G), group)
                    coeffs, _ = cls.lde_columns_scatter(src, rate_bits, G, is_values, ctx, peers.ptrs, c0)
                    sharding.stream_barrier(dev, group)
                    stream.synchronize()
                    peers.close_peers()  # before the cap all-gather: nobody maps a buffer whose owner has returned
                    b = cls.from_lde_rows((peers.own, W, peers), log_n, rate_bits, cap_height, g, G, ctx)
                elif exchange == "nccl":
                    coeffs, send = cls.lde_columns_scatter(src, rate_bits, G, is_values, ctx)
                    rows = sharding.exchange_rows(send, W, group)
                    b = cls.from_lde_rows(rows, log_n, rate_bits, cap_height, g, G, ctx)
                    del send
                else:
                    raise NlcError("exchange must be 'nccl' or 'p2p'")
                b.column_range, b.column_coeffs = (c0, nc), coeffs
                stream.synchronize()
                del src
        finally:
            ctx.set_stream(None)
        b.merkle_tree._cap = sharding.gather_cap(b.merkle_tree.local_cap, cap_height, group, device=ctx.device)
        return b

    @classmethod
    def from_device(cls, dev_ptr: int, n_polys: int, degree_log: int, rate_bits: int, cap_height: int,
                    is_values: bool = True, ctx=None):
        """Input already in HBM as an n_polys x n column-major matrix (asynchronous: no host sync).  The commit is
        ordered after torch's current stream (Context.wait_torch); the caller keeps the buffer alive until a
        synchronising accessor (merkle_tree.cap, ...) or ctx.synchronize() has returned."""
        ctx = ctx or default_context()
        ctx.wait_torch()
        h = vp()
        fn = _ffi.lib().nlc_commit_from_values_dev if is_values else _ffi.lib().nlc_commit_from_coeffs_dev
        ctx._check(fn(ctx._h, vp(dev_ptr), n_polys, degree_log, rate_bits, cap_height, C.byref(h)))
        return cls(ctx, h, n_polys, degree_log, rate_bits, cap_height)

    @classmethod
    def import_serialized(cls, polynomials, rate_bits: int, cap_height: int, cap=None, digests=None, ctx=None):
        """A PolynomialBatch deserialised from circuit data (constants_sigmas_commitment of
        /root/reference/scripts/build-circuit.sh:10-18): the LDE and the tree are recomputed on the GPU from
        `polynomials`; `cap` / `digests` (the serialised merkle_tree fields), when given, must match or this raises."""
        ctx = ctx or default_context()
        table, arrs, n = _col_table(polynomials)
        log_n = _log2_strict(n)
        cap_a = None if cap is None else _a(cap).reshape(-1)
        dig_a = None if digests is None else _a(digests).reshape(-1)
        if cap_a is not None and cap_a.size != 4 << cap_height:
            raise NlcError("serialised cap has the wrong size")
        if dig_a is not None and dig_a.size != 8 * ((n << rate_bits) - (1 << cap_height)):
            raise NlcError("serialised digests have the wrong size")
        h = vp()
        ctx._check(_ffi.lib().nlc_batch_import(ctx._h, table, len(arrs), log_n, rate_bits, cap_height,
                                               None if cap_a is None else _p(cap_a),
                                               None if dig_a is None else _p(dig_a), C.byref(h)))
        return cls(ctx, h, len(arrs), log_n, rate_bits, cap_height)

    @classmethod
    def from_trace_rows(cls, rows, col0: int, n_cols: int, rate_bits: int, cap_height: int, ctx=None):
        """Commit columns [col0, col0+n_cols) of a row-major AirTrace (starkyx trace/mod.rs:19-24)."""
        ctx = ctx or default_context()
        r = _a(rows)
        log_n = _log2_strict(r.shape[0])
        h = vp()
        ctx._check(_ffi.lib().nlc_commit_from_trace_rows(ctx._h, _p(r), r.shape[1], col0, n_cols, log_n, rate_bits,
                                                         cap_height, C.byref(h)))
        return cls(ctx, h, n_cols, log_n, rate_bits, cap_height)

    @classmethod
    def from_quotient_values(cls, quotient_values, degree_log: int, quotient_degree_factor: int, rate_bits: int,
                             cap_height: int, divide_by_zh: bool = False, ctx=None):
        """What starkyx does with the constraint evaluations (prover.rs:336-357, 158-178):
        `quotient_values` = size x num_challenges row-major (size = n << log2_ceil(quotient_degree_factor)) ->
        per challenge coset_ifft(F::coset_shift()), trim_to_len(n * quotient_degree_factor), chunks(n),
        from_coeffs(chunks, rate_bits, false, cap_height).  divide_by_zh: rows are raw accumulators, multiply by
        ZeroPolyOnCoset::eval_inverse first.  Panics like the reference when the quotient does not divide."""
        ctx = ctx or default_context()
        q = _a(quotient_values)
        qdb = max(quotient_degree_factor - 1, 0).bit_length()  # log2_ceil
        if q.ndim != 2 or q.shape[0] != 1 << (degree_log + qdb):
            raise NlcError("quotient_values must be (n << log2_ceil(quotient_degree_factor)) x num_challenges")
        h = vp()
        ctx._check(_ffi.lib().nlc_quotient_commit(ctx._h, _p(q), q.shape[1], degree_log, qdb, quotient_degree_factor,
                                                  int(divide_by_zh), rate_bits, cap_height, C.byref(h)))
        return cls(ctx, h, q.shape[1] * quotient_degree_factor, degree_log, rate_bits, cap_height)

    # -- fields ---------------------------------------------------------------------------------
    @property
    def lde_size(self) -> int:
        return 1 << (self.degree_log + self.rate_bits)

    @property
    def polynomials(self):
        if self._polys is None:
            n = 1 << self.degree_log
            out = np.zeros((self.n_polys, n), np.uint64)
            table = (u64p * self.n_polys)(*[_p(out[j]) for j in range(self.n_polys)])
            self._ctx._check(_ffi.lib().nlc_batch_coeffs(self._h, table))
            self._coeff_matrix = out
            self._polys = [PolynomialCoeffs(out[j]) for j in range(self.n_polys)]
        return self._polys

    @property
    def coeff_matrix(self) -> np.ndarray:
        self.polynomials
        return self._coeff_matrix

    def get_lde_values(self, index: int, step: int = 1) -> np.ndarray:
        """leaves[reverse_bits(index * step, lde_bits)] (fri/oracle.rs get_lde_values)."""
        return self.get_lde_rows([index], step)[0]

    def get_lde_rows(self, indices, step: int = 1) -> np.ndarray:
        idx = _a(indices).reshape(-1)
        out = np.zeros((idx.size, self.n_polys), np.uint64)
        self._ctx._check(_ffi.lib().nlc_batch_lde_rows(self._h, _p(idx), idx.size, step, _p(out)))
        return out

    def get_lde_rows_range(self, index_start: int, count: int, step: int = 1) -> np.ndarray:
        """get_lde_values(i, step) for i = index_start .. index_start + count - 1 in one launch and one D2H."""
        out = np.zeros((count, self.n_polys), np.uint64)
        self._ctx._check(_ffi.lib().nlc_batch_lde_rows_range(self._h, index_start, count, step, _p(out)))
        return out

    def get_lde_values_packed(self, index_start: int, step: int, width: int = 1) -> np.ndarray:
        """P::WIDTH consecutive rows, transposed to n_polys x width like PackedField::from_slice per poly."""
        return self.get_lde_rows_range(index_start, width, step).T.copy()

    def eval_at(self, z) -> np.ndarray:
        """Every polynomial at the extension point z -> (n_polys, 2) (StarkOpeningSet::new, proof.rs:307-310)."""
        out = np.zeros((self.n_polys, 2), np.uint64)
        self._ctx._check(_ffi.lib().nlc_batch_eval(self._h, _p(_a(z)), _p(out)))
        return out

    def device_pointers(self):
        L = _ffi.lib()
        return {"lde": L.nlc_batch_lde_dev(self._h), "coeffs": L.nlc_batch_coeffs_dev(self._h),
                "cap": L.nlc_batch_cap_dev(self._h)}

    # -- openings -------------------------------------------------------------------------------
    @staticmethod
    def prove_openings(instance: FriInstanceInfo, oracles, challenger: Challenger, fri_params: FriParams,
                       timing=None) -> FriProof:
        """PolynomialBatch::prove_openings(instance, oracles, challenger, fri_params, timing)."""
        if not oracles:
            raise NlcError("prove_openings: no oracles")
        ctx = oracles[0]._ctx
        handles = (vp * len(oracles))(*[o._h for o in oracles])
        p = fri_params.to_c()
        ci = instance.to_c()
        words = _ffi.lib().nlc_fri_proof_words(C.byref(p), handles, len(oracles))
        out = np.zeros(words, np.uint64)
        ctx._check(_ffi.lib().nlc_prove_openings(ctx._h, C.byref(ci), handles, len(oracles), C.byref(challenger.c),
                                                 C.byref(p), _p(out)))
        return FriProof.parse(out, fri_params, [o.n_polys for o in oracles])

    def free(self):
        self._s.free()


class FriCommitPhase:
    """Rust-driven commit phase (transcript stays with the caller): nlc_fri_begin / commit_round /
    fold_round / final_poly / queries -- the steps of fri_proof() in fri/prover.rs."""

    def __init__(self, instance: FriInstanceInfo, oracles, alpha, fri_params: FriParams):
        self._ctx = oracles[0]._ctx
        self._oracles = oracles
        self.params = fri_params
        handles = (vp * len(oracles))(*[o._h for o in oracles])
        p = fri_params.to_c()
        ci = instance.to_c()
        self._h = vp()
        self._ctx._check(_ffi.lib().nlc_fri_begin(self._ctx._h, C.byref(ci), handles, len(oracles), _p(_a(alpha)),
                                                  C.byref(p), C.byref(self._h)))
        p2 = fri_params.to_c()
        total = _ffi.lib().nlc_fri_proof_words(C.byref(p2), handles, len(oracles))
        nq = fri_params.config.num_query_rounds
        fixed = len(fri_params.reduction_arity_bits) * (4 << fri_params.config.cap_height) + 2 * fri_params.final_poly_len() + 1
        self.query_words = (total - fixed) // nq if nq else 0

    def commit_round(self) -> np.ndarray:
        cap = np.zeros((1 << self.params.config.cap_height, 4), np.uint64)
        self._ctx._check(_ffi.lib().nlc_fri_commit_round(self._h, _p(cap)))
        return cap

    def fold_round(self, beta):
        self._ctx._check(_ffi.lib().nlc_fri_fold_round(self._h, _p(_a(beta))))

    def final_poly(self) -> np.ndarray:
        out = np.zeros((self.params.final_poly_len(), 2), np.uint64)
        self._ctx._check(_ffi.lib().nlc_fri_final_poly(self._h, _p(out)))
        return out

    def queries(self, x_indices) -> np.ndarray:
        x = _a(x_indices).reshape(-1)
        out = np.zeros((x.size, self.query_words), np.uint64)
        self._ctx._check(_ffi.lib().nlc_fri_queries(self._h, _p(x), x.size, _p(out)))
        return out

    def __del__(self):
        try:
            if self._h and self._ctx._h:
                _ffi.lib().nlc_fri_state_free(self._h)
            self._h = None
        except Exception:
            pass


def fri_proof_of_work(challenger: Challenger, proof_of_work_bits: int) -> int:
    """fri_proof_of_work's search only (smallest valid witness); the caller observes it afterwards."""
    out = np.zeros(1, np.uint64)
    ctx = challenger._ctx
    ctx._check(_ffi.lib().nlc_fri_pow(ctx._h, C.byref(challenger.c), proof_of_work_bits, _p(out)))
    return int(out[0])

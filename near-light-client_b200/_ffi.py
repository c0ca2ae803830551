"""ctypes binding of libnlc_commit.so (include/nlc_commit.h) -- the same C ABI the Rust FFI crate binds
(INTEGRATION.md).  No compute happens in Python and there is no fallback: if the shared library is
missing or no CUDA device is usable, loading / ``Context()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NLC_COMMIT_LIB") or os.path.join(_HERE, "libnlc_commit.so")

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
vp = C.c_void_p


class NlcError(RuntimeError):
    """A non-zero return code of the C ABI (plonky2's contract at this boundary is panic)."""


class CChallenger(C.Structure):
    _fields_ = [("sponge_state", C.c_uint64 * 12), ("input_buffer", C.c_uint64 * 8), ("input_len", C.c_uint32),
                ("output_buffer", C.c_uint64 * 8), ("output_len", C.c_uint32)]


class CTiming(C.Structure):
    _fields_ = [("ifft_ms", C.c_double), ("lde_ms", C.c_double), ("leaf_hash_ms", C.c_double),
                ("merkle_levels_ms", C.c_double), ("commits", C.c_uint64)]


class CFriParams(C.Structure):
    _fields_ = [("degree_bits", C.c_int), ("rate_bits", C.c_int), ("cap_height", C.c_int),
                ("proof_of_work_bits", C.c_int), ("num_query_rounds", C.c_int), ("n_reduction_rounds", C.c_int),
                ("reduction_arity_bits", C.c_int * 32)]  # NLC_MAX_FRI_ROUNDS


class CAirProgram(C.Structure):
    _fields_ = [("code", u32p), ("n_ops", C.c_uint32), ("n_regs", C.c_uint32), ("pool", u64p), ("n_pool", C.c_uint32),
                ("n_public", C.c_uint32), ("n_global", C.c_uint32), ("n_challenge", C.c_uint32)]


class CPlonkCircuit(C.Structure):
    _fields_ = [("num_constants", C.c_size_t), ("num_routed_wires", C.c_size_t), ("num_challenges", C.c_size_t),
                ("num_partial_products", C.c_size_t), ("quotient_degree_factor", C.c_int), ("k_is", u64p)]


class CFriInstance(C.Structure):
    _fields_ = [("n_batches", C.c_int), ("points", u64p), ("batch_offsets", u32p), ("poly_oracle", u32p),
                ("poly_index", u32p)]


# name -> (restype, argtypes); every symbol include/nlc_commit.h declares
SIGNATURES = {
    "nlc_init": (C.c_int, [C.c_int, C.POINTER(vp)]),
    "nlc_destroy": (None, [vp]),
    "nlc_last_error": (C.c_char_p, [vp]),
    "nlc_device": (C.c_int, [vp]),
    "nlc_set_stream": (C.c_int, [vp, vp]),
    "nlc_synchronize": (C.c_int, [vp]),
    "nlc_wait_stream": (C.c_int, [vp, vp]),
    "nlc_cache_stats": (C.c_int, [vp, u64p]),
    "nlc_launch_count": (C.c_uint64, [vp]),
    "nlc_timing_enable": (C.c_int, [vp, C.c_int]),
    "nlc_timing_read": (C.c_int, [vp, C.POINTER(CTiming)]),
    "nlc_commit_from_values": (C.c_int, [vp, C.POINTER(u64p), C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
    "nlc_commit_from_coeffs": (C.c_int, [vp, C.POINTER(u64p), C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
    "nlc_commit_from_values_materialised": (C.c_int, [vp, C.POINTER(u64p), C.c_size_t, C.c_int, C.c_int, C.c_int,
                                                      C.POINTER(u64p), u64p, u64p, u64p, C.POINTER(vp)]),
    "nlc_commit_from_values_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
    "nlc_commit_from_values_shard": (C.c_int, [vp, C.POINTER(u64p), C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                               C.c_size_t, C.POINTER(vp)]),
    "nlc_commit_from_values_shard_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                                   C.c_size_t, C.POINTER(vp)]),
    "nlc_ifft_columns_bcast_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_size_t, C.c_size_t, C.POINTER(vp)]),
    "nlc_commit_from_coeffs_shard_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t,
                                                   C.POINTER(vp)]),
    "nlc_lde_columns_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, C.c_int, vp, vp]),
    "nlc_commit_from_lde_rows_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t,
                                               C.POINTER(vp)]),
    "nlc_lde_columns_scatter_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, C.c_int, vp, vp, C.c_size_t,
                                              C.POINTER(vp), C.c_size_t]),
    "nlc_peer_alloc": (C.c_int, [vp, C.c_size_t, C.POINTER(vp), C.c_char_p]),
    "nlc_peer_open": (C.c_int, [vp, C.c_char_p, C.POINTER(vp)]),
    "nlc_peer_close": (C.c_int, [vp, vp]),
    "nlc_peer_free": (C.c_int, [vp, vp]),
    "nlc_batch_shard": (C.c_int, [vp, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "nlc_commit_from_coeffs_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
    "nlc_commit_from_trace_rows": (C.c_int, [vp, u64p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int,
                                             C.POINTER(vp)]),
    "nlc_quotient_domain": (C.c_int, [vp, C.c_int, C.c_int, u64p, u64p, u64p, u64p]),
    "nlc_quotient_commit": (C.c_int, [vp, u64p, C.c_size_t, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_int,
                                      C.POINTER(vp)]),
    "nlc_quotient_commit_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                          C.c_int, C.POINTER(vp)]),
    "nlc_air_eval_dev": (C.c_int, [vp, C.POINTER(CAirProgram), C.POINTER(vp), C.c_int, u64p, u64p, u64p, u64p, C.c_size_t,
                                   C.c_int, vp]),
    "nlc_air_eval": (C.c_int, [vp, C.POINTER(CAirProgram), C.POINTER(vp), C.c_int, u64p, u64p, u64p, u64p, C.c_size_t,
                               C.c_int, u64p]),
    "nlc_stark_quotient_commit": (C.c_int, [vp, C.POINTER(CAirProgram), C.POINTER(vp), C.c_int, u64p, u64p, u64p, u64p,
                                            C.c_size_t, C.c_size_t, C.c_int, C.POINTER(vp)]),
    "nlc_plonk_zs_partial_products_commit": (C.c_int, [vp, C.POINTER(CPlonkCircuit), vp, vp, u64p, u64p, C.c_int, C.POINTER(vp)]),
    "nlc_plonk_vanishing_eval": (C.c_int, [vp, C.POINTER(CPlonkCircuit), C.POINTER(CAirProgram), vp, vp, vp, u64p, u64p, u64p,
                                           u64p, u64p]),
    "nlc_plonk_quotient_commit": (C.c_int, [vp, C.POINTER(CPlonkCircuit), C.POINTER(CAirProgram), vp, vp, vp, u64p, u64p, u64p,
                                            u64p, C.c_int, C.POINTER(vp)]),
    "nlc_batch_free": (None, [vp]),
    "nlc_batch_info": (C.c_int, [vp, C.POINTER(C.c_size_t), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "nlc_batch_cap": (C.c_int, [vp, u64p]),
    "nlc_batch_coeffs": (C.c_int, [vp, C.POINTER(u64p)]),
    "nlc_batch_leaves": (C.c_int, [vp, u64p]),
    "nlc_batch_digests": (C.c_int, [vp, u64p]),
    "nlc_batch_lde_rows": (C.c_int, [vp, u64p, C.c_size_t, C.c_size_t, u64p]),
    "nlc_batch_lde_rows_range": (C.c_int, [vp, C.c_size_t, C.c_size_t, C.c_size_t, u64p]),
    "nlc_batch_import": (C.c_int, [vp, C.POINTER(u64p), C.c_size_t, C.c_int, C.c_int, C.c_int, u64p, u64p, C.POINTER(vp)]),
    "nlc_batch_open": (C.c_int, [vp, C.c_size_t, u64p, u64p]),
    "nlc_batch_lde_dev": (vp, [vp]),
    "nlc_batch_coeffs_dev": (vp, [vp]),
    "nlc_batch_cap_dev": (vp, [vp]),
    "nlc_batch_eval": (C.c_int, [vp, u64p, u64p]),
    "nlc_fft": (C.c_int, [vp, u64p, C.c_size_t, C.c_int]),
    "nlc_ifft": (C.c_int, [vp, u64p, C.c_size_t, C.c_int]),
    "nlc_coset_fft": (C.c_int, [vp, u64p, C.c_size_t, C.c_int, C.c_uint64]),
    "nlc_coset_ifft": (C.c_int, [vp, u64p, C.c_size_t, C.c_int, C.c_uint64]),
    "nlc_lde_onto_coset": (C.c_int, [vp, u64p, u64p, C.c_size_t, C.c_int, C.c_int]),
    "nlc_poseidon_permute": (C.c_int, [vp, u64p, C.c_size_t]),
    "nlc_hash_no_pad": (C.c_int, [vp, u64p, C.c_size_t, C.c_size_t, u64p]),
    "nlc_hash_or_noop": (C.c_int, [vp, u64p, C.c_size_t, C.c_size_t, u64p]),
    "nlc_two_to_one": (C.c_int, [vp, u64p, u64p, C.c_size_t, u64p]),
    "nlc_merkle_cap_static_dev": (C.c_int, [vp, vp, C.c_size_t, C.c_size_t, C.c_int, vp]),
    "nlc_merkle_new": (C.c_int, [vp, u64p, C.c_size_t, C.c_size_t, C.c_int, C.POINTER(vp)]),
    "nlc_tree_cap": (C.c_int, [vp, u64p]),
    "nlc_tree_digests": (C.c_int, [vp, u64p]),
    "nlc_tree_open": (C.c_int, [vp, C.c_size_t, u64p, u64p]),
    "nlc_tree_free": (None, [vp]),
    "nlc_challenger_init": (None, [C.POINTER(CChallenger)]),
    "nlc_challenger_observe": (C.c_int, [vp, C.POINTER(CChallenger), u64p, C.c_size_t]),
    "nlc_challenger_get": (C.c_int, [vp, C.POINTER(CChallenger), C.c_size_t, u64p]),
    "nlc_fri_params_constant_arity": (C.c_int, [C.POINTER(CFriParams)] + [C.c_int] * 7),
    "nlc_fri_proof_words": (C.c_size_t, [C.POINTER(CFriParams), C.POINTER(vp), C.c_int]),
    "nlc_prove_openings": (C.c_int, [vp, C.POINTER(CFriInstance), C.POINTER(vp), C.c_int, C.POINTER(CChallenger),
                                     C.POINTER(CFriParams), u64p]),
    "nlc_fri_begin": (C.c_int, [vp, C.POINTER(CFriInstance), C.POINTER(vp), C.c_int, u64p, C.POINTER(CFriParams),
                                C.POINTER(vp)]),
    "nlc_fri_commit_round": (C.c_int, [vp, u64p]),
    "nlc_fri_fold_round": (C.c_int, [vp, u64p]),
    "nlc_fri_final_poly": (C.c_int, [vp, u64p]),
    "nlc_fri_pow": (C.c_int, [vp, C.POINTER(CChallenger), C.c_int, u64p]),
    "nlc_fri_queries": (C.c_int, [vp, u64p, C.c_size_t, u64p]),
    "nlc_fri_state_free": (None, [vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libnlc_commit.so and type every entry point.  Raises if the library is absent (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NlcError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib

"""CPU suite: the C-ABI library loads, exports every symbol include/nlc_commit.h declares, and fails loudly
(no fallback) without a CUDA device.  No compute calls here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "nlc_commit.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nlc_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg._ffi.lib()
    names = declared_symbols()
    assert len(names) >= 50
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/nlc_commit.h but not exported"
    assert sorted(pkg._ffi.SIGNATURES) == names  # the ctypes table binds exactly the header


def test_rust_sys_crate_is_complete_and_current():
    """bindings/nlc-commit-sys/src/lib.rs is GENERATED from the header (tools/gen_sys_rs.py): every declared symbol is
    bound, with the header's arity, and the committed file is what the generator produces now (no rustc here, so this
    is the check that the Rust surface cannot drift or be hand-elided)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("gen_sys_rs", os.path.join(ROOT, "tools", "gen_sys_rs.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    header = open(os.path.join(ROOT, "include", "nlc_commit.h")).read()
    defines, structs, funcs = gen.parse_header(header)
    rs = open(os.path.join(ROOT, "bindings", "nlc-commit-sys", "src", "lib.rs")).read()
    assert rs == gen.render(defines, structs, funcs), "stale: run python tools/gen_sys_rs.py"
    assert sorted(f[0] for f in funcs) == declared_symbols()
    plain = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    for name, args, _ret in funcs:
        m = re.search(r"\b%s\s*\(([^()]*)\)" % name, plain)
        c_arity = len([a for a in m.group(1).split(",") if a.strip() and a.strip() != "void"])
        r = re.search(r"pub fn %s\(([^)]*)\)" % name, rs)
        assert r, name
        assert len([a for a in r.group(1).split(",") if a.strip()]) == c_arity == len(args), name
    for sname, fields in structs:
        assert f"pub struct {sname} {{" in rs and all(f"pub {f}:" in rs for f, _ in fields)


def test_struct_layouts_match_header(pkg):
    assert ctypes.sizeof(pkg._ffi.CChallenger) == 12 * 8 + 8 * 8 + 8 + 8 * 8 + 8
    assert ctypes.sizeof(pkg._ffi.CFriParams) == 4 * (6 + 32)  # NLC_MAX_FRI_ROUNDS = 32
    assert pkg._ffi.CChallenger.output_buffer.offset == 12 * 8 + 8 * 8 + 8


def test_host_only_entry_points(pkg):
    """fri_params / proof_words / challenger_init are pure host logic: callable without a GPU."""
    P2 = pkg.plonky2
    p = P2.FriConfig.standard_fast_config().fri_params(16)
    assert p.reduction_arity_bits == [4, 4, 4] and p.final_poly_len() == 16 and p.lde_bits() == 17
    assert P2.FriConfig.standard_recursion_config().fri_params(12).reduction_arity_bits == [4, 4]
    assert P2.FriConfig.standard_fast_config().fri_params(5).reduction_arity_bits == []
    # plonky2's ConstantArityBits loop is unbounded; never hand back fewer rounds than it computes
    p = P2.FriConfig(1, 4, 16, P2.FriReductionStrategy.ConstantArityBits(1, 5), 84).fri_params(24)
    assert p.reduction_arity_bits == [1] * 19 and p.final_poly_len() == 32
    raw = pkg._ffi.CFriParams()
    assert pkg._ffi.lib().nlc_fri_params_constant_arity(ctypes.byref(raw), 40, 1, 0, 16, 84, 1, 0) != 0  # > 32 rounds
    c = pkg._ffi.CChallenger()
    c.input_len = 5
    pkg._ffi.lib().nlc_challenger_init(ctypes.byref(c))
    assert c.input_len == 0 and c.output_len == 0


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.NlcError, match="no CPU fallback"):
        pkg.plonky2.Context()


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "near-light-client_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle/" not in src.replace("fri/oracle.rs", "") and "liboracle" not in src, f
                assert "import oracle" not in src and "from oracle" not in src, f

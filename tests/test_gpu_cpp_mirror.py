"""The compiled-language host mirror (include/nlc_plonky2.hpp: plonky2's interface in C++ over the C ABI): the
driver tests/cpp/test_plonky2_mirror.cpp follows starkyx's prover.rs:50-225 and checks every artefact against the
oracle.  CPU: it must compile and link against every symbol it uses; GPU: it must run green."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG, ORC = os.path.join(ROOT, "near-light-client_b200"), os.path.join(ROOT, "oracle")


def build(tmp_path) -> str:
    import __graft_entry__ as graft
    if not (os.path.exists(os.path.join(PKG, "libnlc_commit.so")) and os.path.exists(os.path.join(ORC, "liboracle.so"))):
        graft.build()
    exe = str(tmp_path / "test_plonky2_mirror")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", ORC,
           os.path.join(ROOT, "tests", "cpp", "test_plonky2_mirror.cpp"), "-o", exe, "-L", PKG, "-lnlc_commit", "-L", ORC,
           "-loracle", f"-Wl,-rpath,{PKG}", f"-Wl,-rpath,{ORC}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return exe


def test_cpp_mirror_compiles_and_links(tmp_path):
    build(tmp_path)


@pytest.mark.gpu
def test_cpp_mirror_runs(tmp_path):
    r = subprocess.run([build(tmp_path)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "all checks passed" in r.stdout

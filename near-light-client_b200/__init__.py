"""nlc-commit: B200-native commitment backend (Goldilocks coset LDE -> Poseidon Merkle commit -> FRI)
for the plonky2 / starkyx provers of near/near-light-client.  The product is the CUDA library
libnlc_commit.so (csrc/, C ABI in include/nlc_commit.h); this package is the thin host-side mirror of the
plonky2 interface used by the parity tests and the bench.  The directory name carries a hyphen, so load it
with ``__graft_entry__.load_package()`` (alias ``nlc_b200``).
"""
from . import _ffi  # noqa: F401
from ._ffi import NlcError  # noqa: F401
from . import plonky2  # noqa: F401
from . import sharding  # noqa: F401
from . import starky  # noqa: F401
from . import plonk  # noqa: F401

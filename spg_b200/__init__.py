"""spg_b200 — Python host-side mirror over the spgx C ABI (include/spgx.h).

The product is the CUDA library spg_b200/lib/libspgx.so (hand-written sm_100a kernels behind a C ABI) plus
the C++ facade include/spgx/spgx.hpp.  This package is only the thin ctypes binding the tests, bench.py and
__graft_entry__ drive it through; it contains no solver arithmetic and no CPU fallback: if the shared
library is missing, importing `spg_b200.api` raises.
"""
from . import scenes  # noqa: F401

__all__ = ["scenes"]

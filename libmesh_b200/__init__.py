"""libmesh_b200 — B200-native finite-element assembly hot path for libMesh (CUDA sm_100a behind a C ABI).

Only the assembly path lives here: ``csrc/`` (CUDA kernels + the C ABI of include/feasm.h), ``shim/`` (the C++
SparseMatrix / NumericVector adaptors a libMesh build compiles), and thin Python host helpers used by the tests
and bench.py.  Importing this package never touches ``oracle/``.
"""
from . import capi  # noqa: F401

__all__ = ["capi"]

"""proton_b200 — B200-native HHO / cutHHO hot path (local operators -> static condensation -> CSR -> PCG).

This package is a thin ctypes binding of the C ABI in include/proton_b200.h (libproton_b200.so, built in-tree by
proton_b200/csrc/Makefile).  There is no CPU implementation behind it: importing works without a GPU (so the
symbol table can be checked), every compute call needs a CUDA device and raises otherwise.
"""
from ._binding import (PB, Context, ProtonError, abi_symbols, cg_params, hho_degree_info, lib, lib_path,
                       mesh_init_params, strip_layout, strip_rows, dist_unique_id)

__all__ = ["PB", "Context", "ProtonError", "abi_symbols", "cg_params", "hho_degree_info", "lib", "lib_path", "mesh_init_params", "strip_layout", "strip_rows", "dist_unique_id"]

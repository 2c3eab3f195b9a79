"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/proton_b200.h
declares, and refuses to compute without a CUDA device (no CPU fallback)."""
import ctypes

import pytest

import proton_b200 as pb


def test_library_exports_every_declared_symbol():
    L = pb.lib()
    syms = pb.abi_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/proton_b200.h but not exported"
    assert L.pb_abi_version() == 1


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pb.ProtonError, match="no CUDA device|no CPU fallback"):
        pb.Context(0)


def test_product_does_not_import_oracle():
    import os, re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "proton_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"(from|import)\s+oracle|oracle/|orc_", txt), f"{f} references the oracle"


def test_degree_info_reverts_like_reference():
    # utils.hpp:75-95
    assert (pb.hho_degree_info(3, 1).cd, pb.hho_degree_info(3, 1).fd) == (1, 1)
    assert (pb.hho_degree_info(1, 0).cd, pb.hho_degree_info(0, 1).cd) == (1, 0)
    assert pb.hho_degree_info(2, 1).rbs == 6

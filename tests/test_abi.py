"""CPU checks: the C-ABI library loads and exports every symbol the header declares."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "gmcore_b200/lib/libgmcore_b200.so")


def declared_symbols():
    txt = open(os.path.join(ROOT, "include/gmcore_gpu.h")).read()
    return sorted(set(re.findall(r"\b(gmc_[a-z_]+)\s*\(", txt)))


@pytest.mark.skipif(not os.path.exists(LIB), reason="CUDA library not built")
def test_library_exports_header_symbols():
    lib = C.CDLL(LIB)
    syms = declared_symbols()
    assert len(syms) >= 17
    for s in syms:
        assert hasattr(lib, s), s


@pytest.mark.skipif(not os.path.exists(LIB), reason="CUDA library not built")
def test_struct_sizes_match_ctypes():
    from gmcore_b200 import abi
    lib = C.CDLL(LIB)
    a, b = C.c_int(), C.c_int()
    lib.gmc_struct_sizes(C.byref(a), C.byref(b))
    assert a.value == C.sizeof(abi.GmcConfig)
    assert b.value == C.sizeof(abi.GmcMeshVectors)


@pytest.mark.skipif(not os.path.exists(LIB), reason="CUDA library not built")
def test_no_cpu_fallback_without_device():
    """Without a GPU gmc_create must fail loudly (no CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from gmcore_b200.host import Host
    from gmcore_b200.model import GmcoreGPU
    host = Host(32, 17, 1, 600.0, "rossby_haurwitz_wave")
    with pytest.raises(RuntimeError, match="no CUDA device|CPU fallback"):
        GmcoreGPU(host)


def test_product_never_imports_oracle():
    for dp, _, fs in os.walk(os.path.join(ROOT, "gmcore_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp")):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle" not in txt.replace("-ffp-contract=off oracle", ""), os.path.join(dp, f)

"""CPU-only checks of the drop-in boundary: libefgp_b200.so loads, exports every symbol that
include/efgp_b200.h declares, and the product refuses to run without CUDA tensors (no CPU fallback)."""
import os
import re

import pytest
import torch

from conftest import ROOT


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "efgp_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(efgp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from efgp_b200 import _lib
    names = _declared_symbols()
    assert len(names) >= 13
    lib = _lib.lib()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/efgp_b200.h but not exported"
        assert n in _lib.EXPORTED_SYMBOLS, f"{n} has no ctypes signature"
    assert lib.efgp_version() >= 1


def test_no_cpu_fallback():
    from efgp_b200 import ops, _lib
    x = torch.rand(10, 2, dtype=torch.float64)
    with pytest.raises(_lib.EfgpError):
        ops.NufftPlan(x, 0.3)
    import efgpnd
    with pytest.raises(_lib.EfgpError):
        efgpnd.NUFFT(x, torch.zeros(2, dtype=torch.float64), 0.3, 1e-6)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "gp-quadrature_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                s = open(os.path.join(dp, f)).read()
                assert "efgp_oracle" not in s and "fast_nufft" not in s and "cnufft" not in s, os.path.join(dp, f)

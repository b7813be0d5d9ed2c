"""The C-ABI library loads and exports every symbol include/kgcuda.h declares (no compute calls: no GPU here)."""
import os
import re

from kineticgas_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_match_header():
    hdr = open(os.path.join(ROOT, "include", "kgcuda.h")).read()
    declared = set(re.findall(r"\b(kg_[a-z0-9_]+)\s*\(", hdr))
    L = _lib.lib()
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/kgcuda.h but not exported"
    assert declared == set(_lib.EXPORTS)


def test_no_cpu_fallback():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from kineticgas_b200 import DeviceModel, mie_desc
    with pytest.raises(RuntimeError, match="no CUDA device"):
        DeviceModel([mie_desc(3e-10, 1e-21, 6, 12)], [1e-26], True)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "kineticgas_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "oracle" not in src.replace("# oracle-free", ""), f"{f} mentions the oracle"

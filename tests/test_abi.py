"""CPU tests of the boundary: the C-ABI library loads, exports every symbol the header declares,
and fails loudly (no CPU fallback) when there is no CUDA device."""
import os
import re
import pytest
import psascan_b200 as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    h = open(os.path.join(ROOT, "include", "psascan_cuda.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(psc_[a-z0-9_]+)\s*\(", h)))


def test_library_exports_every_header_symbol():
    lib = P.load_library()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), s
    assert sorted(P.EXPORTS) == syms


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(P.PscError) as e:
        P.Context()
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_product_never_touches_oracle():
    for d, _, files in os.walk(os.path.join(ROOT, "psascan_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                src = open(os.path.join(d, f)).read()
                assert "oracle" not in src.lower() or f == "__init__.py" and False, os.path.join(d, f)

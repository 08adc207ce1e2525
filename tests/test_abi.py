"""CPU-side checks of the C-ABI boundary: the library loads and exports every declared symbol."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    h = open(os.path.join(ROOT, "include", "cuclark_b200.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(ccb_[a-z0-9_]+)\s*\(", h)))


def test_header_symbols_exported_and_bound():
    from jn_cuclark_b200 import lib
    L = lib.load_library()
    names = _declared()
    assert len(names) >= 19
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/cuclark_b200.h but not exported"
        assert n in lib.SYMBOLS, f"{n} has no ctypes prototype"
    assert L.ccb_version().decode().endswith("sm_100a")


def test_no_gpu_fails_loudly():
    """Without a CUDA device the product refuses to run (no CPU fallback)."""
    import torch
    import pytest
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from jn_cuclark_b200 import CuClarkDB, CcbError
    with pytest.raises(CcbError) as e:
        CuClarkDB(1, 31, 1, 4)
    assert "No CUDA devices found" in str(e.value)


def test_product_does_not_reference_oracle():
    bad = []
    for d, _, fs in os.walk(os.path.join(ROOT, "jn_cuclark_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", "Makefile")):
                s = open(os.path.join(d, f), errors="ignore").read()
                if re.search(r"oracle|ccbo_", s):
                    bad.append(os.path.join(d, f))
    assert not bad, bad

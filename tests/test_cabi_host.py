"""Host-side checks that need no GPU: the C-ABI library loads and exports every symbol include/magritte_b200.h
declares, and the product refuses to run without a CUDA device (no CPU fallback)."""
import os
import re

import pytest

import util


def _declared():
    src = open(os.path.join(util.ROOT, "include", "magritte_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mgb_[A-Za-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from magritte_b200 import cabi
    lib = cabi.load()
    names = _declared()
    assert len(names) >= 24
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/magritte_b200.h but not exported"
    assert sorted(cabi.EXPORTS) == names
    assert lib.mgb_abi_version() == 2


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from magritte_b200 import cabi
    with pytest.raises(cabi.MgbError, match="no CUDA device"):
        cabi.Context(0)


def test_product_does_not_touch_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py may touch oracle/."""
    pkg = os.path.join(util.ROOT, "magritte_b200")
    for root, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                txt = open(os.path.join(root, fn)).read()
                assert "feautrier_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, fn

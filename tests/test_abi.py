"""CPU tests: the C-ABI library loads, exports every symbol include/krd_b200.h declares, and
refuses to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "krd_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(krd_[A-Za-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    import gappa_b200
    L = gappa_b200.lib()
    syms = _declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/krd_b200.h but not exported"
    assert sorted(gappa_b200.EXPORTS) == syms


def test_no_cpu_fallback_without_gpu():
    import torch
    import gappa_b200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(gappa_b200.KrdError) as ei:
        gappa_b200.KrdPlan(device=0)
    assert ei.value.status == gappa_b200.krd.KRD_ERR_NO_DEVICE
    assert "no CPU fallback" in str(ei.value)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "gappa_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "libkrd_oracle" not in src and "libkrd_ref" not in src, f

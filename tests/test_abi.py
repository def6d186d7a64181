"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/proton_em.h declares, and
fails loudly (no CPU fallback) when no GPU is present."""
import ctypes
import re
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parents[1]


def _built():
    from proton_b200 import build
    return build.build_library()


def test_library_exports_every_declared_symbol():
    lib_path = _built()
    lib = ctypes.CDLL(str(lib_path))
    header = (REPO / "include" / "proton_em.h").read_text()
    declared = sorted(set(re.findall(r"\b(proton_em_[a-z_0-9]+)\s*\(", header)))
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/proton_em.h but not exported"
    from proton_b200 import _lib
    assert sorted(_lib.EXPORTS) == declared


def test_no_silent_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    _built()
    from proton_b200.api import Context, ProtonEmError
    with pytest.raises(ProtonEmError) as ei:
        Context(0)
    assert ei.value.code == -3 and "no CPU path" in ei.value.message


def test_product_never_imports_the_oracle():
    for p in (REPO / "proton_b200").rglob("*"):
        if p.suffix in (".py", ".cu", ".cuh", ".cpp", ".h"):
            txt = p.read_text()
            assert "oracle" not in txt.replace("no oracle", ""), f"{p} mentions the oracle"

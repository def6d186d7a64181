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


def test_host_only_entry_points_and_argument_checks_need_no_gpu(tmp_path):
    """The parts of the ABI that are host logic answer without a device: the class of a line length, the discretiser,
    plan accessors on NULL, and the file-level entry points' argument checks (no context -> PROTON_EM_ECUDA with the
    message in `msg`, nothing read or written)."""
    lib = ctypes.CDLL(str(_built()))
    k, lr, rows = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
    assert lib.proton_em_debug_class_of(19087, ctypes.byref(k), ctypes.byref(lr), ctypes.byref(rows)) == 0
    assert (k.value, lr.value, rows.value) == (5, 12, 10)                  # long Bluestein path, 10 rows of 4096 points
    assert lib.proton_em_debug_class_of(0, ctypes.byref(k), ctypes.byref(lr), ctypes.byref(rows)) == -1
    lib.proton_em_plan_line_name.restype = ctypes.c_char_p
    assert lib.proton_em_plan_line_name(None, ctypes.c_int64(0)) is None
    lib.proton_em_plan_num_lines.restype = ctypes.c_int64
    assert lib.proton_em_plan_num_lines(None) == 0 and lib.proton_em_plan_line_nodes(None, ctypes.c_int64(0)) == 0
    assert lib.proton_em_formulate_segments(None, None, ctypes.c_double(1.0), ctypes.c_double(1.0), ctypes.c_double(1.0),
                                            None, None, None, None, None, None, None, None) == -1
    msg = ctypes.create_string_buffer(256)
    plan, skipped = ctypes.c_void_p(), ctypes.c_int64(7)
    (tmp_path / "input" / "M1_x" / "CuDD_378.0_0.1").mkdir(parents=True)
    rc = lib.proton_em_pack_from_project(None, str(tmp_path).encode(), None, ctypes.byref(plan), ctypes.byref(skipped),
                                         msg, ctypes.c_size_t(len(msg)))
    assert rc == -3 and b"No CUDA context" in msg.value and not plan.value and skipped.value == 0
    rc = lib.proton_em_run_cfgs(None, None, ctypes.c_int64(0), None, msg, ctypes.c_size_t(len(msg)))
    assert rc == -3 and b"No CUDA context" in msg.value
    rc = lib.proton_em_run_cfgs(None, None, ctypes.c_int64(3), None, msg, ctypes.c_size_t(len(msg)))
    assert rc == -1 and b"bad arguments" in msg.value

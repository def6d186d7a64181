"""ctypes binding of libproton_em.so (include/proton_em.h).  There is no fallback: if the CUDA library is
missing or no GPU is usable, every call fails loudly."""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libproton_em.so"

OK, EINVAL, ENOMEM, ECUDA, EIO = 0, -1, -2, -3, -4
HOST, DEVICE = 0, 1
ALG_AUTO, ALG_DIRECT = 0, 1

c_i32p = ctypes.POINTER(ctypes.c_int32)
c_i64p = ctypes.POINTER(ctypes.c_int64)
c_dp = ctypes.POINTER(ctypes.c_double)
c_u8p = ctypes.POINTER(ctypes.c_uint8)


class Batch(ctypes.Structure):
    _fields_ = [("num_lines", ctypes.c_int64), ("num_nodes", ctypes.c_void_p), ("a_coeff", ctypes.c_void_p),
                ("trip_off", ctypes.c_void_p), ("trip_row", ctypes.c_void_p), ("trip_col", ctypes.c_void_p),
                ("trip_val", ctypes.c_void_p), ("u_off", ctypes.c_void_p), ("curden", ctypes.c_void_p)]


class Segments(ctypes.Structure):          # proton_em_segments
    _fields_ = [("num_lines", ctypes.c_int64), ("seg_off", ctypes.c_void_p), ("length_um", ctypes.c_void_p),
                ("current", ctypes.c_void_p), ("lcoord", ctypes.c_void_p), ("rcoord", ctypes.c_void_p),
                ("named", ctypes.c_void_p)]


FORM_CORRUPT, FORM_NOT_DENSE, FORM_INDEX = 1, 2, 3


class Opts(ctypes.Structure):
    _fields_ = [("critical_stress", ctypes.c_double), ("algorithm", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class Result(ctypes.Structure):
    _fields_ = [("stress", ctypes.c_void_p), ("max_stress", ctypes.c_void_p), ("argmax", ctypes.c_void_p),
                ("flag", ctypes.c_void_p)]


class ProtonEmError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"proton_em error {code}: {message}")
        self.code = code
        self.message = message


_lib = None

EXPORTS = ["proton_em_create", "proton_em_destroy", "proton_em_last_error", "proton_em_version", "proton_em_pack",
           "proton_em_plan_destroy", "proton_em_plan_total_nodes", "proton_em_plan_num_lines",
           "proton_em_plan_scratch_bytes", "proton_em_eval", "proton_em_synchronize", "proton_em_stream",
           "proton_em_launch_count", "proton_em_analyze_host", "proton_em_run_cfg", "proton_em_debug_exp",
           "proton_em_eval_line_noalloc", "proton_em_debug_expm1", "proton_em_profile", "proton_em_profile_read",
           "proton_em_analyze_history_host", "proton_em_debug_line_table", "proton_em_pack_sigma0", "proton_em_run_cfgs",
           "proton_em_pack_from_project", "proton_em_plan_line_nodes", "proton_em_plan_set_line_names",
           "proton_em_plan_line_name", "proton_em_formulate_segments",
           "proton_em_debug_class_of"]


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    path = Path(os.environ.get("PROTON_EM_LIB", LIB_PATH))      # (override: A/B timing of two builds on one box)
    if not path.exists():
        raise ProtonEmError(ECUDA, f"{path} is not built (run `python -m proton_b200.build`); "
                                   "there is no CPU fallback")
    lib = ctypes.CDLL(str(path))
    vp = ctypes.c_void_p
    lib.proton_em_create.argtypes = [ctypes.POINTER(vp), ctypes.c_int]
    lib.proton_em_destroy.argtypes = [vp]
    lib.proton_em_last_error.restype = ctypes.c_char_p
    lib.proton_em_version.restype = ctypes.c_char_p
    lib.proton_em_pack.argtypes = [vp, ctypes.POINTER(Batch), ctypes.c_int, ctypes.POINTER(vp)]
    if hasattr(lib, "proton_em_pack_sigma0"):
        lib.proton_em_pack_sigma0.argtypes = [vp, ctypes.POINTER(Batch), vp, ctypes.c_int, ctypes.POINTER(vp)]
    lib.proton_em_plan_destroy.argtypes = [vp]
    for f in (lib.proton_em_plan_total_nodes, lib.proton_em_plan_num_lines, lib.proton_em_plan_scratch_bytes,
              lib.proton_em_launch_count):
        f.argtypes = [vp]
        f.restype = ctypes.c_int64
    lib.proton_em_eval.argtypes = [vp, vp, ctypes.c_int32, ctypes.POINTER(Opts), ctypes.POINTER(Result), ctypes.c_int]
    lib.proton_em_synchronize.argtypes = [vp]
    lib.proton_em_stream.argtypes = [vp]
    lib.proton_em_stream.restype = vp
    lib.proton_em_analyze_host.argtypes = [vp, ctypes.POINTER(Batch), vp, ctypes.c_int32, ctypes.POINTER(Opts),
                                           ctypes.POINTER(Result), ctypes.c_int64]
    lib.proton_em_run_cfg.argtypes = [vp, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_size_t]
    if hasattr(lib, "proton_em_run_cfgs"):
        lib.proton_em_run_cfgs.argtypes = [vp, ctypes.POINTER(ctypes.c_char_p), ctypes.c_int64, c_i32p, ctypes.c_char_p,
                                           ctypes.c_size_t]
    if hasattr(lib, "proton_em_pack_from_project"):
        lib.proton_em_pack_from_project.argtypes = [vp, ctypes.c_char_p, ctypes.c_char_p, ctypes.POINTER(vp),
                                                    ctypes.POINTER(ctypes.c_int64), ctypes.c_char_p, ctypes.c_size_t]
        lib.proton_em_plan_line_nodes.argtypes = [vp, ctypes.c_int64]
        lib.proton_em_plan_line_nodes.restype = ctypes.c_int32
        lib.proton_em_plan_set_line_names.argtypes = [vp, ctypes.POINTER(ctypes.c_char_p), ctypes.c_int64]
        lib.proton_em_plan_line_name.argtypes = [vp, ctypes.c_int64]
        lib.proton_em_plan_line_name.restype = ctypes.c_char_p
    if hasattr(lib, "proton_em_formulate_segments"):
        lib.proton_em_formulate_segments.argtypes = [ctypes.POINTER(Segments), vp, ctypes.c_double, ctypes.c_double,
                                                     ctypes.c_double, vp, vp, vp, vp, vp, vp, vp,
                                                     ctypes.POINTER(ctypes.c_int64)]
    if hasattr(lib, "proton_em_debug_class_of"):
        lib.proton_em_debug_class_of.argtypes = [ctypes.c_int32, c_i32p, c_i32p, c_i32p]
    lib.proton_em_debug_exp.argtypes = [vp, vp, vp, ctypes.c_int64]
    lib.proton_em_debug_expm1.argtypes = [vp, vp, vp, ctypes.c_int64]
    lib.proton_em_eval_line_noalloc.argtypes = [ctypes.c_int, vp, ctypes.c_int32, ctypes.c_double, ctypes.c_int32, vp, vp,
                                                vp, ctypes.c_int32, vp, vp, ctypes.c_int32, ctypes.c_double, vp, vp, vp, vp]
    lib.proton_em_analyze_history_host.argtypes = [vp, ctypes.POINTER(Batch), ctypes.POINTER(vp), vp, ctypes.c_int32, vp,
                                                   ctypes.c_int32, ctypes.POINTER(Opts), ctypes.POINTER(Result)]
    lib.proton_em_profile.argtypes = [vp, ctypes.c_int]
    lib.proton_em_profile_read.argtypes = [vp, ctypes.c_int, ctypes.POINTER(ctypes.c_char_p),
                                           ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64)]
    if hasattr(lib, "proton_em_debug_line_table"):
        lib.proton_em_debug_line_table.argtypes = [vp, ctypes.c_int64, ctypes.c_int, vp, ctypes.c_int64,
                                                   ctypes.POINTER(ctypes.c_int64)]
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise ProtonEmError(rc, load().proton_em_last_error().decode(errors="replace"))


KINDS = {0: "dense", 1: "fft_pow2", 2: "fft_blue", 3: "fft_mix", 4: "long_pow2", 5: "long_blue", -1: "none"}


def class_of(n: int):
    """(family, log2 of the row length, rows) of the kernel that evaluates a line of n nodes — the library's own
    classify(), host logic only."""
    k, lr, rows = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
    check(load().proton_em_debug_class_of(int(n), ctypes.byref(k), ctypes.byref(lr), ctypes.byref(rows)))
    return KINDS[k.value], lr.value, rows.value

"""oracle/restatement.py — TEST INFRASTRUCTURE ONLY.

ctypes front-end of oracle/em_oracle.c (the CPU restatement of bin/EMtool_analytical's arithmetic,
SURVEY.md §8a rows 5-12) plus a driver that runs the UNMODIFIED reference binary under the shim
libraries in oracle/_ref.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs import this module; the product package never does.
"""
from __future__ import annotations

import ctypes
import os
import re
import subprocess
import tempfile
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "liboracle.so"
REF_DIR = HERE / "_ref"
REF_BIN = REF_DIR / "EMtool_analytical"

_lib = None


def build(quiet: bool = True) -> None:
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    out = subprocess.run(["make", "-C", str(HERE)], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if not quiet:
        print(out.stdout)


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        build()
    lib = ctypes.CDLL(str(LIB_PATH))
    c_i32p = ctypes.POINTER(ctypes.c_int32)
    c_i64p = ctypes.POINTER(ctypes.c_int64)
    c_dp = ctypes.POINTER(ctypes.c_double)
    lib.em_oracle_r2r.argtypes = [ctypes.c_int, c_dp, c_dp, ctypes.c_int, ctypes.c_int]
    lib.em_oracle_r2r.restype = ctypes.c_int
    lib.em_oracle_spmv.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_i32p, c_i32p, c_dp, c_dp, c_dp]
    lib.em_oracle_spmv.restype = ctypes.c_int
    lib.em_oracle_line.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_int, c_i32p, c_i32p,
                                   c_dp, c_dp, ctypes.c_int, c_dp, c_dp, c_dp, c_i32p, ctypes.c_int]
    lib.em_oracle_line.restype = ctypes.c_int
    lib.em_oracle_line_r.argtypes = [ctypes.c_int, ctypes.c_double, c_dp, ctypes.c_int, c_dp, c_dp, c_dp,
                                     c_i32p, ctypes.c_int]
    lib.em_oracle_line_r.restype = ctypes.c_int
    lib.em_oracle_batch.argtypes = [ctypes.c_int64, c_i32p, c_dp, c_i64p, c_i32p, c_i32p, c_dp, c_i64p, c_dp,
                                    ctypes.c_int, c_dp, c_dp, c_dp, c_i32p, ctypes.c_int, ctypes.c_int]
    lib.em_oracle_batch.restype = ctypes.c_int
    lib.em_oracle_max_threads.restype = ctypes.c_int
    lib.em_oracle_exp_array.argtypes = [ctypes.c_int64, c_dp, c_dp]
    lib.em_oracle_lambda.argtypes = [ctypes.c_int, c_dp]
    _lib = lib
    return lib


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double)) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)) if a is not None else None


def _lp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)) if a is not None else None


def r2r(x, kind: int, use_double: bool = False) -> np.ndarray:
    """FFTW-style unnormalised REDFT10 (kind=5) / REDFT01 (kind=4)."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty_like(x)
    rc = load().em_oracle_r2r(len(x), _dp(x), _dp(y), kind, int(use_double))
    if rc:
        raise ValueError(f"em_oracle_r2r rc={rc}")
    return y


def spmv(N, S, rows, cols, vals, u) -> np.ndarray:
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    vals = np.ascontiguousarray(vals, dtype=np.float64)
    u = np.ascontiguousarray(u, dtype=np.float64)
    r = np.empty(N, dtype=np.float64)
    rc = load().em_oracle_spmv(N, S, len(rows), _ip(rows), _ip(cols), _dp(vals), _dp(u), _dp(r))
    if rc:
        raise ValueError(f"em_oracle_spmv rc={rc}")
    return r


def line(N, a, rows, cols, vals, u, ts, use_double: bool = False):
    """One line at T timestamps -> (stress[T,N], max[T], argmax[T])."""
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    vals = np.ascontiguousarray(vals, dtype=np.float64)
    u = np.ascontiguousarray(u, dtype=np.float64)
    ts = np.ascontiguousarray(np.atleast_1d(ts), dtype=np.float64)
    T = len(ts)
    out = np.empty((T, N), dtype=np.float64)
    mx = np.empty(T, dtype=np.float64)
    amx = np.empty(T, dtype=np.int32)
    rc = load().em_oracle_line(N, float(a), len(u), len(rows), _ip(rows), _ip(cols), _dp(vals), _dp(u), T, _dp(ts),
                               _dp(out), _dp(mx), _ip(amx), int(use_double))
    if rc:
        raise ValueError(f"em_oracle_line rc={rc}")
    return out, mx, amx


def batch(num_nodes, a_coeff, trip_off, rows, cols, vals, u_off, u, ts, want_stress=True, use_double=False,
          nthreads=0):
    """Ragged batch in the product ABI's layout -> (stress or None, max[L,T], argmax[L,T])."""
    num_nodes = np.ascontiguousarray(num_nodes, dtype=np.int32)
    a_coeff = np.ascontiguousarray(a_coeff, dtype=np.float64)
    trip_off = np.ascontiguousarray(trip_off, dtype=np.int64)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    vals = np.ascontiguousarray(vals, dtype=np.float64)
    u_off = np.ascontiguousarray(u_off, dtype=np.int64)
    u = np.ascontiguousarray(u, dtype=np.float64)
    ts = np.ascontiguousarray(np.atleast_1d(ts), dtype=np.float64)
    L, T = len(num_nodes), len(ts)
    total = int(num_nodes.astype(np.int64).sum())
    stress = np.empty(total * T, dtype=np.float64) if want_stress else None
    mx = np.empty((L, T), dtype=np.float64)
    amx = np.empty((L, T), dtype=np.int32)
    rc = load().em_oracle_batch(L, _ip(num_nodes), _dp(a_coeff), _lp(trip_off), _ip(rows), _ip(cols), _dp(vals),
                                _lp(u_off), _dp(u), T, _dp(ts), _dp(stress), _dp(mx), _ip(amx), int(use_double),
                                int(nthreads))
    if rc:
        raise ValueError(f"em_oracle_batch rc={rc}")
    return stress, mx, amx


def history_batch(num_nodes, a_coeff, trip_off, rows, cols, vals, u_off, curden_segments, seg_start, ts,
                  use_double=False):
    """Piecewise-constant current history (product: proton_em_analyze_history_host; no counterpart in the reference).
    Superposition of the reference closed form over the current increments, each started at its switching time:
    sigma(t) = sum_s sigma_ref[u_s - u_{s-1}](max(t - tau_s, 0)).  Independently pinned by tests/test_history.py
    against a stiff ODE integration of d sigma/dt = a L sigma + B u(t).  -> (stress, max[L,T], argmax[L,T])"""
    ts = np.ascontiguousarray(np.atleast_1d(ts), dtype=np.float64)
    num_nodes = np.ascontiguousarray(num_nodes, dtype=np.int32)
    total = None
    prev = None
    for u, tau in zip(curden_segments, seg_start):
        u = np.ascontiguousarray(u, dtype=np.float64)
        incr = u if prev is None else u - prev
        prev = u
        st, _, _ = batch(num_nodes, a_coeff, trip_off, rows, cols, vals, u_off, incr, np.maximum(ts - tau, 0.0),
                         use_double=use_double)
        total = st if total is None else total + st
    L, T = len(num_nodes), len(ts)
    off = np.concatenate([[0], np.cumsum(num_nodes.astype(np.int64))])
    mx = np.empty((L, T)); amx = np.empty((L, T), np.int32)
    for l in range(L):
        blk = total[off[l] * T:off[l + 1] * T].reshape(T, int(num_nodes[l]))
        mx[l] = blk.max(axis=1); amx[l] = blk.argmax(axis=1)
    return total, mx, amx


def sigma0_batch(num_nodes, a_coeff, trip_off, rows, cols, vals, u_off, curden, sigma0, ts, use_double=False):
    """Evolution from an initial stress field (product: proton_em_pack_sigma0; no counterpart in the reference, whose
    kernel starts from zero stress):  sigma(t) = V exp(Lambda a t) V^T sigma0 + the reference closed form for the
    currents, V = the DCT-II basis, lambda_k as the reference computes it.  The homogeneous part is restated with
    scipy.fft (DCT-II forward, DCT-III back) — another transform code than either the GPU path or em_oracle.c — and
    pinned in tests/test_sigma0.py against a stiff ODE integration of d sigma/dt = a L sigma + B u from sigma0.
    -> (stress, max[L,T], argmax[L,T])"""
    from scipy.fft import dct
    ts = np.ascontiguousarray(np.atleast_1d(ts), dtype=np.float64)
    num_nodes = np.ascontiguousarray(num_nodes, dtype=np.int32)
    a_coeff = np.ascontiguousarray(a_coeff, dtype=np.float64)
    sigma0 = np.ascontiguousarray(sigma0, dtype=np.float64).reshape(-1)
    total, _, _ = batch(num_nodes, a_coeff, trip_off, rows, cols, vals, u_off, curden, ts, use_double=use_double)
    total = total.copy()
    L, T = len(num_nodes), len(ts)
    off = np.concatenate([[0], np.cumsum(num_nodes.astype(np.int64))])
    mx = np.empty((L, T)); amx = np.empty((L, T), np.int32)
    for l in range(L):
        N = int(num_nodes[l])
        lam = reference_lambda(N)
        y = dct(sigma0[off[l]:off[l + 1]], type=2) / (2.0 * N)            # y_k = (1/N) sum_n x_n cos(..)
        blk = total[off[l] * T:off[l + 1] * T].reshape(T, N)
        for j, t in enumerate(ts):
            blk[j] += dct(y * np.exp(lam * (a_coeff[l] * t)), type=3)    # x_0 + 2 sum_{k>=1} x_k cos(..)
        mx[l] = blk.max(axis=1); amx[l] = blk.argmax(axis=1)
    return total, mx, amx


def max_threads() -> int:
    return int(load().em_oracle_max_threads())


def libm_exp(x) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty_like(x)
    load().em_oracle_exp_array(x.size, _dp(x), _dp(y))
    return y


def reference_lambda(N: int) -> np.ndarray:
    lam = np.empty(N, dtype=np.float64)
    load().em_oracle_lambda(N, _dp(lam))
    return lam


# ----------------------------------------------------------------------------------------------
# The reference binary under shims (oracle/_ref), driven the way analytical_class_cli.py:53-68 does
# minus the oneAPI `. vars.sh;` prefix.
# ----------------------------------------------------------------------------------------------
def ref_binary_available() -> bool:
    return REF_BIN.exists() and os.access(REF_BIN, os.X_OK) and (REF_DIR / "libmkl_intel_lp64.so.2").exists()


def write_line_inputs(input_dir: Path, N: int, a, rows, cols, vals, u, t) -> Path:
    """Write B.csv / curden.csv / analytical.cfg like matrix_formulation_func_cli.py:240-286 and
    analytical_class_cli.py:53-57 do (Python repr of each float round-trips exactly)."""
    input_dir = Path(input_dir)
    input_dir.mkdir(parents=True, exist_ok=True)
    with open(input_dir / "B.csv", "w") as f:
        for i, j, v in zip(rows, cols, vals):
            f.write(f"{int(i)},{int(j)},{float(v)!r}\n")
    with open(input_dir / "curden.csv", "w") as f:
        for v in u:
            f.write(f"{float(v)!r}\n")
    cfg = input_dir / "analytical.cfg"
    with open(cfg, "w") as f:
        f.write(f"input_files {input_dir}/\nnum_nodes {int(N)}\n")
        f.write(f"A_coeff {float(a)!r}\nsimulation_time {t}\n")
    return cfg


def stress_filename(t: float) -> str:
    """std::to_string(double) == '%f' (bin/EMtool_analytical@0xbca3)."""
    return "stress_" + ("%f" % float(t)) + ".txt"


def run_binary(cfg: Path, exe: Path = REF_BIN, fft: str = "ld", timeout: float = 600.0):
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = str(REF_DIR) + os.pathsep + env.get("LD_LIBRARY_PATH", "")
    env["EM_SHIM_FFT"] = "double" if fft == "double" else "ld"
    return subprocess.run([str(exe), str(cfg)], env=env, capture_output=True, text=True, timeout=timeout)


def reference_binary_line(N, a, rows, cols, vals, u, t, workdir=None, fft: str = "ld") -> np.ndarray:
    """Run the reference binary for one (line, t); returns the N stresses parsed from its text file."""
    if not ref_binary_available():
        raise RuntimeError("oracle/_ref is not built (reference binary unavailable)")
    ctx = tempfile.TemporaryDirectory() if workdir is None else None
    base = Path(ctx.name if ctx else workdir)
    try:
        input_dir = base / "input" / "line" / "cfg"
        cfg = write_line_inputs(input_dir, N, a, rows, cols, vals, u, t)
        res = run_binary(cfg, fft=fft)
        if res.returncode != 0:
            raise RuntimeError(f"reference binary rc={res.returncode}: {res.stdout}{res.stderr}")
        out_file = Path(re.sub("input", "output", str(input_dir), count=1)) / stress_filename(float(t))
        return np.array([float(x) for x in out_file.read_text().split()], dtype=np.float64)
    finally:
        if ctx:
            ctx.cleanup()


# ---- CPU twin of the GPU arithmetic (oracle/em_twin.c) -----------------------------------------------------------
TWIN_PATH = HERE / "libtwin.so"
_twin = None


def load_twin() -> ctypes.CDLL:
    global _twin
    if _twin is None:
        if not TWIN_PATH.exists():
            build()
        lib = ctypes.CDLL(str(TWIN_PATH))
        c_dp = ctypes.POINTER(ctypes.c_double)
        for f in (lib.em_twin_dense, lib.em_twin_pow2):
            f.argtypes = [ctypes.c_int, ctypes.c_double, c_dp, c_dp, c_dp, ctypes.c_int, c_dp, c_dp]
            f.restype = ctypes.c_int
        lib.em_twin_blue.argtypes = [ctypes.c_int, ctypes.c_double, c_dp, c_dp, c_dp, c_dp, c_dp, ctypes.c_int, c_dp, c_dp]
        lib.em_twin_blue.restype = ctypes.c_int
        lib.em_twin_rows.argtypes = [ctypes.c_int, ctypes.c_double, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, ctypes.c_int,
                                     ctypes.c_int, ctypes.c_int, ctypes.c_int, c_dp, c_dp]
        lib.em_twin_rows.restype = ctypes.c_int
        lib.em_twin_long_pow2.argtypes = [ctypes.c_int, ctypes.c_double, c_dp, c_dp, c_dp, c_dp, ctypes.c_int, c_dp, c_dp]
        lib.em_twin_long_pow2.restype = ctypes.c_int
        _twin = lib
    return _twin


def twin_line(kind: str, N: int, a: float, lam, W, table, ts) -> np.ndarray:
    """Stress [T, N] of one line exactly as the GPU kernel of `kind` ("dense" | "pow2") computes it, given the device's
    own tables: lambda, W, and `table` = the cosine table (dense) or the pre-twiddles (pow2)."""
    lam = np.ascontiguousarray(lam, dtype=np.float64)
    W = np.ascontiguousarray(W, dtype=np.float64)
    table = np.ascontiguousarray(table, dtype=np.float64)
    ts = np.ascontiguousarray(np.atleast_1d(ts), dtype=np.float64)
    out = np.empty((len(ts), N), dtype=np.float64)
    fn = load_twin().em_twin_dense if kind == "dense" else load_twin().em_twin_pow2
    rc = fn(int(N), float(a), _dp(lam), _dp(W), _dp(table), len(ts), _dp(ts), _dp(out))
    if rc:
        raise ValueError(f"em_twin_{kind} rc={rc}")
    return out


def twin_bluestein(N: int, a: float, lam, W, pre, post, hhat, ts, twl=None, lr: int = 0, rows: int = 0,
                   long_path: bool = False) -> np.ndarray:
    """Stress [T, N] of one line exactly as a Bluestein class of the GPU computes it: the in-CTA power-of-two class
    (rows == 0), a mixed-radix class (rows x 2^lr points) or the long path (rows = P1, lr = 12, long_path=True), from the
    device's own tables."""
    c = lambda x: np.ascontiguousarray(x, dtype=np.float64)
    lam, W, pre, post, hhat, ts = c(lam), c(W), c(pre), c(post), c(hhat), c(np.atleast_1d(ts))
    out = np.empty((len(ts), N), dtype=np.float64)
    if rows == 0:
        rc = load_twin().em_twin_blue(int(N), float(a), _dp(lam), _dp(W), _dp(pre), _dp(post), _dp(hhat), len(ts), _dp(ts),
                                      _dp(out))
    else:
        twl = c(twl)
        rc = load_twin().em_twin_rows(int(N), float(a), _dp(lam), _dp(W), _dp(pre), _dp(post), _dp(hhat), _dp(twl), int(lr),
                                      int(rows), int(bool(long_path)), len(ts), _dp(ts), _dp(out))
    if rc:
        raise ValueError(f"em_twin bluestein rc={rc}")
    return out


def twin_long_pow2(N: int, a: float, lam, W, pre, twl, ts) -> np.ndarray:
    """Stress [T, N] of a power-of-two line above 8192 nodes exactly as the GPU's scatter / row / column kernels compute it."""
    c = lambda x: np.ascontiguousarray(x, dtype=np.float64)
    lam, W, pre, twl, ts = c(lam), c(W), c(pre), c(twl), c(np.atleast_1d(ts))
    out = np.empty((len(ts), N), dtype=np.float64)
    rc = load_twin().em_twin_long_pow2(int(N), float(a), _dp(lam), _dp(W), _dp(pre), _dp(twl), len(ts), _dp(ts), _dp(out))
    if rc:
        raise ValueError(f"em_twin_long_pow2 rc={rc}")
    return out

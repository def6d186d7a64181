"""Python front-end of the batched line-analysis library (libproton_em.so).

`LineBatch` is the ragged batch the C ABI consumes (include/proton_em.h: proton_em_batch) — per line the
contents of the reference's `analytical.cfg`, `B.csv` and `curden.csv`.  `Context.pack` uploads it and runs the
packing and forward-DCT kernels; `Plan.eval` evaluates it at any number of timestamps.  Host (numpy) and
device (torch CUDA tensor) entry points are separate so nothing is copied behind the caller's back.
"""
from __future__ import annotations

import ctypes
import os
import weakref
from dataclasses import dataclass
from typing import Iterable, Optional, Sequence

import numpy as np

from . import _lib
from ._lib import ALG_AUTO, ALG_DIRECT, DEVICE, HOST, Batch, Opts, ProtonEmError, Result, check


@dataclass
class LineBatch:
    num_nodes: np.ndarray   # int32 [L]
    a_coeff: np.ndarray     # float64 [L]
    trip_off: np.ndarray    # int64 [L+1]
    trip_row: np.ndarray    # int32
    trip_col: np.ndarray    # int32
    trip_val: np.ndarray    # float64
    u_off: np.ndarray       # int64 [L+1]
    curden: np.ndarray      # float64
    names: Optional[list] = None

    def __post_init__(self):
        self.num_nodes = np.ascontiguousarray(self.num_nodes, dtype=np.int32)
        self.a_coeff = np.ascontiguousarray(self.a_coeff, dtype=np.float64)
        self.trip_off = np.ascontiguousarray(self.trip_off, dtype=np.int64)
        self.trip_row = np.ascontiguousarray(self.trip_row, dtype=np.int32)
        self.trip_col = np.ascontiguousarray(self.trip_col, dtype=np.int32)
        self.trip_val = np.ascontiguousarray(self.trip_val, dtype=np.float64)
        self.u_off = np.ascontiguousarray(self.u_off, dtype=np.int64)
        self.curden = np.ascontiguousarray(self.curden, dtype=np.float64)
        L = len(self.num_nodes)
        if len(self.a_coeff) != L or len(self.trip_off) != L + 1 or len(self.u_off) != L + 1:
            raise ValueError("LineBatch: inconsistent array lengths")

    @property
    def num_lines(self) -> int:
        return len(self.num_nodes)

    @property
    def node_off(self) -> np.ndarray:
        off = np.zeros(self.num_lines + 1, dtype=np.int64)
        np.cumsum(self.num_nodes, out=off[1:])
        return off

    @property
    def total_nodes(self) -> int:
        return int(self.num_nodes.astype(np.int64).sum())

    @staticmethod
    def from_lines(lines: Iterable) -> "LineBatch":
        """From objects with num_nodes, a_coeff, rows, cols, vals, curden (proton_b200.formulation.LineInputs)."""
        lines = list(lines)
        trip_off = np.zeros(len(lines) + 1, dtype=np.int64)
        u_off = np.zeros(len(lines) + 1, dtype=np.int64)
        for i, li in enumerate(lines):
            trip_off[i + 1] = trip_off[i] + len(li.rows)
            u_off[i + 1] = u_off[i] + len(li.curden)
        cat = lambda xs, dt: (np.concatenate(xs).astype(dt) if xs else np.zeros(0, dt))
        return LineBatch(num_nodes=[li.num_nodes for li in lines], a_coeff=[li.a_coeff for li in lines],
                         trip_off=trip_off, trip_row=cat([np.asarray(li.rows) for li in lines], np.int32),
                         trip_col=cat([np.asarray(li.cols) for li in lines], np.int32),
                         trip_val=cat([np.asarray(li.vals) for li in lines], np.float64), u_off=u_off,
                         curden=cat([np.asarray(li.curden) for li in lines], np.float64),
                         names=[getattr(li, "name", str(i)) for i, li in enumerate(lines)])

    def subset(self, idx: Sequence[int]) -> "LineBatch":
        """The lines `idx` (any order, repeats allowed) as a batch of their own."""
        idx = np.asarray(idx, dtype=np.int64)
        tl = self.trip_off[idx + 1] - self.trip_off[idx]
        ul = self.u_off[idx + 1] - self.u_off[idx]
        trip_off = np.concatenate([[0], np.cumsum(tl)]).astype(np.int64)
        u_off = np.concatenate([[0], np.cumsum(ul)]).astype(np.int64)
        # element e of the new triplet array comes from old position start_of_its_line + (e - new_start_of_its_line)
        tsel = np.repeat(self.trip_off[idx] - trip_off[:-1], tl) + np.arange(int(trip_off[-1]), dtype=np.int64)
        usel = np.repeat(self.u_off[idx] - u_off[:-1], ul) + np.arange(int(u_off[-1]), dtype=np.int64)
        return LineBatch(self.num_nodes[idx], self.a_coeff[idx], trip_off, self.trip_row[tsel], self.trip_col[tsel],
                         self.trip_val[tsel], u_off, self.curden[usel],
                         names=[self.names[i] for i in idx] if self.names else None)

    def slice(self, l0: int, l1: int) -> "LineBatch":
        """The contiguous lines [l0, l1) (views where possible)."""
        t0, t1 = int(self.trip_off[l0]), int(self.trip_off[l1])
        c0, c1 = int(self.u_off[l0]), int(self.u_off[l1])
        return LineBatch(self.num_nodes[l0:l1], self.a_coeff[l0:l1], self.trip_off[l0:l1 + 1] - t0, self.trip_row[t0:t1],
                         self.trip_col[t0:t1], self.trip_val[t0:t1], self.u_off[l0:l1 + 1] - c0, self.curden[c0:c1],
                         names=self.names[l0:l1] if self.names else None)

    def _as_c(self) -> Batch:
        p = lambda a: a.ctypes.data if a.size else None
        return Batch(self.num_lines, p(self.num_nodes), p(self.a_coeff), self.trip_off.ctypes.data, p(self.trip_row),
                     p(self.trip_col), p(self.trip_val), self.u_off.ctypes.data, p(self.curden))


@dataclass
class EvalResult:
    stress: Optional[np.ndarray]   # flat; line l at node_off[l]*T, [T][N_l]
    max_stress: np.ndarray         # [L, T]
    argmax: np.ndarray             # [L, T]
    flag: np.ndarray               # [L, T] uint8
    node_off: np.ndarray
    T: int

    def line(self, l: int) -> np.ndarray:
        a, b = int(self.node_off[l]), int(self.node_off[l + 1])
        return self.stress[a * self.T:b * self.T].reshape(self.T, b - a)


def _opts(critical, algorithm) -> Opts:
    return Opts(float("nan") if critical is None else float(critical), int(algorithm), 0)


class Context:
    """One per (process, GPU)."""

    def __init__(self, device: int = 0):
        self._lib = _lib.load()
        self._h = ctypes.c_void_p()
        check(self._lib.proton_em_create(ctypes.byref(self._h), int(device)))
        self.device = int(device)
        self._plans = weakref.WeakSet()     # live plans: they hold pointers into the context and must go first

    def close(self):
        if self._h:
            for plan in list(self._plans):
                plan.close()
            self._lib.proton_em_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self) -> int:
        return int(self._lib.proton_em_launch_count(self._h))

    def profile(self, enable: bool) -> None:
        """Start (clears the totals) or stop the per-kernel CUDA-event timing of the evaluation path."""
        check(self._lib.proton_em_profile(self._h, 1 if enable else 0))

    def profile_read(self) -> dict:
        """{kernel name: (total ms, launches)} accumulated since profile(True)."""
        out, kind = {}, 0
        while True:
            name, ms, n = ctypes.c_char_p(), ctypes.c_double(), ctypes.c_int64()
            if self._lib.proton_em_profile_read(self._h, kind, ctypes.byref(name), ctypes.byref(ms), ctypes.byref(n)) != 0:
                return out
            out[name.value.decode()] = (ms.value, int(n.value))
            kind += 1

    @property
    def stream(self) -> int:
        return int(self._lib.proton_em_stream(self._h) or 0)

    def synchronize(self):
        check(self._lib.proton_em_synchronize(self._h))

    def pack(self, batch: LineBatch, sigma0=None) -> "Plan":
        """`sigma0`: optional initial stress, one value per node (line l at node_off[l]); see proton_em_pack_sigma0."""
        h = ctypes.c_void_p()
        cb = batch._as_c()
        if sigma0 is None:
            check(self._lib.proton_em_pack(self._h, ctypes.byref(cb), HOST, ctypes.byref(h)))
        else:
            s0 = np.ascontiguousarray(sigma0, dtype=np.float64).reshape(-1)
            if s0.size != batch.total_nodes:
                raise ValueError("sigma0 needs one value per node of the batch")
            check(self._lib.proton_em_pack_sigma0(self._h, ctypes.byref(cb), s0.ctypes.data, HOST, ctypes.byref(h)))
        return Plan(self, h, batch.node_off, keep=None)

    def pack_project(self, project_dir: str, variant: Optional[str] = None) -> "Plan":
        """Every discretised line of a PROTON project directory (`<project>/input/<line>/<variant>/analytical.txt,
        B.csv, curden.csv`) in one plan, read and parsed inside the library; see proton_em_pack_from_project.
        `plan.names` are the line names, `plan.skipped` the number of lines left out, `plan.message` the text of the
        first failure."""
        h, skipped = ctypes.c_void_p(), ctypes.c_int64()
        msg = ctypes.create_string_buffer(1024)
        rc = self._lib.proton_em_pack_from_project(self._h, os.fsencode(project_dir),
                                                   variant.encode() if variant else None, ctypes.byref(h),
                                                   ctypes.byref(skipped), msg, len(msg))
        if rc:
            raise ProtonEmError(rc, msg.value.decode(errors="replace").strip())
        L = int(self._lib.proton_em_plan_num_lines(h))
        node_off = np.zeros(L + 1, dtype=np.int64)
        np.cumsum([self._lib.proton_em_plan_line_nodes(h, l) for l in range(L)], out=node_off[1:])
        plan = Plan(self, h, node_off, keep=None)
        plan.skipped, plan.message = int(skipped.value), msg.value.decode(errors="replace")
        return plan

    def pack_device(self, num_nodes, a_coeff, trip_off, trip_row, trip_col, trip_val, u_off, curden) -> "Plan":
        """All arguments are torch CUDA tensors (int32 / float64 / int64 as in proton_em_batch); the plan keeps
        references to them because the kernels read num_nodes / a_coeff / trip_off in place."""
        keep = (num_nodes, a_coeff, trip_off, trip_row, trip_col, trip_val, u_off, curden)
        for t in keep:
            if not t.is_cuda or not t.is_contiguous():
                raise ValueError("pack_device needs contiguous CUDA tensors")
        cb = Batch(int(num_nodes.numel()), num_nodes.data_ptr(), a_coeff.data_ptr(), trip_off.data_ptr(),
                   trip_row.data_ptr(), trip_col.data_ptr(), trip_val.data_ptr(), u_off.data_ptr(), curden.data_ptr())
        h = ctypes.c_void_p()
        check(self._lib.proton_em_pack(self._h, ctypes.byref(cb), DEVICE, ctypes.byref(h)))
        node_off = np.zeros(int(num_nodes.numel()) + 1, dtype=np.int64)
        np.cumsum(num_nodes.cpu().numpy(), out=node_off[1:])
        return Plan(self, h, node_off, keep=keep)

    def analyze_host(self, batch: LineBatch, ts, critical=None, want_stress=True, algorithm=ALG_AUTO,
                     max_device_bytes: int = 0, out: Optional[EvalResult] = None) -> EvalResult:
        """pack + eval + copy back in one C call (the end-to-end path).  `out` may hold preallocated (e.g.
        pinned) arrays to write into."""
        ts = np.ascontiguousarray(np.atleast_1d(ts), dtype=np.float64)
        T, L = len(ts), batch.num_lines
        if out is None:
            out = EvalResult(np.empty(batch.total_nodes * T) if want_stress else None, np.empty((L, T)),
                             np.empty((L, T), np.int32), np.empty((L, T), np.uint8), batch.node_off, T)
        res = Result(out.stress.ctypes.data if out.stress is not None else None, out.max_stress.ctypes.data,
                     out.argmax.ctypes.data, out.flag.ctypes.data)
        cb = batch._as_c()
        o = _opts(critical, algorithm)
        check(self._lib.proton_em_analyze_host(self._h, ctypes.byref(cb), ts.ctypes.data, T, ctypes.byref(o),
                                               ctypes.byref(res), int(max_device_bytes)))
        return out

    def analyze_history(self, batch: LineBatch, curden_segments, seg_start, ts, critical=None, want_stress=True,
                        algorithm=ALG_AUTO) -> EvalResult:
        """Stress under a piecewise-constant current history: segment s applies `curden_segments[s]` (each laid out
        like `batch.curden`) from `seg_start[s]` on (seg_start[0] == 0).  See proton_em_analyze_history_host."""
        ts = np.ascontiguousarray(np.atleast_1d(ts), dtype=np.float64)
        seg_start = np.ascontiguousarray(np.atleast_1d(seg_start), dtype=np.float64)
        segs = [np.ascontiguousarray(c, dtype=np.float64) for c in curden_segments]
        if len(segs) != len(seg_start):
            raise ValueError("one start time per current segment")
        for c in segs:
            if c.shape != np.asarray(batch.curden).shape:
                raise ValueError("every segment needs one current density per entry of batch.curden")
        T, L = len(ts), batch.num_lines
        out = EvalResult(np.empty(batch.total_nodes * T) if want_stress else None, np.empty((L, T)),
                         np.empty((L, T), np.int32), np.empty((L, T), np.uint8), batch.node_off, T)
        res = Result(out.stress.ctypes.data if out.stress is not None else None, out.max_stress.ctypes.data,
                     out.argmax.ctypes.data, out.flag.ctypes.data)
        ptrs = (ctypes.c_void_p * len(segs))(*[c.ctypes.data for c in segs])
        cb = batch._as_c()
        o = _opts(critical, algorithm)
        check(self._lib.proton_em_analyze_history_host(self._h, ctypes.byref(cb), ptrs, seg_start.ctypes.data, len(segs),
                                                       ts.ctypes.data, T, ctypes.byref(o), ctypes.byref(res)))
        return out

    def debug_exp(self, x) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty_like(x)
        check(self._lib.proton_em_debug_exp(self._h, x.ctypes.data, y.ctypes.data, x.size))
        return y

    def debug_expm1(self, x) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty_like(x)
        check(self._lib.proton_em_debug_expm1(self._h, x.ctypes.data, y.ctypes.data, x.size))
        return y

    def run_cfg(self, cfg_path: str):
        """The reference process for one analytical.cfg -> (return code, stdout text of the reference)."""
        buf = ctypes.create_string_buffer(2048)
        rc = self._lib.proton_em_run_cfg(self._h, str(cfg_path).encode(), buf, len(buf))
        return rc, buf.value.decode(errors="replace")


    def run_cfgs(self, cfg_paths):
        """Many analytical.cfg files as one batch (proton_em_run_cfgs) -> (failed count or negative code, per-cfg status
        array, stdout text of the first failure)."""
        paths = [str(p).encode() for p in cfg_paths]
        arr = (ctypes.c_char_p * len(paths))(*paths)
        status = np.zeros(len(paths), dtype=np.int32)
        buf = ctypes.create_string_buffer(2048)
        rc = self._lib.proton_em_run_cfgs(self._h, arr, len(paths), status.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
                                          buf, len(buf))
        return rc, status, buf.value.decode(errors="replace")


class Plan:
    def __init__(self, ctx: Context, handle, node_off: np.ndarray, keep):
        self.ctx = ctx
        self._h = handle
        self.node_off = node_off
        self._keep = keep
        ctx._plans.add(self)

    @property
    def num_lines(self) -> int:
        return len(self.node_off) - 1

    @property
    def total_nodes(self) -> int:
        return int(self.node_off[-1])

    @property
    def scratch_bytes(self) -> int:
        return int(self.ctx._lib.proton_em_plan_scratch_bytes(self._h))

    @property
    def names(self) -> Optional[list]:
        """Line names held by the plan (proton_em_plan_set_line_names / pack_project), or None."""
        first = self.ctx._lib.proton_em_plan_line_name(self._h, 0)
        if first is None:
            return None
        return [self.ctx._lib.proton_em_plan_line_name(self._h, l).decode() for l in range(self.num_lines)]

    @names.setter
    def names(self, names) -> None:
        arr = (ctypes.c_char_p * len(names))(*[str(n).encode() for n in names])
        check(self.ctx._lib.proton_em_plan_set_line_names(self._h, arr, len(names)))

    def close(self):
        if self._h:
            if self.ctx._h:                      # a context that is already gone took its plans with it
                self.ctx._lib.proton_em_plan_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    TABLES = {"lambda": 0, "W": 1, "pre": 2, "post": 3, "costab": 4, "hhat": 5, "twl": 6}

    def debug_table(self, line: int, which: str) -> np.ndarray:
        """Test hook: one of the device tables the kernels read for `line` (float64; complex tables interleaved)."""
        N = int(self.node_off[line + 1] - self.node_off[line])
        out = np.empty(max(4 * N + 8, 2 * 131072 + 8), dtype=np.float64)      # (H of the longest class: 131072 complex)
        n = ctypes.c_int64()
        check(self.ctx._lib.proton_em_debug_line_table(self._h, int(line), self.TABLES[which], out.ctypes.data, out.size,
                                                       ctypes.byref(n)))
        return out[: n.value].copy()

    def eval(self, ts, critical=None, want_stress=True, algorithm=ALG_AUTO) -> EvalResult:
        """Host call: timestamps in, numpy results out (copies and synchronises)."""
        ts = np.ascontiguousarray(np.atleast_1d(ts), dtype=np.float64)
        T, L = len(ts), self.num_lines
        out = EvalResult(np.empty(self.total_nodes * T) if want_stress else None, np.empty((L, T)),
                         np.empty((L, T), np.int32), np.empty((L, T), np.uint8), self.node_off, T)
        res = Result(out.stress.ctypes.data if want_stress else None, out.max_stress.ctypes.data,
                     out.argmax.ctypes.data, out.flag.ctypes.data)
        o = _opts(critical, algorithm)
        check(self.ctx._lib.proton_em_eval(self._h, ts.ctypes.data, T, ctypes.byref(o), ctypes.byref(res), HOST))
        return out

    def eval_device(self, ts, stress, max_stress, argmax, flag, critical=None, algorithm=ALG_AUTO) -> None:
        """Device call: all arguments are torch CUDA tensors (stress may be None); work is queued on the
        context's stream and NOT synchronised."""
        T = int(ts.numel())
        res = Result(stress.data_ptr() if stress is not None else None, max_stress.data_ptr(), argmax.data_ptr(),
                     flag.data_ptr())
        o = _opts(critical, algorithm)
        check(self.ctx._lib.proton_em_eval(self._h, ts.data_ptr(), T, ctypes.byref(o), ctypes.byref(res), DEVICE))

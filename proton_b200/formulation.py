"""Per-line FDM formulation: the in-memory twin of PROTON's `Matrix_Formulation_Class`.

The reference (matrix_formulation_func_cli.py:24-299) builds dense N x N / N x S Python lists and
writes `B.csv`, `curden.csv` and `analytical.txt` for the C++ kernel.  This module produces the same
numbers — N, `Acoef`, the non-zero triplets of `B` in the order the reference writes them, and the
float32-rounded current densities — without the O(N^2) dense build and without touching disk, so a
whole power grid can be handed to the batched GPU path.  `read_line_inputs` reads the files the
reference emitted instead, for callers that already ran it.

Reference quirks reproduced on purpose (each one changes the numbers the kernel sees):
  * constants exactly as written, incl. `b = Ea*Z*rho/Omega`            (:48-73)
  * points mode: dx = round(total/(points + S - 1), 2) um               (:86-102)
  * nodes per segment = round(len/dx) with Python's banker's rounding   (:133)
  * a segment is joined to its predecessor (shares a via node) only when the coordinate that is
    constant along the CURRENT segment equals the predecessor's                (:148-157)
  * `B` is assigned, not accumulated: later assignments overwrite earlier ones, row index -1
    wraps to the last row (Python list indexing), and the closing assignment
    B[N-1][S-1] = -1 overwrites whatever was there (e.g. the N == 1 case)      (:191-204)
  * current density is rounded to float32 (`array.array('f')`)                 (:121,180)
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

NOT_DENSE_MSG = "The given discretization is not dense enough for the geometry of the given line."
NO_METHOD_MSG = "Error no discretization method is choosen!"
CORRUPT_MSG = "Corrupted CSV segments file."
NOT_FOUND_MSG = "CSV file not found."


class FormulationError(ValueError):
    """Raised with the same message text the reference returns/prints for this condition."""


@dataclass
class LineInputs:
    """What the analytical kernel consumes for one line."""
    name: str
    num_nodes: int            # nx_total  -> analytical.cfg num_nodes
    a_coeff: float            # Acoef     -> analytical.cfg A_coeff
    rows: np.ndarray          # B.csv triplets, int32, in file order (row-major over the dense B)
    cols: np.ndarray
    vals: np.ndarray          # float64 (= +-bcoef)
    curden: np.ndarray        # float64 holding float32-rounded values, length S
    dx: float = 0.0
    bcoef: float = 0.0


def technology_constants(technology: str, temperature: float) -> Tuple[float, float]:
    """(kappa, b) for 'CuDD' or anything else (= the reference's Al branch), :48-73."""
    T = float(temperature)
    if technology == "CuDD":
        D0, Ea, kB, B0, Omega, rho, Z = 1.3e-9, 1.609e-19, 1.3806e-23, 28e9, 1.182e-29, 2.25e-8, 1
    else:
        D0, Ea, kB, B0, Omega, rho, Z = 1e-4, 1.92e-29, 1.3806e-23, 80e9, 2e-29, 3.16e-8, 1
    Da = D0 * math.exp(-Ea / (kB * T))
    kappa = (Da * B0 * Omega) / (kB * T)
    b = Ea * Z * rho / Omega
    return kappa, b


def parse_line_csv(csv_file: str) -> Tuple[List[Tuple[str, str, float, float]], int]:
    """Rows of a `<layer>_<net>_<k>.csv` (spice_function_cli.py:263,275) that start with 'R':
    (left_node, right_node, length_um, current_A).  Node names keep the leading blank the reference's
    `split(',')` leaves in place.  Also returns the raw line count (used by points mode)."""
    try:
        with open(csv_file) as f:
            lines = f.readlines()
    except FileNotFoundError:
        raise FormulationError(NOT_FOUND_MSG)
    segs = []
    try:
        for line in lines:
            if line and line[0] == "R":
                c = line.split(",")
                segs.append((c[1], c[2], float(c[3]), float(c[4])))
    except (ValueError, IndexError):
        raise FormulationError(CORRUPT_MSG)
    return segs, len(lines)


def formulate_segments(segs: Sequence[Tuple[str, str, float, float]], *, sp_step=None, given_disc_point=None,
                       technology="CuDD", temperature=378.0, givenWidth=0.1, name="line",
                       n_file_lines: Optional[int] = None) -> LineInputs:
    """Discretise one line given its segments (left_node, right_node, length_um, current_A)."""
    if sp_step is None and given_disc_point is None:
        raise FormulationError(NO_METHOD_MSG)
    kappa, b = technology_constants(technology, temperature)
    S_all = len(segs)
    dx = 0.0
    if given_disc_point is not None:
        total_length = 0.0
        for s in segs:
            total_length = total_length + float(s[2])
        nl = (S_all + 1) if n_file_lines is None else n_file_lines
        gdp = float(given_disc_point) + (nl - 2)
        dx = round(float(total_length / gdp), 2)
        dx = dx * 1e-6
    w = float(givenWidth)
    nx_total = 0
    num_segments = 0
    curden_list = []
    vias_nodes: List[int] = []
    vias = {}
    left_node = right_node = None
    for (ln, rn, length_um, current) in segs:
        num_segments += 1
        length = float(length_um) * 1e-6
        curden_list.append(float(current) / (w * 1e-6 * w * 1e-6))
        if sp_step is not None:
            dx = float(sp_step) * 1e-6
        if length < dx:
            raise FormulationError(NOT_DENSE_MSG)
        disc_points = round(length / dx)
        nx_total += disc_points
        prev_left, prev_right = left_node, right_node
        left_node, right_node = ln, rn
        lc, rc = left_node.split("_"), right_node.split("_")
        same = 2 if lc[2] == rc[2] else 1
        if prev_left and prev_right and prev_left.split("_")[same] == lc[same]:
            key = nx_total - disc_points
            vias[key] = num_segments
            vias_nodes.append(key)
            nx_total -= 1
    N, S = nx_total, num_segments
    if N <= 0 or S <= 0:
        raise FormulationError(CORRUPT_MSG)
    entries = {}

    def assign(i: int, j: int, v: float) -> None:
        if i < 0:
            i += N
        if j < 0:
            j += S
        if not (0 <= i < N and 0 <= j < S):
            raise IndexError("list assignment index out of range")
        entries[(i, j)] = v

    assign(0, 0, 1.0)
    for node in vias_nodes:
        segment = vias[node]
        assign(node - 1, segment - 2, -1.0)
        assign(node - 1, segment - 1, 1.0)
    assign(N - 1, S - 1, -1.0)
    acoef = kappa / (dx * dx)
    bcoef = kappa * b / dx
    keys = sorted(entries)
    rows = np.array([k[0] for k in keys], dtype=np.int32)
    cols = np.array([k[1] for k in keys], dtype=np.int32)
    vals = np.array([entries[k] * bcoef for k in keys], dtype=np.float64)
    curden = np.asarray(curden_list, dtype=np.float64).astype(np.float32).astype(np.float64)
    return LineInputs(name=name, num_nodes=int(N), a_coeff=float(acoef), rows=rows, cols=cols, vals=vals,
                      curden=curden, dx=float(dx), bcoef=float(bcoef))


def formulate_csv(csv_file: str, *, sp_step=None, given_disc_point=None, technology="CuDD", temperature=378.0,
                  givenWidth=0.1) -> LineInputs:
    """`Matrix_Formulation_Class(csv_file, ..).matrix_formulation()` without the files."""
    segs, nlines = parse_line_csv(csv_file)
    name = os.path.splitext(os.path.basename(csv_file))[0]
    return formulate_segments(segs, sp_step=sp_step, given_disc_point=given_disc_point, technology=technology,
                              temperature=temperature, givenWidth=givenWidth, name=name, n_file_lines=nlines)


def analyze_policy_step(segs: Sequence[Tuple[str, str, float, float]]) -> float:
    """The whole-grid driver's choice of step: 0.99 x the shortest segment (PROTON_cli.py:1005-1012)."""
    return float(min(float(s[2]) for s in segs) * 0.99)


def read_line_inputs(input_dir: str, name: Optional[str] = None) -> LineInputs:
    """Read what the reference's formulation wrote for one line: `analytical.txt`
    (analytical_class_cli.py:31-41), `B.csv` (int,int,double triplets; separators ',', blank, tab as in
    tripletParser, bin/EMtool_analytical@0xb010) and `curden.csv` (one double per line, @0xba39)."""
    nx_total, acoef = None, None
    with open(os.path.join(input_dir, "analytical.txt")) as f:
        for line in f:
            words = line.split()
            if "Acoef" in words:
                acoef = float(words[2])
            if "nx_total" in words:
                nx_total = int(words[2])
    if nx_total is None or acoef is None:
        raise FormulationError(f"analytical.txt in {input_dir} lacks nx_total/Acoef")
    rows, cols, vals = [], [], []
    with open(os.path.join(input_dir, "B.csv")) as f:
        for k, line in enumerate(f, start=1):
            tok = line.replace(",", " ").split()
            if not tok:
                continue
            if len(tok) < 2:
                raise FormulationError(f"Required second value at line {k} of the file B.csv")
            if len(tok) < 3:
                raise FormulationError(f"Required third value at line {k} of the file B.csv")
            rows.append(int(tok[0])), cols.append(int(tok[1])), vals.append(float(tok[2]))
    with open(os.path.join(input_dir, "curden.csv")) as f:
        curden = [float(x) for x in f.read().split()]
    return LineInputs(name=name or os.path.basename(os.path.normpath(input_dir)), num_nodes=nx_total,
                      a_coeff=acoef, rows=np.asarray(rows, np.int32), cols=np.asarray(cols, np.int32),
                      vals=np.asarray(vals, np.float64), curden=np.asarray(curden, np.float64))


# ---------------------------------------------------------------------------------------------------------------
# The same discretisation for a whole grid at once: a flat segment table in, a ragged LineBatch out.
# ---------------------------------------------------------------------------------------------------------------
@dataclass
class SegmentTable:
    """All segments of many lines, flat.  Segment k belongs to line `line_of[k]`; `seg_off[l]:seg_off[l+1]` are the
    segments of line l in file order.  `lcoord[k] / rcoord[k]` are integer ids of the two `_`-separated coordinates
    of the left / right node name (equal id <=> equal string, which is what the reference compares, :148-157)."""
    names: list
    seg_off: np.ndarray       # int64 [L+1]
    length_um: np.ndarray     # float64 [K]
    current: np.ndarray       # float64 [K]
    lcoord: np.ndarray        # int64 [K, 2]   ids of left_node.split("_")[1], [2]
    rcoord: np.ndarray        # int64 [K, 2]
    named: np.ndarray         # bool [K]: both node names non-empty (the reference's truthiness test)

    @property
    def num_lines(self) -> int:
        return len(self.seg_off) - 1

    @staticmethod
    def from_lines(lines: Sequence[Tuple[str, Sequence[Tuple[str, str, float, float]]]]) -> "SegmentTable":
        """From (name, [(left_node, right_node, length_um, current_A), ...]) per line (parse_line_csv's rows)."""
        names = [n for n, _ in lines]
        counts = np.fromiter((len(s) for _, s in lines), dtype=np.int64, count=len(lines))
        seg_off = np.zeros(len(lines) + 1, dtype=np.int64)
        np.cumsum(counts, out=seg_off[1:])
        flat = [s for _, segs in lines for s in segs]
        K = len(flat)
        length = np.fromiter((float(s[2]) for s in flat), dtype=np.float64, count=K)
        current = np.fromiter((float(s[3]) for s in flat), dtype=np.float64, count=K)
        words = {}

        def ident(w):
            return words.setdefault(w, len(words))
        lc = np.empty((K, 2), dtype=np.int64)
        rc = np.empty((K, 2), dtype=np.int64)
        named = np.empty(K, dtype=bool)
        for k, s in enumerate(flat):
            a, b = s[0].split("_"), s[1].split("_")
            lc[k, 0], lc[k, 1] = ident(a[1]), ident(a[2])
            rc[k, 0], rc[k, 1] = ident(b[1]), ident(b[2])
            named[k] = bool(s[0]) and bool(s[1])
        return SegmentTable(names, seg_off, length, current, lc, rc, named)


def formulate_table(tab: SegmentTable, *, sp_step=None, technology="CuDD", temperature=378.0, givenWidth=0.1,
                    engine: str = "native"):
    """Discretise every line of `tab` in step mode in one pass over all segments; returns a `proton_b200.api.LineBatch`
    whose arrays are bit-identical to running `formulate_segments` line by line (same N, Acoef, B triplets in the
    same order, float32-rounded currents).  `sp_step`: one step in um for all lines, an array with one per line, or
    None for the whole-grid driver's policy 0.99 x the line's shortest segment (PROTON_cli.py:1005-1012).
    `engine`: "native" = proton_em_formulate_segments of libproton_em.so (host code, no GPU needed; 0.2 us per line),
    "numpy" = the same in NumPy array operations (2.6 us per line; the cross-check of tests/test_formulation.py)."""
    if engine == "native":
        return _formulate_table_native(tab, sp_step, technology, temperature, givenWidth)
    if engine != "numpy":
        raise ValueError("engine must be 'native' or 'numpy'")
    from .api import LineBatch
    kappa, b = technology_constants(technology, temperature)
    L, K = tab.num_lines, len(tab.length_um)
    counts = np.diff(tab.seg_off)
    if L == 0 or np.any(counts <= 0):
        raise FormulationError(CORRUPT_MSG)
    line_of = np.repeat(np.arange(L, dtype=np.int64), counts)
    first = np.zeros(K, dtype=bool)
    first[tab.seg_off[:-1]] = True
    if sp_step is None:
        step_um = np.minimum.reduceat(tab.length_um, tab.seg_off[:-1]) * 0.99
    else:
        step_um = np.broadcast_to(np.asarray(sp_step, dtype=np.float64), (L,)).copy()
    dx_line = step_um * 1e-6
    dx = dx_line[line_of]
    length = tab.length_um * 1e-6
    if np.any(length < dx):
        raise FormulationError(NOT_DENSE_MSG)
    w = float(givenWidth)
    curden = (tab.current / (w * 1e-6 * w * 1e-6)).astype(np.float32).astype(np.float64)
    d = np.rint(length / dx).astype(np.int64)                  # Python's round(): half to even
    # joined to the predecessor: the coordinate that is constant along THIS segment equals the predecessor's
    same2 = tab.lcoord[:, 1] == tab.rcoord[:, 1]              # split("_")[2] equal -> compare index 2, else index 1
    cur = np.where(same2, tab.lcoord[:, 1], tab.lcoord[:, 0])
    prev = np.empty(K, dtype=np.int64)
    prev[1:] = np.where(same2[1:], tab.lcoord[:-1, 1], tab.lcoord[:-1, 0])
    prev[0] = -1
    prev_named = np.empty(K, dtype=bool)
    prev_named[1:] = tab.named[:-1]
    prev_named[0] = False
    join = (~first) & prev_named & (prev == cur)
    net = d - join
    csum = np.cumsum(net)
    before_line = csum[tab.seg_off[:-1]] - net[tab.seg_off[:-1]]          # running sum before each line's first segment
    after = csum - np.repeat(before_line, counts)              # nx_total after each segment, within its line
    N = after[tab.seg_off[1:] - 1]
    S = counts
    if np.any(N <= 0):
        raise FormulationError(CORRUPT_MSG)
    v = np.nonzero(join)[0]
    nv = len(v)
    vline = line_of[v]
    vkey = after[v] - d[v] + 1                                 # nx_total before the segment (the via's node)
    vseg = v - tab.seg_off[vline] + 1                          # 1-based segment number
    # the reference keeps `vias[key] = segment` in a dict: a repeated key keeps the LAST segment number, and the
    # repeated vias then assign the same two entries again.  Keys do not decrease along a line, so equal keys are
    # adjacent: only the last via of each run matters.
    if nv:
        run_end = np.ones(nv, dtype=bool)
        run_end[:-1] = (vline[1:] != vline[:-1]) | (vkey[1:] != vkey[:-1])
        if not run_end.all():
            vline, vkey, vseg = vline[run_end], vkey[run_end], vseg[run_end]
            nv = len(vline)
    # Assignments in the reference's order: (0, 0, +1); per via (key-1, seg-2, -1), (key-1, seg-1, +1); (N-1, S-1, -1),
    # laid out line after line.  Rows increase strictly from via to via, so the list is already sorted by (row, col)
    # and two assignments to one position (the first / last entry against a via's) are neighbours: the later one wins.
    vcount = np.bincount(vline, minlength=L) if nv else np.zeros(L, dtype=np.int64)
    vb = np.zeros(L + 1, dtype=np.int64)
    np.cumsum(vcount, out=vb[1:])
    total = 2 * L + 2 * nv
    lines_all = np.arange(L, dtype=np.int64)
    a_line = np.repeat(lines_all, 2 + 2 * vcount)
    a_i = np.empty(total, dtype=np.int64)
    a_j = np.empty(total, dtype=np.int64)
    a_v = np.empty(total, dtype=np.float64)
    i_first = 2 * lines_all + 2 * vb[:-1]
    i_last = 2 * lines_all + 2 * vb[1:] + 1
    i_via = 2 * np.arange(nv, dtype=np.int64) + 2 * vline + 1
    a_i[i_first], a_j[i_first], a_v[i_first] = 0, 0, 1.0
    a_i[i_via], a_j[i_via], a_v[i_via] = vkey - 1, vseg - 2, -1.0
    a_i[i_via + 1], a_j[i_via + 1], a_v[i_via + 1] = vkey - 1, vseg - 1, 1.0
    a_i[i_last], a_j[i_last], a_v[i_last] = N - 1, S - 1, -1.0
    if nv and (vkey.min() < 1 or vseg.min() < 2):
        # never seen (a via sits after at least one node and one segment); kept for exactness: Python list indexing
        # wraps negative indices once, and then the list is no longer sorted
        Ni, Si = N[a_line], S[a_line]
        a_i = np.where(a_i < 0, a_i + Ni, a_i)
        a_j = np.where(a_j < 0, a_j + Si, a_j)
        if np.any((a_i < 0) | (a_i >= Ni) | (a_j < 0) | (a_j >= Si)):
            raise IndexError("list assignment index out of range")
        pos_key = (a_line * (int(N.max()) + 1) + a_i) * (int(S.max()) + 1) + a_j
        order = np.argsort(pos_key, kind="stable")
        a_line, a_i, a_j, a_v = a_line[order], a_i[order], a_j[order], a_v[order]
    keep = np.ones(total, dtype=bool)                         # later assignments overwrite earlier ones
    keep[:-1] = (a_i[1:] != a_i[:-1]) | (a_j[1:] != a_j[:-1]) | (a_line[1:] != a_line[:-1])
    if not keep.all():
        a_line, a_i, a_j, a_v = a_line[keep], a_i[keep], a_j[keep], a_v[keep]
    acoef = kappa / (dx_line * dx_line)
    bcoef = kappa * b / dx_line
    trip_off = np.zeros(L + 1, dtype=np.int64)
    np.cumsum(np.bincount(a_line, minlength=L), out=trip_off[1:])
    return LineBatch(N.astype(np.int32), acoef, trip_off, a_i.astype(np.int32), a_j.astype(np.int32),
                     a_v * bcoef[a_line], tab.seg_off.copy(), curden, names=list(tab.names))


def _formulate_table_native(tab: SegmentTable, sp_step, technology, temperature, givenWidth):
    """formulate_table through the C ABI (include/proton_em.h: proton_em_formulate_segments)."""
    import ctypes
    from . import _lib
    from .api import LineBatch
    lib = _lib.load()
    kappa, b = technology_constants(technology, temperature)
    L, K = tab.num_lines, len(tab.length_um)
    c = lambda a, dt: np.ascontiguousarray(a, dtype=dt)
    seg_off, length_um, current = c(tab.seg_off, np.int64), c(tab.length_um, np.float64), c(tab.current, np.float64)
    lcoord, rcoord, named = c(tab.lcoord, np.int64), c(tab.rcoord, np.int64), c(tab.named, np.uint8)
    step = None if sp_step is None else np.broadcast_to(np.asarray(sp_step, dtype=np.float64), (L,)).copy()
    w = float(givenWidth)
    segs = _lib.Segments(L, seg_off.ctypes.data, length_um.ctypes.data, current.ctypes.data, lcoord.ctypes.data,
                         rcoord.ctypes.data, named.ctypes.data)
    num_nodes, a_coeff = np.empty(max(L, 1), np.int32), np.empty(max(L, 1), np.float64)
    trip_off = np.zeros(L + 1, np.int64)
    rows, cols, vals = np.empty(2 * K + 2, np.int32), np.empty(2 * K + 2, np.int32), np.empty(2 * K + 2, np.float64)
    curden = np.empty(max(K, 1), np.float64)
    bad = ctypes.c_int64(-1)
    rc = lib.proton_em_formulate_segments(ctypes.byref(segs), None if step is None else step.ctypes.data, kappa, b,
                                          w * 1e-6 * w * 1e-6, num_nodes.ctypes.data, a_coeff.ctypes.data,
                                          trip_off.ctypes.data, rows.ctypes.data, cols.ctypes.data, vals.ctypes.data,
                                          curden.ctypes.data, ctypes.byref(bad))
    if rc == _lib.FORM_INDEX:        # Python's negative list indices: only the NumPy formulation reproduces them
        return formulate_table(tab, sp_step=sp_step, technology=technology, temperature=temperature,
                               givenWidth=givenWidth, engine="numpy")
    if rc == _lib.FORM_NOT_DENSE:
        raise FormulationError(NOT_DENSE_MSG)
    if rc != 0:
        raise FormulationError(CORRUPT_MSG)
    n = int(trip_off[L])
    return LineBatch(num_nodes[:L], a_coeff[:L], trip_off, rows[:n].copy(), cols[:n].copy(), vals[:n].copy(),
                     seg_off.copy(), curden[:K], names=list(tab.names))

"""Whole-grid EM check: the batched twin of PROTON's `analyze <t> --critical s --sample k` command.

The reference walks `<project>/<benchmark>/**/<layer>_<net>_<k>.csv`, and per line (one thread + one process each)
discretises with dx = 0.99 x the shortest segment, runs the analytical kernel, re-reads the stress file, takes the
maximum and compares it with the critical stress (PROTON_cli.py:992-1073, :1076-1214).  Here the same lines are
formulated in memory (proton_b200.formulation), packed once and evaluated in ONE batched GPU call; the per-line maxima
and flags come straight from the fused epilogue.  Optionally the reference's on-disk outputs
(`output/<line>/<tech>_<T>_<W>/stress_<%f>.txt`, one `%.18g` per line) are written so that `report_line_stress`, the
GUI's `add_plot` and `export_function` keep working.

The reference sets a module-global discretisation step from many pool threads (a race); here dx is a
local of each line, so results are deterministic.
"""
from __future__ import annotations

import os
import random
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import formulation as F
from .api import Context, LineBatch


@dataclass
class AnalyzeReport:
    lines: List[str]
    discarded: Dict[str, str]                 # line -> reason (reference prints "The line was discarded")
    max_stress: np.ndarray                    # [L, T]
    argmax: np.ndarray                        # [L, T]
    flag: np.ndarray                          # [L, T]
    problematic_lines: Dict[str, float]       # line -> max stress at the LAST timestamp, if above critical
    seconds: float
    timestamps: np.ndarray = field(default_factory=lambda: np.zeros(0))

    @property
    def max_stress_line(self) -> Optional[str]:
        """The reference's final report: the line with the largest stress (PROTON_cli.py:1197-1213)."""
        if not self.problematic_lines:
            return None
        return max(self.problematic_lines, key=self.problematic_lines.get)

    def message(self) -> str:
        return (f"EM stress analysis of {len(self.lines)} lines was successfully performed in {self.seconds:.3f} seconds; "
                f"{len(self.problematic_lines)} line(s) exceed the critical stress.")


def find_line_csvs(benchmark_dir: str) -> List[str]:
    """`<benchmark>/<layer>/<layer>_<net>_<k>.csv` (spice_function_cli.py:253-263), sorted for determinism."""
    out = []
    for root, _dirs, files in os.walk(benchmark_dir):
        for f in files:
            if f.endswith(".csv"):
                out.append(os.path.join(root, f))
    return sorted(out)


def formulate_for_analyze(csv_files: Sequence[str], technology: str, temperature: float, width: float):
    """dx policy of analyze_line_item: step = 0.99 * min length, up to 10 retries with min length * 0.7
    (PROTON_cli.py:1000-1014).  All lines that parse are discretised in one vectorised pass
    (formulation.formulate_table); a line that the first pass rejects goes through the reference's retry loop on its
    own.  Returns (LineBatch with names, {discarded line: reason})."""
    parsed, discarded = [], {}
    for path in csv_files:
        name = os.path.splitext(os.path.basename(path))[0]
        try:
            segs, nlines = F.parse_line_csv(path)
        except F.FormulationError as e:
            discarded[name] = str(e)
            continue
        if not segs:
            discarded[name] = F.CORRUPT_MSG
            continue
        parsed.append((name, segs, nlines))
    if not parsed:
        return LineBatch.from_lines([]), discarded
    try:
        table = F.SegmentTable.from_lines([(n, s) for n, s, _ in parsed])
        return F.formulate_table(table, technology=technology, temperature=temperature, givenWidth=width), discarded
    except (F.FormulationError, IndexError):
        pass                                   # some line needs the retry loop or is malformed: line by line below
    lines = []
    for name, segs, nlines in parsed:
        try:
            min_length = min(float(s[2]) for s in segs)
            li = None
            for _ in range(10):
                try:
                    li = F.formulate_segments(segs, sp_step=min_length * 0.99, technology=technology,
                                              temperature=temperature, givenWidth=width, name=name, n_file_lines=nlines)
                    break
                except F.FormulationError as e:
                    if str(e) != F.NOT_DENSE_MSG:
                        raise
                    min_length *= 0.7
                    if min_length <= 0:
                        break
            if li is None:
                discarded[name] = "The line was discarded"
            else:
                lines.append(li)
        except (F.FormulationError, IndexError) as e:
            discarded[name] = str(e)
    return LineBatch.from_lines(lines), discarded


def write_reference_outputs(project_dir: str, names, result, timestamps, technology, temperature, width) -> int:
    """stress_<%f>.txt per (line, t) exactly where and how the reference binary writes them; returns the number of files.
    The file name carries `std::to_string(t)`, i.e. six decimals (bin/EMtool_analytical@0xbca3): timestamps that print
    the same way share one file, and the reference — one process per (line, t), run one after the other — leaves the
    LAST of them in it (the readers match `stress_([\\d.]+)\\.` and compare floats,
    PROTON_cli.py:1039-1048).  Same here: only the last timestamp of each file name is written."""
    ts = [float(t) for t in timestamps]
    fname = ["stress_" + ("%f" % t) + ".txt" for t in ts]
    last = {f: j for j, f in enumerate(fname)}             # a later timestamp with the same name wins
    n = 0
    for l, name in enumerate(names):
        out_dir = os.path.join(project_dir, "output", name, f"{technology}_{temperature}_{width}")
        os.makedirs(out_dir, exist_ok=True)
        block = result.line(l)
        for f, j in last.items():
            with open(os.path.join(out_dir, f), "w") as fh:
                fh.write("".join("%.18g\n" % v for v in block[j]))
            n += 1
    return n


def analyze(project_dir: str, benchmark: str, simulation_time, critical_stress: Optional[float] = None,
            sample_lines: Optional[int] = None, technology: str = "CuDD", temperature: float = 378.0, width: float = 0.1,
            write_outputs: bool = False, ctx: Optional[Context] = None, seed: Optional[int] = None) -> AnalyzeReport:
    """`simulation_time` may be one time (the reference's form) or a sequence of timestamps (seconds)."""
    t0 = time.time()
    ts = np.atleast_1d(np.asarray(simulation_time, dtype=np.float64))
    csvs = find_line_csvs(os.path.join(project_dir, benchmark))
    if sample_lines is not None and sample_lines < len(csvs):
        rnd = random.Random(seed)
        csvs = rnd.sample(csvs, sample_lines)          # PROTON_cli.py:1146-1149
    batch, discarded = formulate_for_analyze(csvs, technology, float(temperature), float(width))
    names = list(batch.names or [])
    own = ctx is None
    ctx = ctx or Context(0)
    try:
        res = ctx.analyze_host(batch, ts, critical=critical_stress, want_stress=write_outputs)
    finally:
        if own:
            ctx.close()
    if write_outputs:
        write_reference_outputs(project_dir, names, res, ts, technology, float(temperature), float(width))
    problematic = {}
    if critical_stress is not None:
        for l, name in enumerate(names):
            if res.flag[l, -1]:
                problematic[name] = float(res.max_stress[l, -1])
    return AnalyzeReport(names, discarded, res.max_stress, res.argmax, res.flag, problematic, time.time() - t0, ts)

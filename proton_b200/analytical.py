"""In-process stand-in for PROTON's kernel launcher class.

PROTON's GUI and CLI construct `Analytical_Class(sim_time, directory, selected_line, TECHNOLOGY, TEMPERATURE, WIDTH,
IS_LINUX, INSTALLATION_FOLDER)` and call `analytical_function()`, which turns the formulation's `analytical.txt` into an
`analytical.cfg`, starts `bin/EMtool_analytical` as a child process and maps its exit status to a message
(analytical_class_cli.py:6-91).  Callers only look for the word "seconds" in the returned text (PROTON_cli.py:680,
GUI_auxiliary.py:1438).  This class keeps that surface but runs the kernel through `proton_em_run_cfg` in the calling
process: no shell, no `. /opt/intel/oneapi/.../vars.sh` prefix (which aborts dash on boxes without oneAPI),
one CUDA context for any number of lines.
"""
from __future__ import annotations

import re
import time
from pathlib import Path
from typing import Optional, Tuple

from .api import Context

_KEYVAL = re.compile(r"^\s*(nx_total|Acoef)\s*=\s*(\S+)", re.MULTILINE)


def _read_formulation_summary(path: Path) -> Tuple[str, str]:
    """(`nx_total`, `Acoef`) as the strings `matrix_formulation` wrote (matrix_formulation_func_cli.py:283-286)."""
    found = dict(_KEYVAL.findall(path.read_text()))
    return found.get("nx_total", ""), found.get("Acoef", "")


class Analytical_Class:
    def __init__(self, sim_time, directory, selected_line, TECHNOLOGY, TEMPERATURE, WIDTH, IS_LINUX=True,
                 INSTALLATION_FOLDER="", ctx: Optional[Context] = None):
        self.sim_time, self.directory, self.selected_line = sim_time, directory, selected_line
        self.TECHNOLOGY, self.TEMPERATURE, self.WIDTH = TECHNOLOGY, TEMPERATURE, WIDTH
        self.IS_LINUX, self.INSTALLATION_FOLDER = IS_LINUX, INSTALLATION_FOLDER
        self._ctx = ctx

    # -- helpers ---------------------------------------------------------------------------------------------
    def _config_tag(self) -> str:
        return f"{self.TECHNOLOGY}_{self.TEMPERATURE}_{self.WIDTH}"

    def _input_dir(self) -> Path:
        return Path(self.directory) / "input" / str(self.selected_line) / self._config_tag()

    def _write_cfg(self, nodes: str, acoef: str) -> Path:
        folder = f"{self.directory}/input/{self.selected_line}/{self._config_tag()}/"
        cfg = Path(folder) / "analytical.cfg"
        cfg.write_text("".join(f"{k} {v}\n" for k, v in (("input_files", folder), ("num_nodes", nodes),
                                                          ("A_coeff", acoef), ("simulation_time", self.sim_time))))
        return cfg

    # -- the reference's entry point ---------------------------------------------------------------------------
    def analytical_function(self) -> str:
        if not self.selected_line:
            return "No line was selected and discretized prior to simulation."
        folder = self._input_dir()
        try:
            nodes, acoef = _read_formulation_summary(folder / "analytical.txt")
        except FileNotFoundError:
            return (f"No discretization has been performed for line {self.selected_line} with params: "
                    f"{self.TECHNOLOGY}, {self.TEMPERATURE}, {self.WIDTH}.")
        if not folder.is_dir():
            return f"Input files directory for {self.selected_line} does not exist."
        cfg = self._write_cfg(nodes, acoef)
        started = time.time()
        ctx = self._ctx if self._ctx is not None else Context(0)
        try:
            status, kernel_text = ctx.run_cfg(str(cfg))
        finally:
            if self._ctx is None:
                ctx.close()
        took = time.time() - started
        if status == 0:
            return (f"EM stress analysis was successfully performed in {took:.3f} seconds for the line "
                    f"{self.selected_line} (Configuration: {self.TECHNOLOGY}, {self.TEMPERATURE}K, {self.WIDTH}um).")
        return f"Line analysis terminated with unknown error {status}. {kernel_text.strip()}"

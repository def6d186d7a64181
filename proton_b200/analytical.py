"""`Analytical_Class`: the in-process twin of the reference launcher class of the same name
(analytical_class_cli.py:6-91).  Same constructor arguments, same `analytical_function()` contract: it reads the
`analytical.txt` the formulation wrote, writes `analytical.cfg` with the four keys the kernel reads, runs the kernel —
here `proton_em_run_cfg` inside this process instead of `os.system(bin/EMtool_analytical ...)` — and returns the same
message texts (callers only test for the substring "seconds", PROTON_cli.py:680)."""
from __future__ import annotations

import os
import time

from .api import Context


class Analytical_Class():

  def __init__(self, sim_time, directory, selected_line, TECHNOLOGY, TEMPERATURE, WIDTH, IS_LINUX=True,
               INSTALLATION_FOLDER="", ctx=None):
    self.sim_time = sim_time
    self.directory = directory
    self.selected_line = selected_line
    self.TECHNOLOGY = TECHNOLOGY
    self.TEMPERATURE = TEMPERATURE
    self.WIDTH = WIDTH
    self.IS_LINUX = IS_LINUX
    self.INSTALLATION_FOLDER = INSTALLATION_FOLDER
    self._ctx = ctx

  def analytical_function(self):
    Acoef = ""
    nxTotal = ""
    if not self.selected_line:
      return('No line was selected and discretized prior to simulation.')
    line_path = self.selected_line+"/"+self.TECHNOLOGY+"_"+str(self.TEMPERATURE)+"_"+str(self.WIDTH)+"/"
    analytical_file = self.directory+"/"+"input"+"/"+line_path+"analytical.txt"
    try:
      with open(analytical_file) as f_anal:
        for line in f_anal.readlines():
          words = line.split()
          if 'Acoef' in words: Acoef = words[2]
          if 'nx_total' in words: nxTotal = words[2]
    except FileNotFoundError:
      return(f"No discretization has been performed for line {self.selected_line} with params: {self.TECHNOLOGY}, {self.TEMPERATURE}, {self.WIDTH}.")
    input_files = self.directory + "/input/"+line_path
    if not os.path.exists(input_files):
      return(f'Input files directory for {self.selected_line} does not exist.')
    config_file = self.directory+"/input/"+line_path+"analytical.cfg"
    with open(config_file, "w") as f:
      f.write(f"input_files {input_files}\nnum_nodes {nxTotal}\n")
      f.write(f"A_coeff {Acoef}\nsimulation_time {self.sim_time}\n")
    start_time = time.time()
    own = self._ctx is None
    ctx = self._ctx or Context(0)
    try:
      return_value, text = ctx.run_cfg(config_file)
    finally:
      if own:
        ctx.close()
    elapsed_time = time.time() - start_time
    if return_value == 0:
      return(f"EM stress analysis was successfully performed in {elapsed_time:.3f} seconds for the line {self.selected_line} (Configuration: {self.TECHNOLOGY}, {self.TEMPERATURE}K, {self.WIDTH}um).")
    return(f"Line analysis terminated with unknown error {return_value}. {text.strip()}")

import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch, bench
from proton_b200.api import Context
b, T = bench.build_global("ibmpg2_shaped")
dev = torch.device("cuda:0")
ctx = Context(0)
tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
d_in = [tt(b.num_nodes), tt(b.a_coeff), tt(b.trip_off), tt(b.trip_row), tt(b.trip_col), tt(b.trip_val), tt(b.u_off), tt(b.curden)]
for i in range(3):
    print("---- pack", i, flush=True)
    plan = ctx.pack_device(*d_in)
    plan.close()

"""tools/profile_step.py — per-length-class timing of the evaluation step (CUDA events), and a short run for ncu.

usage: python tools/profile_step.py [--workload ibmpg2_shaped] [--classes] [--steps 2]
"""
import argparse
import sys
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))


def class_of(n):
    if n <= 32:
        return "small"
    if n & (n - 1) == 0:
        return f"pow2_{n}" if n <= 8192 else f"Lpow2_{n}"
    p = 1
    while p < 2 * n - 1:
        p *= 2
    return f"blue_{p}" if n <= 4096 else f"Lblue_{p}"


def main():
    import torch
    import bench
    from proton_b200.api import Context
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="ibmpg2_shaped")
    ap.add_argument("--classes", action="store_true")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--uniform", type=int, nargs="*", default=None, help="line lengths; one uniform batch per value")
    ap.add_argument("--lines", type=int, default=0, help="lines per uniform batch (0: ~3e6 nodes)")
    ap.add_argument("--T", type=int, default=100)
    args = ap.parse_args()
    if args.uniform:
        from proton_b200.workloads import synthetic_ragged
        batch, desc = None, {"T": args.T}
    else:
        if args.workload == "synthetic_ragged20k":
            batch, T0 = bench.build_global("synthetic_ragged", 20000)
        else:
            batch, T0 = bench.build_global(args.workload)
        desc = {"T": T0}
    T = desc["T"]
    ts = bench.timestamps(T)
    dev = torch.device("cuda:0")
    ctx = Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)

    def run(b, label):
        d_in = [tt(b.num_nodes), tt(b.a_coeff), tt(b.trip_off), tt(b.trip_row), tt(b.trip_col), tt(b.trip_val),
                tt(b.u_off), tt(b.curden)]
        plan = ctx.pack_device(*d_in)
        L = b.num_lines
        st = torch.empty(b.total_nodes * T, dtype=torch.float64, device=dev)
        mx = torch.empty(L * T, dtype=torch.float64, device=dev)
        amx = torch.empty(L * T, dtype=torch.int32, device=dev)
        fl = torch.empty(L * T, dtype=torch.uint8, device=dev)
        d_ts = tt(ts)
        plan.eval_device(d_ts, st, mx, amx, fl, critical=40e6)
        ctx.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = ctx.launch_count
        e0.record(stream)
        for _ in range(args.steps):
            plan.eval_device(d_ts, st, mx, amx, fl, critical=40e6)
        e1.record(stream)
        ctx.synchronize()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        ev = b.total_nodes * T
        print(f"{label:>14s}: lines {L:6d} nodes {b.total_nodes:9d}  {ms:9.3f} ms  {ev / ms / 1e6:9.2f} Gevals/s  "
              f"launches/step {(ctx.launch_count - l0) // args.steps}", flush=True)
        plan.close()
        del st

    if args.uniform:
        for n in args.uniform:
            L = args.lines or max(8, 3_000_000 // n)
            run(synthetic_ragged(0, seed=1, sizes=[n] * L), f"N={n}")
        return
    run(batch, "all")
    if args.classes:
        cls = np.array([class_of(int(n)) for n in batch.num_nodes])
        for c in sorted(set(cls), key=lambda s: (s.split("_")[0], int(s.split("_")[1]) if "_" in s else 0)):
            run(batch.subset(np.nonzero(cls == c)[0]), c)


if __name__ == "__main__":
    main()

"""tools/kernel_times.py — per-kernel in-situ durations (proton_em_profile) of one uniform batch or a bench workload.
usage: python tools/kernel_times.py [--uniform N] [--lines L] [--T 100] [--workload ibmpg2_shaped]"""
import argparse
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))


def main():
    import torch
    import bench
    from proton_b200.api import Context
    from proton_b200.workloads import synthetic_ragged
    ap = argparse.ArgumentParser()
    ap.add_argument("--uniform", type=int, default=0)
    ap.add_argument("--lines", type=int, default=64)
    ap.add_argument("--T", type=int, default=100)
    ap.add_argument("--workload", default="ibmpg2_shaped")
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    if args.uniform:
        b, T = synthetic_ragged(0, seed=1, sizes=[args.uniform] * args.lines), args.T
    else:
        b, T = bench.build_global(args.workload, 20000 if args.workload == "synthetic_ragged" else 0)
    ctx = Context(0)
    dev = torch.device("cuda:0")
    sh = bench.Shard(ctx, torch, dev, b, T)
    sh.pack()
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
    for _ in range(2):
        sh.eval()
    ms = bench.time_steps(torch, stream, sh.eval, args.steps, barrier)
    ctx.profile(True)
    for _ in range(args.steps):
        sh.eval()
    ctx.profile(False)
    print(f"step {ms:.3f} ms, {b.total_nodes * T / ms / 1e6:.2f} Gevals/s")
    for k, (tot, n) in ctx.profile_read().items():
        if n:
            print(f"  {k:34s} {tot / args.steps:8.3f} ms/step  {n // args.steps:5d} launches  {1e3 * tot / n:8.2f} us avg")


if __name__ == "__main__":
    main()

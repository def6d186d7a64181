"""tools/one_step.py — exactly one evaluation step of a bench workload inside cudaProfilerStart/Stop, for
    ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,\
smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,\
smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__inst_executed_pipe_fp64.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python tools/one_step.py --workload ibmpg2_shaped
(the launch list tools/launch_summary.py turns into profiles/r02_fp64_counters_<workload>.json)."""
import argparse
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))


def main():
    import torch
    import bench
    from proton_b200.api import Context
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="ibmpg2_shaped")
    ap.add_argument("--lines", type=int, default=0)
    args = ap.parse_args()
    b, T = bench.build_global(args.workload, args.lines)
    ctx = Context(0)
    sh = bench.Shard(ctx, torch, torch.device("cuda:0"), b, T)
    sh.pack()
    sh.eval()
    ctx.synchronize()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    sh.eval()
    ctx.synchronize()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print(f"{args.workload}: {b.num_lines} lines, {b.total_nodes} nodes, T = {T}, evals per step {b.total_nodes * T}")


if __name__ == "__main__":
    main()

"""tools/launch_summary.py <launches.csv> [--last N | --range A B] [--json out.json] — per-kernel totals of an ncu launch list.

The list comes from
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python tools/profile_step.py --steps 1
(per-launch times are serialised and, with ncu's default cache control, cold-cache: the SHARES are meaningful, the
absolute sum is not the step time).  `--last N` keeps only the final N launches (= the last evaluation step).
When the list also carries the FP64 instruction counters
    smsp__sass_thread_inst_executed_op_{dadd,dmul,dfma}_pred_on.sum, sm__inst_executed_pipe_fp64.sum
they are summed per kernel too (executed FP64 thread-instructions; bench.py's roofline.fp64_executed reads the
--json output committed under profiles/).
"""
import csv
import json
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else 0
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rows = list(csv.DictReader(lines))
    per_id = defaultdict(dict)
    order = []
    for r in rows:
        if not r.get("ID", "").isdigit():
            continue
        i = int(r["ID"])
        if i not in per_id:
            order.append(i)
        per_id[i]["name"] = r["Kernel Name"]
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        unit = r["Metric Unit"]
        m = r["Metric Name"]
        if m == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)      # -> us
        elif m.startswith("dram__bytes"):
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        elif unit in ("Kinst", "Minst", "Ginst"):
            v *= {"Kinst": 1e3, "Minst": 1e6, "Ginst": 1e9}[unit]
        per_id[i][m] = v
    if last:
        order = order[-last:]
    if "--range" in sys.argv:                      # launches [A, B) in launch order (e.g. one evaluation step)
        a0 = int(sys.argv[sys.argv.index("--range") + 1])
        b0 = int(sys.argv[sys.argv.index("--range") + 2])
        order = order[a0:b0]
    FP = ("smsp__sass_thread_inst_executed_op_dadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
          "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "sm__inst_executed_pipe_fp64.sum")
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0])
    have_fp = False
    for i in order:
        d = per_id[i]
        name = d["name"].split("(")[0]
        a = agg[name]
        a[0] += 1
        a[1] += d.get("gpu__time_duration.sum", 0.0)
        a[2] += d.get("dram__bytes_read.sum", 0.0)
        a[3] += d.get("dram__bytes_write.sum", 0.0)
        for q, m in enumerate(FP):
            if m in d:
                have_fp = True
                a[4 + q] += d[m]
    tot = sum(a[1] for a in agg.values())
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        line = (f"{name[:52]:52s} n={a[0]:5d} total={a[1] / 1e3:8.3f} ms share={100 * a[1] / tot:5.1f}% avg={a[1] / a[0]:8.1f} us"
                f"  dram rd={a[2] / 1e9:7.3f} GB wr={a[3] / 1e9:7.3f} GB")
        if have_fp:
            line += f"  fp64 thread-inst dadd={a[4]:.4g} dmul={a[5]:.4g} dfma={a[6]:.4g} warp-inst(pipe)={a[7]:.4g}"
        print(line)
    print(f"sum: {tot / 1e3:.3f} ms over {len(order)} launches; dram read {sum(a[2] for a in agg.values()) / 1e9:.3f} GB"
          f" + write {sum(a[3] for a in agg.values()) / 1e9:.3f} GB")
    if have_fp:
        th = sum(a[4] + a[5] + a[6] for a in agg.values())
        print(f"executed FP64 thread-instructions (dadd + dmul + dfma, predicated on): {th:.6g}; "
              f"FP64-pipe warp-instructions: {sum(a[7] for a in agg.values()):.6g}")
    if "--json" in sys.argv:
        out = {"launches": len(order), "sum_ms": tot / 1e3,
               "kernels": {name: {"launches": a[0], "total_us": a[1], "dram_read": a[2], "dram_write": a[3],
                                  "dadd": a[4], "dmul": a[5], "dfma": a[6], "pipe_fp64_warp_inst": a[7]}
                           for name, a in agg.items()}}
        if have_fp:
            out["fp64_thread_inst"] = sum(a[4] + a[5] + a[6] for a in agg.values())
            out["fp64_pipe_warp_inst"] = sum(a[7] for a in agg.values())
        with open(sys.argv[sys.argv.index("--json") + 1], "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()

"""tools/launch_summary.py <launches.csv> [--last N] — per-kernel totals of an ncu launch list.

The list comes from
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python tools/profile_step.py --steps 1
(per-launch times are serialised and, with ncu's default cache control, cold-cache: the SHARES are meaningful, the
absolute sum is not the step time).  `--last N` keeps only the final N launches (= the last evaluation step).
"""
import csv
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
        per_id[i][m] = v
    if last:
        order = order[-last:]
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for i in order:
        d = per_id[i]
        name = d["name"].split("(")[0]
        a = agg[name]
        a[0] += 1
        a[1] += d.get("gpu__time_duration.sum", 0.0)
        a[2] += d.get("dram__bytes_read.sum", 0.0)
        a[3] += d.get("dram__bytes_write.sum", 0.0)
    tot = sum(a[1] for a in agg.values())
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name[:52]:52s} n={a[0]:5d} total={a[1] / 1e3:8.3f} ms share={100 * a[1] / tot:5.1f}% avg={a[1] / a[0]:8.1f} us"
              f"  dram rd={a[2] / 1e9:7.3f} GB wr={a[3] / 1e9:7.3f} GB")
    print(f"sum: {tot / 1e3:.3f} ms over {len(order)} launches; dram read {sum(a[2] for a in agg.values()) / 1e9:.3f} GB"
          f" + write {sum(a[3] for a in agg.values()) / 1e9:.3f} GB")


if __name__ == "__main__":
    main()

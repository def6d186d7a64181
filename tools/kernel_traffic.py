"""tools/kernel_traffic.py <ncu_long_summary.txt> <ncu_incta_summary.txt> > profiles/r02_kernel_traffic.json

DRAM bytes per launch of the dominant kernels, from the two `ncu --set full` summaries (tools/ncu_summary.py output):
the long-path capture holds one launch per kernel, the in-CTA capture every launch of one synthetic-ragged step (summed
per kernel family).  bench.py's roofline.traffic reads the result."""
import json
import re
import sys


def blocks(path):
    txt = open(path).read()
    for blk in txt.split("==== ")[1:]:
        name = blk.splitlines()[0].strip()
        def g(k):
            m = re.search(re.escape(k) + r"\s+([\d.]+)\s*(\S*)", blk)
            if not m:
                return 0.0
            v = float(m.group(1))
            return v * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1.0, "ms": 1e3}.get(m.group(2), 1.0)
        yield name, g("gpu__time_duration.sum"), g("dram__bytes_read.sum"), g("dram__bytes_write.sum")


def family(n):
    n = n.replace("void ", "")
    if n.startswith("em_eval_fft_kernel"):
        return "em_eval_fft_kernel<bluestein>" if re.search(r",\s*1>", n) else "em_eval_fft_kernel<pow2>"
    return n.split("<")[0].split("(")[0]


def main():
    out = {}
    for name, us, rd, wr in blocks(sys.argv[1]):
        out[family(name)] = {"capture": "profiles/r02_ncu_long_full.txt (--cache-control none: the L2-resident scratch stays in L2 as in production)",
                             "duration_us": us, "dram_read": rd, "dram_write": wr, "dram_bytes_per_launch": rd + wr}
    cap = ("profiles/r02_ncu_incta_full.txt (synthetic ragged 20 000 lines, T = 256: totals over the launches of one step; "
           "algorithmic: 8 B x 14.7e6 nodes x 256 timestamps = 30.2 GB of stress per step over all families)")
    for name, us, rd, wr in blocks(sys.argv[2]):
        f = family(name)
        d = out.setdefault(f, {"launches": 0, "dram_read": 0.0, "dram_write": 0.0, "capture": cap})
        d["launches"] += 1
        d["dram_read"] += rd
        d["dram_write"] += wr
        d["dram_bytes_per_launch"] = (d["dram_read"] + d["dram_write"]) / d["launches"]
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()

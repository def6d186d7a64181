#!/usr/bin/env python
"""bench.py — node·timestamp stress evaluations per second of the line-analysis hot path on N B200s.

One "step" = one pass of the hot path (modal evaluation + inverse DCT + stress write-back + per-line max / argmax / flag)
over the whole batch of lines at all T timestamps.  `value` times the step with the packed batch resident in HBM (CUDA
events on the library's stream); `value_with_pack` adds the packing and forward-DCT kernels; `e2e` times the public
host-buffer call (pack + upload + eval + download into pinned host memory).  `--impl reference` times the reference's
own CPU implementation on the host cores.

Workloads (`--workload`, BASELINE.json configs):
  ibmpg2_shaped (default, config 2)  benchmarks/ibmpg2.spice is a missing blob, so it is ibmpg1's real line population
                                     (reference parser) replicated 7x with seeded jitter, T = 100 log-spaced timestamps
  ibmpg1 (config 1)                  the real ibmpg1 line population, T = 100
  synthetic_ragged (config 3)        1 000 000 lines of 16-4096 nodes (log-uniform), T = 256; the 1.5 TB stress field is
                                     written through a reused device ring buffer (--lines scales it down)
  config4 (config 4)                 10 M nodes with the ibmpg length mix, T = 100
  long_lines (config 5)              8 lines of 65 536 nodes, T = 1000

Multi-GPU runs are STRONG scaling: the same global batch is split over the ranks by the segment-balanced LPT split
(workloads.partition_lines), no data-path collective; per step the per-line reductions {max, argmax, flag} are
all-gathered (NCCL) and, after the timed region, rank 0 checks lines owned by other ranks against the CPU oracle.  With
N > 1 the default workload also carries a `config4` sub-record (10 M nodes split N ways: per-rank time, imbalance of
time, speed-up over one GPU, gather time).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO))

METRIC = "segment-timestamp stress evals/sec"
UNIT = "evals/s"
CRITICAL = 40e6
RING_BYTES = 8 << 30          # device ring buffer for fields that do not fit (config 3)
FIELD_LIMIT = 64 << 30        # larger per-GPU fields go through the ring


def timestamps(T):
    return np.logspace(6.0, np.log10(6.38e8), T)


# ------------------------------------------------------------------------------------------------------------------
# workloads
# ------------------------------------------------------------------------------------------------------------------
def build_global(name: str, lines_override: int = 0):
    """The whole batch of a workload (identical on every rank) -> (LineBatch, T)."""
    from proton_b200.api import LineBatch
    from proton_b200 import workloads as W
    if name == "ibmpg2_shaped":
        return W.ibmpg2_shaped_batch(replicas=7, seed=2), 100
    if name == "ibmpg1":
        return W.ibmpg1_batch(), 100
    if name == "synthetic_ragged":
        return W.synthetic_ragged(lines_override or 1_000_000, seed=0), 256
    if name == "config4":
        return W.config4_batch(), 100
    if name == "long_lines":
        return W.synthetic_ragged(0, seed=0, sizes=[65536] * (lines_override or 8)), 1000
    raise SystemExit(f"unknown workload {name}")


def config_dict(name, batch, T, world, parts):
    """`config` of the JSON line: a function of the workload and N only, so both arms print the same dict."""
    nn = batch.num_nodes.astype(np.int64)
    loads = [int(nn[p].sum()) for p in parts]
    return {"workload": name, "lines": int(batch.num_lines), "nodes": int(nn.sum()), "timestamps": int(T),
            "max_nodes_per_line": int(nn.max()),
            "l2": f"the stress written per step is {nn.sum() * T * 8 / world / 2**30:.2f} GiB per GPU, far larger than the "
                  "126 MB L2 (no flush needed between steps)",
            "parallelism": f"one fixed batch, lines sharded over {world} GPU(s) by the LPT segment-balanced split",
            "nodes_per_gpu_max_over_mean": float(max(loads) / (sum(loads) / len(loads)))}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])), mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def bind_to_gpu_numa(index: int):
    """Run this process on the CPUs of the GPU's NUMA node, so that pinned host buffers are allocated next to it
    (the e2e leg is bound by host<->device copies).  Returns a description, or None if the box exposes no NUMA node."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(index)
        dev = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(Path(f"/sys/bus/pci/devices/{dev}/numa_node").read_text().strip())
        if node < 0:
            return None
        cpus = set()
        for part in Path(f"/sys/devices/system/node/node{node}/cpulist").read_text().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return f"numa node {node} ({len(cpus)} cpus)"
    except Exception:
        return None


def measured_peaks():
    p = REPO / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json, STREAM-style copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def fp64_peak():
    """DFMA peak measured on this pool's B200 by tools/fp64_peak.cu (profiles/r01_fp64_peak.json): thread-level DFMA
    instructions per second, and the same as TFLOP/s (2 flops each)."""
    p = REPO / "profiles" / "r01_fp64_peak.json"
    try:
        rows = [json.loads(ln) for ln in p.read_text().splitlines() if ln.strip().startswith("{")]
        best = max(r["dfma_per_s"] for r in rows if "dfma_per_s" in r)
        return best, 2.0 * best / 1e12, "measured (profiles/r01_fp64_peak.json, tools/fp64_peak.cu)"
    except Exception:
        return 1.8e13, 36.0, "fallback (B200: 64 FP64 lanes x 148 SMs x 1.9 GHz)"


# evals per step of the runs the committed counter files were captured on (tools/one_step.py)
PROFILE_EVALS = {"ibmpg2_shaped": 7334331 * 100, "synthetic_ragged": 14743384 * 256, "long_lines": 524288 * 1000}


def fp64_counters(workload, evals):
    """Executed FP64 thread-instructions per step and per kernel family, from the committed ncu launch list of one step
    of this workload (tools/one_step.py -> tools/launch_summary.py --json), scaled by evals when the run is a larger
    or smaller instance of the same length mix.  None when no counter file exists for the workload."""
    p = REPO / "profiles" / f"r02_fp64_counters_{workload}.json"
    if not p.exists() or workload not in PROFILE_EVALS:
        return None
    d = json.loads(p.read_text())
    scale = float(evals) / PROFILE_EVALS[workload]
    fam = {}
    for k, v in d.get("kernels", {}).items():
        fam.setdefault(kernel_family(k), 0.0)
        fam[kernel_family(k)] += scale * (v.get("dadd", 0.0) + v.get("dmul", 0.0) + v.get("dfma", 0.0))
    return {"file": f"profiles/{p.name}", "per_step": scale * float(d.get("fp64_thread_inst", 0.0)), "per_family": fam,
            "scaled_by": scale}


def kernel_family(ncu_name: str) -> str:
    """ncu kernel name -> the family name proton_em_profile reports."""
    n = ncu_name.replace("void ", "").replace("em::", "").strip()
    if n.startswith("em_eval_fft_kernel"):
        return "em_eval_fft_kernel<bluestein>" if n.rstrip(">").rstrip().endswith("1") or "true" in n else "em_eval_fft_kernel<pow2>"
    return n.split("<")[0].split("(")[0]


def cpu_baseline_port(batch, ts, budget_s=12.0):
    """The CPU oracle port (oracle/em_oracle.c, double-precision DCT, all host threads) on a bounded sample."""
    from oracle import restatement as R
    R.load()
    threads = R.max_threads()
    order = np.argsort(-batch.num_nodes, kind="stable")
    # probe with a few lines to size the sample
    probe = batch.subset(order[:: max(1, len(order) // 64)][:64])
    t0 = time.perf_counter()
    R.batch(probe.num_nodes, probe.a_coeff, probe.trip_off, probe.trip_row, probe.trip_col, probe.trip_val, probe.u_off,
            probe.curden, ts[:4], want_stress=False, use_double=True)
    dt = max(time.perf_counter() - t0, 1e-4)
    rate = probe.total_nodes * 4 / dt
    want = rate * budget_s
    Tn = len(ts)
    stride = max(1, int(np.ceil(batch.total_nodes * Tn / want)))
    for _ in range(2):
        sample = batch.subset(np.sort(order[::stride]))
        t0 = time.perf_counter()
        R.batch(sample.num_nodes, sample.a_coeff, sample.trip_off, sample.trip_row, sample.trip_col, sample.trip_val,
                sample.u_off, sample.curden, ts, want_stress=False, use_double=True)
        dt = time.perf_counter() - t0
        if stride == 1 or dt > budget_s / 3:
            break
        stride = max(1, int(stride * dt / budget_s * 1.2))      # the probe under-estimated the rate: enlarge once
    return {"value": sample.total_nodes * Tn / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"every {stride}-th line by length ({sample.num_lines} lines, {sample.total_nodes} nodes) x {Tn} "
                      f"timestamps, {dt:.1f} s; oracle/em_oracle.c with the double-precision O(N log N) DCT, pthreads"}


def run_reference(args, world, rank):
    """--impl reference: the UNMODIFIED reference binary (oracle/_ref/EMtool_analytical under the MKL/Boost shim
    libraries, double-precision shim FFT), one process per (line, timestamp) from a thread pool exactly like
    PROTON_cli.analyze (PROTON_cli.py:1174-1181); falls back to the oracle port if oracle/_ref is absent."""
    if rank != 0:
        return
    from oracle import restatement as R
    from proton_b200 import workloads as W
    import tempfile
    batch, T = build_global(args.workload, args.lines)
    ts = timestamps(T)
    cores = os.cpu_count() or 1
    config = config_dict(args.workload, batch, T, world, W.partition_lines(batch.num_nodes, world))
    if R.ref_binary_available():
        # bounded sample: the binary is O(N^2) in time and memory, so lines above 8000 nodes are left out
        #  and a stride keeps each step to tens of seconds
        ok = np.nonzero(batch.num_nodes <= 8000)[0]
        order = ok[np.argsort(-batch.num_nodes[ok], kind="stable")]
        sel = np.sort(order[:: max(1, len(order) // 600)])
        sub = batch.subset(sel)
        t_one = float(ts[-1])
        tmp = tempfile.mkdtemp(prefix="proton_ref_")
        cfgs = []
        for l in range(sub.num_lines):
            a, b = sub.trip_off[l], sub.trip_off[l + 1]
            c, d = sub.u_off[l], sub.u_off[l + 1]
            cfgs.append(R.write_line_inputs(Path(tmp) / "input" / f"line{l}" / "cfg", int(sub.num_nodes[l]),
                                            float(sub.a_coeff[l]), sub.trip_row[a:b], sub.trip_col[a:b],
                                            sub.trip_val[a:b], sub.curden[c:d], t_one))

        def one(cfg):
            res = R.run_binary(cfg, fft="double")
            if res.returncode != 0:
                raise RuntimeError(res.stdout + res.stderr)

        def step():
            with ThreadPoolExecutor(max_workers=cores) as ex:
                list(ex.map(one, cfgs))

        for _ in range(args.warmup):
            step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step()
        dt = (time.perf_counter() - t0) / args.steps
        evals = sub.total_nodes
        kind = "reference"
        sample = (f"{sub.num_lines} lines (every {max(1, len(order) // 600)}-th by length among N <= 8000; longer lines need "
                  f">0.5 GB each in the reference), {sub.total_nodes} nodes, 1 timestamp per process, {cores} worker threads; "
                  "bin/EMtool_analytical + shim FFT (MKL not on this box)")
    else:
        cb = cpu_baseline_port(batch, ts)
        dt, evals, kind, sample, cores = 1.0, cb["value"], "port", cb["sample"], cb["cores"]
    value = evals / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------------------------
# one rank's share of a batch, resident on its GPU
# ------------------------------------------------------------------------------------------------------------------
class Shard:
    """The lines of one rank, packed on its GPU, with device outputs.  Fields that fit are written whole; larger ones go
    through a reused ring buffer: the shard is cut into contiguous sub-batches whose fields fit the ring, each packed
    into its own plan, and every sub-batch writes its stress to the same device buffer."""

    def __init__(self, ctx, torch, dev, batch, T):
        self.ctx, self.torch, self.dev, self.batch, self.T = ctx, torch, dev, batch, T
        self.L = batch.num_lines
        self.evals = batch.total_nodes * T
        tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        self.d_ts = tt(timestamps(T))
        field = self.evals * 8
        self.ring = field > FIELD_LIMIT
        if self.ring:
            per = max(1, RING_BYTES // (8 * T))                     # nodes per sub-batch
            off = batch.node_off
            cuts = [0]
            while cuts[-1] < self.L:
                l1 = int(np.searchsorted(off, off[cuts[-1]] + per, side="right") - 1)
                cuts.append(min(self.L, max(l1, cuts[-1] + 1)))
            self.parts = [batch.slice(a, b) for a, b in zip(cuts[:-1], cuts[1:])]
        else:
            self.parts = [batch]
        self.line0 = np.cumsum([0] + [p.num_lines for p in self.parts])
        self.d_in = [[tt(x) for x in (p.num_nodes, p.a_coeff, p.trip_off, p.trip_row, p.trip_col, p.trip_val, p.u_off,
                                      p.curden)] for p in self.parts]
        cap = max(p.total_nodes for p in self.parts) * T
        self.d_stress = torch.empty(cap, dtype=torch.float64, device=dev)
        self.d_mx = torch.empty(self.L * T, dtype=torch.float64, device=dev)
        self.d_amx = torch.empty(self.L * T, dtype=torch.int32, device=dev)
        self.d_flag = torch.empty(self.L * T, dtype=torch.uint8, device=dev)
        self.plans = []

    def pack(self):
        for p in self.plans:
            p.close()
        self.plans = [self.ctx.pack_device(*d) for d in self.d_in]

    def eval(self):
        T = self.T
        for i, plan in enumerate(self.plans):
            a, b = int(self.line0[i]) * T, int(self.line0[i + 1]) * T
            plan.eval_device(self.d_ts, self.d_stress, self.d_mx[a:b], self.d_amx[a:b], self.d_flag[a:b], critical=CRITICAL)

    def close(self):
        for p in self.plans:
            p.close()
        self.plans = []


class Gather:
    """The one exchange of the path: every rank contributes max (f64), argmax (i32) and flag (u8) of its lines at all
    timestamps, packed into one byte buffer of 13 bytes per (line, t) (padded to the longest shard), and ONE NCCL
    all-gather moves them.  The packing copies are queued on the library's stream right behind the evaluation; the
    all-gather runs on a side stream from one of two send / receive buffer pairs, so the exchange of step k proceeds
    under the evaluation of step k + 1 (the library's stream waits for it only before it reuses the pair, two steps
    later, and in `finish()`, which the timed loops call before their closing event).  PROTON_EM_BENCH_GATHER=serial
    keeps everything on the library's stream."""

    def __init__(self, torch, dist, dev, stream, Lmax, T, world):
        self.torch, self.dist, self.stream = torch, dist, stream
        self.n, self.world = Lmax * T, world
        self.overlap = os.environ.get("PROTON_EM_BENCH_GATHER", "overlap") != "serial"
        n = self.n
        self.bufs = []
        for _ in range(2 if self.overlap else 1):
            send = torch.zeros(n * 13, dtype=torch.uint8, device=dev)
            recv = torch.empty(world * n * 13, dtype=torch.uint8, device=dev)
            views = (send[: 8 * n].view(torch.float64), send[8 * n: 12 * n].view(torch.int32), send[12 * n:])
            self.bufs.append((send, recv, views, torch.cuda.Event(), torch.cuda.Event()))
        self.side = torch.cuda.Stream(device=dev) if self.overlap else None
        self.k = 0
        self.recv = self.bufs[0][1]

    def run(self, shard):
        k = shard.L * shard.T
        send, recv, views, ready, done = self.bufs[self.k % len(self.bufs)]
        self.k += 1
        self.recv = recv
        with self.torch.cuda.stream(self.stream):
            if self.overlap:
                self.stream.wait_event(done)          # the pair is free again (no-op until it has been used once)
            for view, src in zip(views, (shard.d_mx, shard.d_amx, shard.d_flag)):
                view[:k].copy_(src)
            if not self.overlap:
                self.dist.all_gather_into_tensor(recv, send)
                return
            ready.record(self.stream)
        with self.torch.cuda.stream(self.side):
            self.side.wait_event(ready)
            self.dist.all_gather_into_tensor(recv, send)
            done.record(self.side)
        self.last_done = done

    def finish(self):
        """The library's stream waits for the exchange that is still in flight."""
        if self.overlap and self.k:
            self.stream.wait_event(self.last_done)

    def unpack(self):
        """-> [max (world*n f64), argmax (world*n i32), flag (world*n u8)] as numpy, rank-major."""
        n = self.n
        raw = self.recv.cpu().numpy().reshape(self.world, 13 * n)
        return [np.ascontiguousarray(raw[:, : 8 * n]).view(np.float64).reshape(-1),
                np.ascontiguousarray(raw[:, 8 * n: 12 * n]).view(np.int32).reshape(-1),
                np.ascontiguousarray(raw[:, 12 * n:]).reshape(-1)]


def time_steps(torch, stream, step, n, barrier, finish=None):
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(n):
        step()
    if finish is not None:
        finish()                      # (work the steps left on other streams: the closing event waits for it)
    ev1.record(stream)
    barrier()
    return ev0.elapsed_time(ev1) / n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ibmpg2_shaped")
    ap.add_argument("--lines", type=int, default=0, help="line count of the synthetic workloads (0: the BASELINE size)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-config4", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, world, rank)
        return

    import torch
    import torch.distributed as dist
    from proton_b200.api import Context, EvalResult, LineBatch
    from proton_b200 import workloads as W
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    all_cpus = os.sched_getaffinity(0)
    numa = bind_to_gpu_numa(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL prints its version banner on stdout when the first communicator comes up; stdout carries the one JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            warm = torch.zeros(1, device=dev)
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)

    gbatch, T = build_global(args.workload, args.lines)
    ts = timestamps(T)
    parts = W.partition_lines(gbatch.num_nodes, world)
    batch = gbatch.subset(parts[rank]) if world > 1 else gbatch
    config = config_dict(args.workload, gbatch, T, world, parts)
    evals_total = float(gbatch.total_nodes) * T
    ctx = Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    shard = Shard(ctx, torch, dev, batch, T)
    L = shard.L
    t0 = time.perf_counter()
    shard.pack()
    ctx.synchronize()
    pack_first_ms = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    shard.pack()
    ctx.synchronize()
    pack_ms = (time.perf_counter() - t0) * 1e3

    # ---- the only exchange of the path: all-gather of {max, argmax, flag} per (line, t) -----------------------------
    Lmax = max(len(p) for p in parts)
    if world > 1:
        gth = Gather(torch, dist, dev, stream, Lmax, T, world)

    def gather():
        gth.run(shard)

    def step():
        shard.eval()
        if world > 1:
            gather()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = ctx.launch_count
    ms_rank = time_steps(torch, stream, step, args.steps, barrier, gth.finish if world > 1 else None)
    launches = ctx.launch_count - launches0
    clocks = sampler.stop()
    ms = allmax(ms_rank)
    value = evals_total / (ms * 1e-3)

    # ---- per-rank figures of the split (N > 1): evaluation alone, the gather alone --------------------------------
    multi = None
    if world > 1:
        n_sub = max(3, min(args.steps, 10))
        ms_eval = time_steps(torch, stream, shard.eval, n_sub, barrier)
        ms_gather = time_steps(torch, stream, gather, n_sub, barrier, gth.finish)
        stats = torch.tensor([ms_eval, ms_gather, float(batch.total_nodes), float(L)], device=dev, dtype=torch.float64)
        allstats = torch.empty(world * 4, device=dev, dtype=torch.float64)
        dist.all_gather_into_tensor(allstats, stats)
        a = allstats.cpu().numpy().reshape(world, 4)
        multi = {"eval_ms_per_rank": [round(float(x), 4) for x in a[:, 0]],
                 "eval_ms_min": float(a[:, 0].min()), "eval_ms_max": float(a[:, 0].max()),
                 "imbalance_of_time_max_over_mean": float(a[:, 0].max() / a[:, 0].mean()),
                 "gather_ms": float(a[:, 1].max()), "gather_bytes_per_rank": int(Lmax * T * 13 * world),
                 "gather_mode": ("side stream, under the next step's evaluation (two buffer pairs)" if gth.overlap
                                 else "on the library's stream, behind the evaluation"),
                 "nodes_per_rank": [int(x) for x in a[:, 2]], "lines_per_rank": [int(x) for x in a[:, 3]]}
        # rank 0 checks the gathered reductions of lines owned by OTHER ranks against the CPU oracle
        barrier()
        if rank == 0:
            multi["gather_check"] = check_gathered(gbatch, parts, ts, gth.unpack(), Lmax, T)

    # ---- roofline (this rank's share) ----------------------------------------------------------------------------
    roofline = build_roofline(args, ctx, shard, batch, T, ms_rank, barrier, W)

    # ---- pack + eval, device resident, cached tables --------------------------------------------------------------
    def pack_and_eval():
        shard.pack()
        shard.eval()
    pack_and_eval()
    barrier()
    n_pe = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(n_pe):
        pack_and_eval()
    ctx.synchronize()
    ms_with_pack = allmax((time.perf_counter() - t0) * 1e3 / n_pe)

    # ---- e2e: public host-buffer call, pinned host memory, copies inside the timed region ---------------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, ctx, torch, dist, dev, batch, T, ts, evals_total, world, barrier, allmax, shard.ring)

    # ---- config 4 sub-record (N > 1, default workload) ----------------------------------------------------------
    config4 = None
    if world > 1 and args.workload == "ibmpg2_shaped" and not args.no_config4:
        shard.close()
        del shard
        torch.cuda.empty_cache()
        config4 = run_config4(args, ctx, torch, dist, dev, stream, W, world, rank, barrier, allmax)
        shard = None

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)          # the CPU baseline gets every host core
        cpu_baseline = cpu_baseline_port(batch, ts)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong" if world > 1 else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "clocks": clocks, "gpu_launches": int(launches), "roofline": roofline, "e2e": e2e,
                "cpu_baseline": cpu_baseline,
                "value_with_pack": {"value": evals_total / (ms_with_pack * 1e-3), "unit": UNIT, "ms_per_step": ms_with_pack,
                                    "note": "proton_em_pack (packing + forward-DCT kernels, per-length tables cached in the "
                                            "context) + proton_em_eval per step, batch arrays resident in HBM, wall clock "
                                            "around a synchronised loop"},
                "pack_ms": {"first_call_with_table_build": pack_first_ms, "cached_tables": pack_ms},
                "host_affinity": numa, "field": "ring buffer" if (shard.ring if shard else False) else "whole",
                "lines_per_gpu": int(L), "nodes_per_gpu": int(batch.total_nodes)}
        if multi:
            line["multi_gpu"] = multi
        if config4:
            line["config4"] = config4
        print(json.dumps(line))
    if shard:
        shard.close()
    if world > 1:
        dist.destroy_process_group()


def check_gathered(gbatch, parts, ts, recv, Lmax, T, per_rank=3, tol=1e-9):
    """Rank 0: oracle check of the gathered {max, argmax, flag} on lines owned by other ranks (the shard -> line mapping
    of partition_lines).  Escapes (arg-max / flag differences explained by a near-tie within tol) are counted."""
    from oracle import restatement as R
    R.load()
    g_mx, g_amx, g_flag = [r.reshape(len(parts), Lmax, T) for r in recv]
    rng = np.random.default_rng(0)
    checked = escapes = 0
    worst = 0.0
    for r in range(1, len(parts)):
        own = parts[r]
        cand = np.nonzero(gbatch.num_nodes[own] <= 30000)[0]
        pick = rng.choice(cand, size=min(per_rank, len(cand)), replace=False)
        sub = gbatch.subset(own[pick])
        ws, wm, wa = R.batch(sub.num_nodes, sub.a_coeff, sub.trip_off, sub.trip_row, sub.trip_col, sub.trip_val, sub.u_off,
                             sub.curden, ts, want_stress=True)
        off = sub.node_off
        for i, li in enumerate(pick):
            w = ws[off[i] * T:off[i + 1] * T].reshape(T, -1)
            scale = np.abs(w).max(axis=1)
            err = np.abs(g_mx[r, li] - wm[i]) / np.maximum(scale, 1e-300)
            worst = max(worst, float(err.max()))
            if not np.all(err <= tol):
                return {"ok": False, "why": f"rank {r} local line {li}: max differs by {err.max():.3e}"}
            for j in range(T):
                if g_amx[r, li, j] != wa[i, j]:
                    escapes += 1
                    if abs(w[j, g_amx[r, li, j]] - w[j, wa[i, j]]) > 2 * tol * scale[j]:
                        return {"ok": False, "why": f"rank {r} local line {li} t {j}: argmax {g_amx[r, li, j]} vs {wa[i, j]}"}
                wf = wm[i, j] > CRITICAL
                if bool(g_flag[r, li, j]) != wf:
                    escapes += 1
                    if abs(wm[i, j] - CRITICAL) > 2 * tol * scale[j]:
                        return {"ok": False, "why": f"rank {r} local line {li} t {j}: flag differs"}
            checked += 1
    return {"ok": True, "lines_checked": checked, "timestamps": T, "worst_rel_error_of_max": worst,
            "near_tie_escapes": escapes, "tolerance": tol}


def build_roofline(args, ctx, shard, batch, T, ms_rank, barrier, W):
    """`roofline` of the JSON line.  The path is FP64-pipe bound, so the headline entry is the dominant
    kernel's FP64 figure; the HBM figures the contract asks for are kept beside it."""
    L = batch.num_lines
    evals_rank = batch.total_nodes * T
    peak, peak_src = measured_peaks()
    dfma_peak, fp64_tf, fp64_src = fp64_peak()
    alg_bytes = evals_rank * 8.0 + batch.total_nodes * 16.0 + L * T * 13.0
    t_rank = ms_rank * 1e-3
    step_gbs = alg_bytes / t_rank / 1e9
    nn = batch.num_nodes.astype(np.int64)
    kinds = np.array([W.kernel_class(int(n)) for n in nn])
    ev_kind = {k: float(nn[kinds == k].sum()) * T for k in set(kinds.tolist())}
    fl_line = nn * (2.5 * np.log2(np.maximum(nn, 2)) + 27.0) * T       # algorithmic flops per line
    fl_kind = {k: float(fl_line[kinds == k].sum()) for k in set(kinds.tolist())}
    long_ev = ev_kind.get("long_blue", 0.0) + ev_kind.get("long_pow2", 0.0)
    long_fl = fl_kind.get("long_blue", 0.0) + fl_kind.get("long_pow2", 0.0)
    through = {"em_eval_dense_kernel": ("dense", 1.0), "em_eval_fft_kernel<pow2>": ("fft_pow2", 1.0),
               "em_eval_fft_kernel<bluestein>": ("fft_blue", 1.0), "em_eval_mix_kernel": ("fft_mix", 1.0)}
    evals_through = {k: ev_kind.get(c, 0.0) for k, (c, _) in through.items()}
    flops_through = {k: fl_kind.get(c, 0.0) for k, (c, _) in through.items()}
    for k in ("em_long_colfwd_paired_kernel", "em_long_row_kernel", "em_long_colinv_kernel", "em_long_scatter_kernel",
              "em_long_reduce_kernel", "em_long_fused_kernel"):
        evals_through[k] = long_ev
        flops_through[k] = long_fl
    # Per-kernel live timing: the same step again with a CUDA event pair around every launch, on the stream it is
    # launched on (proton_em_profile).  Kernels of different lanes overlap, so the per-kernel sums exceed the step.
    n_prof = max(1, min(args.steps, 5))
    barrier()
    ctx.profile(True)
    for _ in range(n_prof):
        shard.eval()
    ctx.profile(False)
    prof = {k: v for k, v in ctx.profile_read().items() if v[1] > 0}
    tot_ms = sum(v[0] for v in prof.values()) or 1.0
    counters = fp64_counters(args.workload, evals_rank)
    kernels = {}
    for k, v in prof.items():
        kernels[k] = {"ms_per_step": v[0] / n_prof, "launches_per_step": v[1] / n_prof, "avg_launch_us": 1e3 * v[0] / v[1],
                      "share_of_kernel_time": v[0] / tot_ms}
        if counters and k in counters["per_family"]:
            inst = counters["per_family"][k]
            kernels[k]["fp64_thread_inst_per_step"] = inst
            kernels[k]["fp64_issue_frac"] = inst / (v[0] / n_prof * 1e-3) / dfma_peak
    dom = max(prof, key=lambda k: prof[k][0])
    dom_ms, dom_n = prof[dom][0] / n_prof, prof[dom][1] / n_prof
    dom_bytes = 8.0 * evals_through.get(dom, 0.0)            # per step: 8 B per eval that passes through the kernel
    hbm_achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
    traffic = None
    for tp in (REPO / "profiles" / "r02_kernel_traffic.json", REPO / "profiles" / "r01_kernel_traffic.json"):
        if tp.exists():
            traffic = json.loads(tp.read_text()).get(dom, {}).get("dram_bytes_per_launch")
            if traffic is not None:
                break
    flops = float(fl_line.sum())
    # dominant kernel, FP64: executed thread-instructions (each holds one DFMA issue slot = 2 flops of the peak) from the
    # committed counters when present, else the algorithmic flops of the lines that pass through it
    if counters and dom in counters["per_family"]:
        dom_tf = 2.0 * counters["per_family"][dom] / (dom_ms * 1e-3) / 1e12
        basis = f"executed FP64 thread-instructions per step of this kernel ({counters['file']}) x 2 flops per issue slot"
    else:
        dom_tf = flops_through.get(dom, 0.0) / (dom_ms * 1e-3) / 1e12
        basis = "algorithmic flops T*N*(2.5 log2 N + 27) of the lines that pass through this kernel (no counter file)"
    roofline = {"bound": "fp64", "kernel": dom, "achieved": dom_tf, "peak": fp64_tf, "unit": "TFLOP/s",
                "frac": dom_tf / fp64_tf, "traffic": traffic, "peak_source": fp64_src, "basis": basis,
                "avg_launch_us": 1e3 * dom_ms / dom_n, "launches_per_step": dom_n,
                "share_of_kernel_time": kernels[dom]["share_of_kernel_time"],
                "algorithmic": {"achieved": flops_through.get(dom, 0.0) / (dom_ms * 1e-3) / 1e12, "unit": "TFLOP/s",
                                "frac": flops_through.get(dom, 0.0) / (dom_ms * 1e-3) / 1e12 / fp64_tf,
                                "flops_per_launch": flops_through.get(dom, 0.0) / dom_n,
                                "note": "SURVEY.md §8d flops of the lines passing through the dominant kernel / its summed "
                                        "launch durations"},
                "hbm": {"bound": "hbm", "achieved": hbm_achieved, "peak": peak, "unit": "GB/s", "frac": hbm_achieved / peak,
                        "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes / dom_n,
                        "note": "8 B x evals passing through the dominant kernel per step / its summed launch durations; "
                                "the kernel works out of the L2-resident scratch, HBM is not what binds it"},
                "step": {"hbm": {"achieved": step_gbs, "frac": step_gbs / peak, "unit": "GB/s",
                                 "note": "all launches of one step, 8 + 16/T + 13/N bytes per eval over the step time"},
                         "fp64_algorithmic": {"achieved": flops / t_rank / 1e12, "peak": fp64_tf, "unit": "TFLOP/s",
                                              "frac": flops / t_rank / 1e12 / fp64_tf}},
                "kernels": kernels,
                "note": "dominant kernel by summed live duration (CUDA events on the launching stream, kernels of other "
                        "lanes running concurrently); the kernels are FP64-pipe / shared-memory bound, not HBM bound"}
    if counters:
        per_eval = counters["per_step"] / float(evals_rank) if evals_rank else 0.0
        roofline["fp64_executed"] = {"thread_inst_per_eval": per_eval, "thread_inst_per_step": counters["per_step"],
                                     "frac_of_dfma_issue_peak_over_step": counters["per_step"] / t_rank / dfma_peak,
                                     "dfma_issue_peak_per_s": dfma_peak, "source": counters["file"],
                                     "note": "sum over all launches of one step of smsp__sass_thread_inst_executed_op_"
                                             "{dadd,dmul,dfma}_pred_on (ncu launch list of this command) / evals per "
                                             "step; fraction = that count / step time / measured DFMA issue rate"}
        roofline["step"]["fp64_executed"] = roofline["fp64_executed"]["frac_of_dfma_issue_peak_over_step"]
    return roofline


def d2h_roof(torch, dist, dev, nbytes, world, barrier, reps=3):
    """Bare pinned device-to-host copy of the same size on every rank at once: the ceiling of the full-field e2e leg."""
    n = max(1, nbytes // 8)
    d = torch.empty(n, dtype=torch.float64, device=dev)
    h = torch.empty(n, dtype=torch.float64).pin_memory()
    h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    if world > 1:
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    del d, h
    return n * 8 * world / dt / 1e9, dt


def run_e2e(args, ctx, torch, dist, dev, batch, T, ts, evals_total, world, barrier, allmax, ring):
    from proton_b200.api import EvalResult, LineBatch
    L = batch.num_lines
    evals_rank = batch.total_nodes * T
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    hp = [pin(x) for x in (batch.num_nodes, batch.a_coeff, batch.trip_off, batch.trip_row, batch.trip_col,
                           batch.trip_val, batch.u_off, batch.curden)]
    hb = LineBatch(*[t.numpy() for t in hp])
    h_mx = torch.empty((L, T), dtype=torch.float64).pin_memory()
    h_amx = torch.empty((L, T), dtype=torch.int32).pin_memory()
    h_flag = torch.empty((L, T), dtype=torch.uint8).pin_memory()
    h_ts = pin(ts).numpy()
    n_e2e = max(2, min(args.steps, 5))
    h2d = sum(t.numel() * t.element_size() for t in hp) + T * 8

    def timed(out, want_stress):
        ctx.analyze_host(hb, h_ts, critical=CRITICAL, want_stress=want_stress, out=out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            ctx.analyze_host(hb, h_ts, critical=CRITICAL, want_stress=want_stress, out=out)
        return allmax((time.perf_counter() - t0) / n_e2e)

    # the whole-grid `analyze` use case consumes only max / argmax / flag per (line, t) (PROTON_cli.py:1058-1066)
    out2 = EvalResult(None, h_mx.numpy(), h_amx.numpy(), h_flag.numpy(), hb.node_off, T)
    dt2 = timed(out2, False)
    reduce_only = {"value": evals_total / dt2, "unit": UNIT, "ms_per_step": dt2 * 1e3,
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(L * T * 13),
                   "note": "proton_em_analyze_host with stress == NULL: only max / argmax / flag come back (the whole-grid "
                           "analyze use case, PROTON_cli.py:1058-1066)"}
    if ring:
        e2e = dict(reduce_only)
        e2e["steps"] = n_e2e
        e2e["note"] = ("the full stress field of this workload does not fit host memory; e2e is the reduce-only host call "
                       "(pack + H2D + eval + D2H of max / argmax / flag), wall clock")
        return e2e
    h_stress = torch.empty(evals_rank, dtype=torch.float64).pin_memory()
    out = EvalResult(h_stress.numpy(), h_mx.numpy(), h_amx.numpy(), h_flag.numpy(), hb.node_off, T)
    dt = timed(out, True)
    d2h = evals_rank * 8 + L * T * 13
    del h_stress, out
    roof_gbs, roof_s = d2h_roof(torch, dist, dev, evals_rank * 8, world, barrier)
    return {"value": evals_total / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "ms_per_step": dt * 1e3, "steps": n_e2e,
            "d2h_roof_gbs": roof_gbs, "d2h_roof_ms": roof_s * 1e3, "frac_of_d2h_roof": roof_s / dt,
            "note": "proton_em_analyze_host through ctypes: pack + H2D + eval + D2H of the full stress field into pinned "
                    "host memory, wall clock, max over ranks; bounded by the device-to-host copy (8 B per eval): "
                    "d2h_roof_gbs is a bare pinned cudaMemcpy of the same size on all ranks at once (aggregate)",
            "reduce_only": reduce_only}


def run_config4(args, ctx, torch, dist, dev, stream, W, world, rank, barrier, allmax):
    """BASELINE config 4: one 10 M-node batch split over the N GPUs; then the whole batch on one GPU for the speed-up."""
    gb, T = build_global("config4")
    parts = W.partition_lines(gb.num_nodes, world)
    mine = gb.subset(parts[rank])
    sh = Shard(ctx, torch, dev, mine, T)
    sh.pack()
    Lmax = max(len(p) for p in parts)
    gth = Gather(torch, dist, dev, stream, Lmax, T, world)

    def gather():
        gth.run(sh)

    def step():
        sh.eval()
        gather()

    n = max(5, min(args.steps, 20))
    for _ in range(3):
        step()
    ms_step = allmax(time_steps(torch, stream, step, n, barrier, gth.finish))
    ms_eval = time_steps(torch, stream, sh.eval, n, barrier)
    ms_gather = allmax(time_steps(torch, stream, gather, n, barrier, gth.finish))
    stats = torch.tensor([ms_eval, float(mine.total_nodes)], device=dev, dtype=torch.float64)
    allstats = torch.empty(world * 2, device=dev, dtype=torch.float64)
    dist.all_gather_into_tensor(allstats, stats)
    a = allstats.cpu().numpy().reshape(world, 2)
    prof_kernels = None
    if rank == 0:
        ctx.profile(True)
        sh.eval()
        ctx.profile(False)
        prof = {k: v for k, v in ctx.profile_read().items() if v[1] > 0}
        prof_kernels = {k: {"ms_per_step": v[0], "launches_per_step": v[1]} for k, v in prof.items()}
    barrier()
    check = check_gathered(gb, parts, timestamps(T), gth.unpack(), Lmax, T, per_rank=2) if rank == 0 else None
    sh.close()
    del sh, gth
    torch.cuda.empty_cache()
    # the whole batch on one GPU (every rank does it on its own GPU; they do not interact)
    whole = Shard(ctx, torch, dev, gb, T)
    whole.pack()
    for _ in range(3):
        whole.eval()
    ms_one = allmax(time_steps(torch, stream, whole.eval, max(3, n // 2), barrier))
    whole.close()
    del whole
    torch.cuda.empty_cache()
    evals = float(gb.total_nodes) * T
    return {"workload": "config4", "lines": int(gb.num_lines), "nodes": int(gb.total_nodes), "timestamps": T,
            "value": evals / (ms_step * 1e-3), "unit": UNIT, "ms_per_step": ms_step,
            "eval_ms_per_rank": [round(float(x), 4) for x in a[:, 0]], "eval_ms_min": float(a[:, 0].min()),
            "eval_ms_max": float(a[:, 0].max()), "imbalance_of_time_max_over_mean": float(a[:, 0].max() / a[:, 0].mean()),
            "nodes_per_rank": [int(x) for x in a[:, 1]], "gather_ms": ms_gather,
            "one_gpu_ms": ms_one, "speedup_vs_one_gpu": ms_one / ms_step,
            "strong_scaling_efficiency": ms_one / ms_step / world, "rank0_kernels": prof_kernels, "gather_check": check,
            "note": "step = evaluation of the rank's shard + all-gather of {max, argmax, flag}; one_gpu_ms = the whole "
                    "10 M-node batch evaluated on one GPU in the same process (max over ranks)"}


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py — node·timestamp stress evaluations per second of the line-analysis hot path on N B200s.

One "step" = one pass of the hot path (modal evaluation + inverse DCT + stress write-back + per-line max/flag) over
the whole batch of lines at all T timestamps.  `value` times the step with the packed batch resident in HBM (CUDA
events on the library's stream); `e2e` times the public host-buffer call (pack + upload + eval + download into pinned
host memory).  `--impl reference` times the reference's own CPU implementation on the host cores.

Workload (config.workload): "ibmpg2_shaped" = BASELINE config 2 — benchmarks/ibmpg2.spice is a missing blob, so it
is ibmpg1's real line population (reference parser) replicated 7x with seeded jitter, T = 100 log-spaced timestamps.
Multi-GPU runs are weak scaling: every rank owns one ibmpg2-shaped grid's worth of lines of a grid N times larger,
assigned by the segment-balanced LPT split, no data-path collective, one all-gather of the per-line reductions.
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
T_DEFAULT = 100
CRITICAL = 40e6


def timestamps(T):
    return np.logspace(6.0, np.log10(6.38e8), T)


def build_workload(name: str, world: int, rank: int):
    """Returns (LineBatch of this rank, description dict)."""
    from proton_b200.api import LineBatch
    from proton_b200 import workloads as W
    if name == "ibmpg2_shaped":
        lines = []
        for r in range(world):       # a grid `world` times larger; every rank builds the same global list
            lines += W.ibmpg2_shaped_lines(replicas=7, seed=2 + r)
        T = T_DEFAULT
    elif name == "ibmpg1":
        lines = W.ibmpg1_lines() * world
        T = T_DEFAULT
    elif name == "synthetic_ragged":
        b = W.synthetic_ragged(20000 * world, seed=0)
        parts = W.partition_lines(b.num_nodes, world)
        return b.subset(parts[rank]), {"lines_total": b.num_lines, "nodes_total": b.total_nodes, "T": 256}
    elif name == "long_lines":
        b = W.synthetic_ragged(0, seed=0, sizes=[65536] * (8 * world))
        parts = W.partition_lines(b.num_nodes, world)
        return b.subset(parts[rank]), {"lines_total": b.num_lines, "nodes_total": b.total_nodes, "T": 1000}
    else:
        raise SystemExit(f"unknown workload {name}")
    nn = np.array([li.num_nodes for li in lines])
    parts = W.partition_lines(nn, world)
    mine = [lines[i] for i in parts[rank]]
    desc = {"lines_total": len(lines), "nodes_total": int(nn.sum()), "T": T,
            "imbalance_max_over_mean": float(max(nn[p].sum() for p in parts) / np.mean([nn[p].sum() for p in parts]))}
    return LineBatch.from_lines(mine), desc


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
    (the e2e leg is bound by host<->device copies).  Returns a description for `config`, or None if unknown."""
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
    import tempfile
    batch, desc = build_workload(args.workload, 1, 0)
    T = desc["T"]
    ts = timestamps(T)
    cores = os.cpu_count() or 1
    config = {"workload": args.workload, "lines": desc["lines_total"], "nodes": desc["nodes_total"], "timestamps": T}
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
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ibmpg2_shaped")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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
    from proton_b200.api import Context, EvalResult
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

    batch, desc = build_workload(args.workload, world, rank)
    T = desc["T"]
    ts = timestamps(T)
    L = batch.num_lines
    evals_rank = batch.total_nodes * T
    ctx = Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    # ---- device-resident inputs and outputs -------------------------------------------------------------
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    d_in = [tt(batch.num_nodes), tt(batch.a_coeff), tt(batch.trip_off), tt(batch.trip_row), tt(batch.trip_col),
            tt(batch.trip_val), tt(batch.u_off), tt(batch.curden)]
    d_ts = tt(ts)
    out_bytes = evals_rank * 8
    chunked = out_bytes > (100 << 30)
    if chunked:
        raise SystemExit("workload output exceeds 100 GB per GPU; use a smaller workload")
    d_stress = torch.empty(evals_rank, dtype=torch.float64, device=dev)
    d_mx = torch.empty(L * T, dtype=torch.float64, device=dev)
    d_amx = torch.empty(L * T, dtype=torch.int32, device=dev)
    d_flag = torch.empty(L * T, dtype=torch.uint8, device=dev)
    Lmax = L
    if world > 1:
        lm = torch.tensor([L], device=dev)
        dist.all_reduce(lm, op=dist.ReduceOp.MAX)
        Lmax = int(lm.item())
    g_send = torch.zeros(Lmax * T, dtype=torch.float64, device=dev)
    g_recv = torch.empty(world * Lmax * T, dtype=torch.float64, device=dev) if world > 1 else None

    t0 = time.perf_counter()
    plan = ctx.pack_device(*d_in)
    ctx.synchronize()
    pack_first_ms = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    plan2 = ctx.pack_device(*d_in)
    ctx.synchronize()
    pack_ms = (time.perf_counter() - t0) * 1e3
    plan2.close()

    def step():
        plan.eval_device(d_ts, d_stress, d_mx, d_amx, d_flag, critical=CRITICAL)
        if world > 1:      # the only exchange of the path: gather the per-line reductions
            with torch.cuda.stream(stream):
                g_send[: L * T].copy_(d_mx)
                dist.all_gather_into_tensor(g_recv, g_send)

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = ctx.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    barrier()
    launches = ctx.launch_count - launches0
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1) / args.steps
    if world > 1:
        tm = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ms = float(tm.item())
        te = torch.tensor([float(evals_rank)], device=dev, dtype=torch.float64)
        dist.all_reduce(te, op=dist.ReduceOp.SUM)
        evals_total = float(te.item())
    else:
        evals_total = float(evals_rank)
    value = evals_total / (ms * 1e-3)

    # ---- roofline (rank 0's share) ---------------------------------------------------------------------------
    # algorithmic bytes per eval: 8 B stress write + the per-line inputs read once per step (16 B per mode:
    # lambda_k and W_k) + 13 B of reductions per (line, t)  ->  8 + 16/T + 13/N  
    peak, peak_src = measured_peaks()
    alg_bytes = evals_rank * 8.0 + batch.total_nodes * 16.0 + L * T * 13.0
    t_rank = ms * 1e-3
    step_gbs = alg_bytes / t_rank / 1e9
    # Per-kernel live timing: the same step again with a CUDA event pair around every launch, on the stream it is
    # launched on (proton_em_profile).  Kernels of different lanes overlap, so the per-kernel sums exceed the step.
    nn = batch.num_nodes.astype(np.int64)
    kinds = np.array([W.kernel_class(int(n)) for n in nn])
    evals_by_kind = {k: float(nn[kinds == k].sum()) * T for k in set(kinds.tolist())}
    long_evals = evals_by_kind.get("long_blue", 0.0) + evals_by_kind.get("long_pow2", 0.0)
    evals_through = {"em_eval_dense_kernel": evals_by_kind.get("dense", 0.0), "em_eval_small_kernel": 0.0,
                     "em_eval_fft_kernel<pow2>": evals_by_kind.get("fft_pow2", 0.0),
                     "em_eval_fft_kernel<bluestein>": evals_by_kind.get("fft_blue", 0.0),
                     "em_long_colfwd_paired_kernel": evals_by_kind.get("long_blue", 0.0),
                     "em_long_row_kernel": long_evals, "em_long_colinv_kernel": long_evals,
                     "em_long_scatter_kernel": evals_by_kind.get("long_pow2", 0.0), "em_long_reduce_kernel": long_evals}
    n_prof = max(1, min(args.steps, 5))
    barrier()
    ctx.profile(True)
    for _ in range(n_prof):
        plan.eval_device(d_ts, d_stress, d_mx, d_amx, d_flag, critical=CRITICAL)
    ctx.profile(False)
    prof = {k: v for k, v in ctx.profile_read().items() if v[1] > 0}
    tot_ms = sum(v[0] for v in prof.values()) or 1.0
    kernels = {k: {"ms_per_step": v[0] / n_prof, "launches_per_step": v[1] / n_prof, "avg_launch_us": 1e3 * v[0] / v[1],
                   "share_of_kernel_time": v[0] / tot_ms} for k, v in prof.items()}
    dom = max(prof, key=lambda k: prof[k][0])
    dom_ms, dom_n = prof[dom][0] / n_prof, prof[dom][1] / n_prof
    dom_bytes = 8.0 * evals_through.get(dom, 0.0)            # per step: 8 B per eval that passes through the kernel
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
    traffic = None
    tp = REPO / "profiles" / "r01_kernel_traffic.json"
    if tp.exists():
        traffic = json.loads(tp.read_text()).get(dom, {}).get("dram_bytes_per_launch")
    # FP64 side: algorithmic flops = sum T N (2.5 log2 N + 5 + 22) against the measured DFMA peak
    flops = float((nn * (2.5 * np.log2(np.maximum(nn, 2)) + 27.0)).sum()) * T
    fp64_peak = 35.9
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": dom_bytes / dom_n, "avg_launch_us": 1e3 * dom_ms / dom_n,
                "launches_per_step": dom_n, "share_of_kernel_time": kernels[dom]["share_of_kernel_time"],
                "step": {"achieved": step_gbs, "frac": step_gbs / peak, "unit": "GB/s",
                         "note": "all launches of one step, 8 + 16/T + 13/N bytes per eval over the step time"},
                "fp64": {"achieved": flops / t_rank / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": flops / t_rank / 1e12 / fp64_peak,
                         "note": "algorithmic flops T*N*(2.5 log2 N + 27) over the step time against the measured DFMA "
                                 "peak (profiles/r01_fp64_peak.json); the kernels are FP64-pipe / shared-memory bound, "
                                 "not HBM bound (DESIGN.md §5)"},
                "kernels": kernels,
                "note": "dominant kernel by summed live duration; achieved = 8 B x evals passing through it per step / its "
                        "summed launch durations (CUDA events on the launching stream, kernels of other lanes running "
                        "concurrently)"}

    # ---- e2e: public host-buffer call, pinned host memory, copies inside the timed region -----------------
    e2e = None
    if not args.no_e2e:
        pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        from proton_b200.api import LineBatch
        hp = [pin(x) for x in (batch.num_nodes, batch.a_coeff, batch.trip_off, batch.trip_row, batch.trip_col,
                               batch.trip_val, batch.u_off, batch.curden)]
        hb = LineBatch(*[t.numpy() for t in hp])
        h_stress = torch.empty(evals_rank, dtype=torch.float64).pin_memory()
        h_mx = torch.empty((L, T), dtype=torch.float64).pin_memory()
        h_amx = torch.empty((L, T), dtype=torch.int32).pin_memory()
        h_flag = torch.empty((L, T), dtype=torch.uint8).pin_memory()
        out = EvalResult(h_stress.numpy(), h_mx.numpy(), h_amx.numpy(), h_flag.numpy(), hb.node_off, T)
        h_ts = pin(ts).numpy()
        n_e2e = max(2, min(args.steps, 5))
        ctx.analyze_host(hb, h_ts, critical=CRITICAL, out=out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            ctx.analyze_host(hb, h_ts, critical=CRITICAL, out=out)
        dt = (time.perf_counter() - t0) / n_e2e
        if world > 1:
            tm = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            dt = float(tm.item())
        h2d = sum(t.numel() * t.element_size() for t in hp) + T * 8
        d2h = evals_rank * 8 + L * T * 13
        e2e = {"value": evals_total / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms_per_step": dt * 1e3, "steps": n_e2e,
               "note": "proton_em_analyze_host through ctypes: pack + H2D + eval + D2H of the full stress field "
                       "into pinned host memory, wall clock; bounded by PCIe (8 B per eval)"}
        # the whole-grid `analyze` use case consumes only max / argmax / flag per (line, t) (PROTON_cli.py:1058-1066)
        out2 = EvalResult(None, h_mx.numpy(), h_amx.numpy(), h_flag.numpy(), hb.node_off, T)
        ctx.analyze_host(hb, h_ts, critical=CRITICAL, want_stress=False, out=out2)
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            ctx.analyze_host(hb, h_ts, critical=CRITICAL, want_stress=False, out=out2)
        dt2 = (time.perf_counter() - t0) / n_e2e
        if world > 1:
            tm = torch.tensor([dt2], device=dev, dtype=torch.float64)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            dt2 = float(tm.item())
        e2e["reduce_only"] = {"value": evals_total / dt2, "unit": UNIT, "ms_per_step": dt2 * 1e3,
                              "d2h_bytes_per_step": int(L * T * 13),
                              "note": "same call with stress == NULL: only max / argmax / flag come back"}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)          # the CPU baseline gets every host core
        cpu_baseline = cpu_baseline_port(batch, ts)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": args.workload, "lines": desc["lines_total"], "nodes": desc["nodes_total"],
                           "timestamps": T, "lines_per_gpu": L, "nodes_per_gpu": batch.total_nodes,
                           "max_nodes_per_line": int(batch.num_nodes.max()),
                           "l2": f"outputs of one step are {out_bytes / 2**30:.1f} GiB per GPU, far larger than the 126 MB L2",
                           "parallelism": f"lines sharded over {world} GPU(s), LPT segment-balanced",
                           "imbalance_max_over_mean": desc.get("imbalance_max_over_mean"), "host_affinity": numa},
                "clocks": clocks, "gpu_launches": int(launches), "roofline": roofline, "e2e": e2e,
                "cpu_baseline": cpu_baseline,
                "pack_ms": {"first_call_with_table_build": pack_first_ms, "cached_tables": pack_ms}}
        print(json.dumps(line))
    plan.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

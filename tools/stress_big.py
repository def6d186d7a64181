"""tools/stress_big.py — config 3 of BASELINE.json at full line count: 1 000 000 synthetic ragged lines (N log-uniform in
16..4096, 7.4e8 nodes), reduce-only (the full field would be 1.5 TB at T = 256).  STRESS_T sets the number of timestamps."""
import os, sys, time, numpy as np
sys.path.insert(0, '.')
from proton_b200.api import Context
from proton_b200.workloads import synthetic_ragged
ctx = Context(0)
t0 = time.time(); b = synthetic_ragged(1_000_000, seed=0); print("gen", time.time() - t0, "lines", b.num_lines, "nodes", b.total_nodes, flush=True)
T = int(os.environ.get('STRESS_T', '8'))
ts = np.logspace(6, np.log10(6.38e8), T)
t0 = time.time(); r = ctx.analyze_host(b, ts, critical=40e6, want_stress=False); dt = time.time() - t0
print("1M lines reduce-only first call %.2f s -> %.3e evals/s" % (dt, b.total_nodes * T / dt), flush=True)
t0 = time.time(); r = ctx.analyze_host(b, ts, critical=40e6, want_stress=False); dt = time.time() - t0
print("1M lines reduce-only second call %.2f s -> %.3e evals/s; flagged %d, max %.3e, nan %d" % (dt, b.total_nodes * T / dt, int(r.flag.sum()), r.max_stress.max(), int(np.isnan(r.max_stress).sum())), flush=True)
sub = b.subset(np.arange(0, 1_000_000, 5))
t0 = time.time(); r2 = ctx.analyze_host(sub, ts[:4], critical=40e6); dt = time.time() - t0
print("200k lines full field T=4 %.2f s, %.1f GB; consistent max: %s" % (dt, r2.stress.nbytes / 1e9, np.array_equal(r2.max_stress[:50, :], np.array([r2.line(l).max(axis=1) for l in range(50)]))), flush=True)
print("same maxima as the big run:", np.allclose(r2.max_stress, r.max_stress[::5, :4], rtol=1e-12))

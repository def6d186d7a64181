"""Initial stress: sigma(t) = V exp(Lambda a t) V^T sigma0 + the reference closed form.

The reference kernel has no sigma_0 input (bin/EMtool_analytical@0xc68c), so there is no reference output to pin
against: the CPU oracle of this feature (oracle.sigma0_batch, scipy DCTs) is pinned here against a stiff ODE integration
of the FDM system from sigma0 — code that shares nothing with either transform path — and the GPU path
(proton_em_pack_sigma0: dense forward DCT-II kernel + the unchanged evaluation kernels) against that oracle."""
import numpy as np
import pytest

from gpu_common import synth_batch, synth_line
from proton_b200.api import LineBatch
from test_history import _laplacian


def _field(rng, batch, scale=3e7):
    """A rough initial field: smooth part + noise, so that every mode is excited."""
    out = []
    for n in batch.num_nodes:
        x = np.linspace(0, 1, int(n))
        out.append(scale * (np.sin(3 * x) - 0.3 + 0.2 * rng.normal(size=int(n))))
    return np.concatenate(out)


@pytest.mark.parametrize("N", [6, 25, 40])
def test_sigma0_oracle_solves_the_ode(oracle, N):
    from scipy.integrate import solve_ivp
    rng = np.random.default_rng(40 + N)
    a = 2.0e-7
    li = synth_line(rng, N, a=a)
    b = LineBatch.from_lines([li])
    s0 = _field(rng, b)
    ts = np.array([1.0e5, 1.0e6, 1.0e7, 6.0e7])
    want, mx, amx = oracle.sigma0_batch(b.num_nodes, b.a_coeff, b.trip_off, b.trip_row, b.trip_col, b.trip_val, b.u_off,
                                        b.curden, s0, ts)
    want = want.reshape(len(ts), N)
    A = a * _laplacian(N)
    r = oracle.spmv(N, len(li.curden), li.rows, li.cols, li.vals, li.curden)
    sol = solve_ivp(lambda t, y: A @ y + r, (0.0, ts.max()), s0, method="Radau", jac=A, rtol=1e-11,
                    atol=1e-15 * np.abs(want).max(), dense_output=True)
    got = np.array([sol.sol(t) for t in ts])
    err = np.abs(got - want).max(axis=1) / np.abs(want).max(axis=1)
    assert err.max() < 1e-7, err
    assert np.array_equal(amx[0], want.argmax(axis=1))
    # zero initial stress is the reference solution
    zero, _, _ = oracle.sigma0_batch(b.num_nodes, b.a_coeff, b.trip_off, b.trip_row, b.trip_col, b.trip_val, b.u_off,
                                     b.curden, np.zeros(N), ts)
    ref, _, _ = oracle.batch(b.num_nodes, b.a_coeff, b.trip_off, b.trip_row, b.trip_col, b.trip_val, b.u_off, b.curden, ts)
    assert np.array_equal(zero, ref)


@pytest.mark.gpu
def test_gpu_initial_stress_against_the_oracle_all_length_classes(oracle):
    from proton_b200.api import Context
    from gpu_common import compare, self_consistent
    rng = np.random.default_rng(50)
    sizes = [1, 2, 5, 33, 64, 88, 100, 256, 300, 1000, 1024, 2049, 4096, 5000, 8192, 9001, 16384, 20001]
    batch = synth_batch(rng, sizes)
    s0 = _field(rng, batch)
    ts = np.array([1e5, 1e6, 3e7, 6.38e8, 1e8])
    ctx = Context(0)
    try:
        plan = ctx.pack(batch, sigma0=s0)
        got = plan.eval(ts, critical=40e6)
        ws, wm, wa = oracle.sigma0_batch(batch.num_nodes, batch.a_coeff, batch.trip_off, batch.trip_row, batch.trip_col,
                                         batch.trip_val, batch.u_off, batch.curden, s0, ts)
        worst = compare(batch, ts, got, ws, wm, wa, critical=40e6)
        self_consistent(batch, ts, got, 40e6)
        print(f"initial stress: worst norm-wise error {worst:.3e}")
        # reduce-only gives the same reductions; zero initial stress gives the plain plan's bits
        red = plan.eval(ts, critical=40e6, want_stress=False)
        assert np.array_equal(red.max_stress, got.max_stress) and np.array_equal(red.argmax, got.argmax)
        assert np.array_equal(red.flag, got.flag)
        plan.close()
        z = ctx.pack(batch, sigma0=np.zeros_like(s0))
        plain = ctx.pack(batch)
        a, b = z.eval(ts, critical=40e6), plain.eval(ts, critical=40e6)
        assert np.array_equal(a.stress, b.stress) and np.array_equal(a.argmax, b.argmax)
        z.close(), plain.close()
    finally:
        ctx.close()


@pytest.mark.gpu
def test_restart_from_the_evaluated_field_equals_the_current_history(oracle):
    """proton_em_analyze_history_host (superposition of current increments) re-expressed as restarts: evaluate to the
    switching time, hand the field back as sigma0 with the next currents, continue.
    Two statements.  (i) The GPU restart equals the CPU oracle's restart from the same field to rounding (1e-12).
    (ii) Restart and superposition agree to 1e-9: they are the same mathematics, but they call the reference's closed
    form at different times (t and t - tau), and that form loses digits through exp(x) - 1.0 (SURVEY.md F10,
    tests/test_truth.py) — the CPU oracles of the two routes differ from each other by the same 7.5e-11 (N = 3000) that
    the GPU routes do, measured on B200."""
    from proton_b200.api import Context
    rng = np.random.default_rng(51)
    batch = synth_batch(rng, [5, 40, 64, 300, 1000, 3000, 5000, 9000])
    u1 = np.asarray(batch.curden)
    u2 = np.float32(rng.normal(0, 1e11, u1.shape)).astype(np.float64)
    u3 = 0.25 * u1
    tau = np.array([0.0, 2.0e7, 2.0e8])
    ts = np.array([2.5e7, 1.9e8, 2.0e8, 3.0e8, 6.38e8])
    T = len(ts)
    off = batch.node_off
    args = (batch.num_nodes, batch.a_coeff, batch.trip_off, batch.trip_row, batch.trip_col, batch.trip_val, batch.u_off)
    ctx = Context(0)
    try:
        hist = ctx.analyze_history(batch, [u1, u2, u3], tau, ts, critical=40e6)
        f1 = ctx.pack(batch).eval(np.array([tau[1]])).stress              # segment 1 from zero to tau_1
        b2 = LineBatch(*args, u2)
        p2 = ctx.pack(b2, sigma0=f1)
        in2 = ts <= tau[2]
        r2 = p2.eval(ts[in2] - tau[1], critical=40e6)
        f2 = p2.eval(np.array([tau[2] - tau[1]])).stress
        r3 = ctx.pack(LineBatch(*args, u3), sigma0=f2).eval(ts[~in2] - tau[2], critical=40e6)
        o2, _, _ = oracle.sigma0_batch(*args, u2, f1, ts[in2] - tau[1])
        o3, _, _ = oracle.sigma0_batch(*args, u3, f2, ts[~in2] - tau[2])
        T2, T3 = int(in2.sum()), int((~in2).sum())
        worst_restart = worst_routes = 0.0
        for l in range(batch.num_lines):
            n = int(batch.num_nodes[l])
            want = hist.line(l)
            got = np.concatenate([r2.line(l), r3.line(l)])
            orc = np.concatenate([o2[off[l] * T2:off[l + 1] * T2].reshape(T2, n), o3[off[l] * T3:off[l + 1] * T3].reshape(T3, n)])
            scale = np.abs(want).max(axis=1)
            worst_restart = max(worst_restart, float((np.abs(got - orc).max(axis=1) / scale).max()))
            worst_routes = max(worst_routes, float((np.abs(got - want).max(axis=1) / scale).max()))
        print(f"restart: GPU vs oracle restart {worst_restart:.2e}; restart vs superposition {worst_routes:.2e}")
        assert worst_restart < 1e-12
        assert worst_routes < 1e-9
    finally:
        ctx.close()

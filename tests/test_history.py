"""Current histories: the superposition rule that proton_em_analyze_history_host implements is
checked here, on the CPU, against something that shares no code with it — a stiff ODE integration of the FDM system
    d sigma / dt = a L sigma + B u(t),   sigma(0) = 0,
with L the Neumann second-difference matrix whose eigenpairs the reference's closed form uses (lambda_k = 2cos(k pi/N) - 2,
DCT-II vectors).  The GPU side of the feature is in tests/test_gpu_parity.py."""
import numpy as np
import pytest

from gpu_common import synth_line
from proton_b200.api import LineBatch


def _laplacian(N):
    Lm = np.zeros((N, N))
    for i in range(N):
        if i > 0:
            Lm[i, i - 1] = 1.0
            Lm[i, i] -= 1.0
        if i < N - 1:
            Lm[i, i + 1] = 1.0
            Lm[i, i] -= 1.0
    return Lm


@pytest.mark.parametrize("N", [7, 24, 40])
def test_superposition_of_current_increments_solves_the_ode(oracle, N):
    from scipy.integrate import solve_ivp
    rng = np.random.default_rng(N)
    a = 2.0e-7                                     # a * t up to ~20: the decaying modes matter
    li = synth_line(rng, N, a=a)
    b = LineBatch.from_lines([li])
    seg_start = np.array([0.0, 2.0e7, 5.5e7])
    segs = [li.curden, li.curden * -0.5 + np.float32(rng.normal(0, 1e11, len(li.curden))), np.zeros_like(li.curden)]
    ts = np.array([1.0e6, 1.9e7, 2.0e7, 3.1e7, 5.5e7, 6.0e7, 1.0e8])
    want, mx, amx = oracle.history_batch(b.num_nodes, b.a_coeff, b.trip_off, b.trip_row, b.trip_col, b.trip_val, b.u_off,
                                         segs, seg_start, ts)
    want = want.reshape(len(ts), N)
    # independent: integrate segment by segment
    A = a * _laplacian(N)
    r = [oracle.spmv(N, len(li.curden), li.rows, li.cols, li.vals, u) for u in segs]
    sigma = np.zeros(N)
    t0 = 0.0
    got = np.zeros_like(want)
    edges = list(seg_start[1:]) + [np.inf]
    for s, t1 in enumerate(edges):
        sel = [j for j, t in enumerate(ts) if t0 < t <= t1]
        t_end = min(t1, ts.max())
        if t_end > t0:
            sol = solve_ivp(lambda t, y: A @ y + r[s], (t0, t_end), sigma, method="Radau", jac=A, rtol=1e-11,
                            atol=1e-9 * max(1.0, np.abs(want).max()) * 1e-6, dense_output=True)
            for j in sel:
                got[j] = sol.sol(ts[j])
            sigma = sol.y[:, -1]
        t0 = t1
        if t0 >= ts.max():
            break
    scale = np.abs(want).max(axis=1)
    err = np.abs(got - want).max(axis=1) / scale
    assert err.max() < 1e-7, err
    assert np.array_equal(amx[0], want.argmax(axis=1)) and np.allclose(mx[0], want.max(axis=1))


def test_history_with_one_segment_is_the_reference_solution(oracle):
    rng = np.random.default_rng(3)
    li = synth_line(rng, 33)
    b = LineBatch.from_lines([li])
    ts = np.array([1e6, 1e8])
    one, _, _ = oracle.history_batch(b.num_nodes, b.a_coeff, b.trip_off, b.trip_row, b.trip_col, b.trip_val, b.u_off,
                                     [b.curden], [0.0], ts)
    ref, _, _ = oracle.batch(b.num_nodes, b.a_coeff, b.trip_off, b.trip_row, b.trip_col, b.trip_val, b.u_off, b.curden, ts)
    assert np.array_equal(one, ref)

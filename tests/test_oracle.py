"""CPU tests that pin the oracle (oracle/em_oracle.c) — SURVEY.md §8c.

The reference has no tests or golden vectors of its own (F9).  The goldens here are outputs of the
unmodified reference binary run under oracle/shim (tools/gen_golden.py); wherever the binary is
present (this container) it is also run live.
"""
import numpy as np
import pytest


def _case(g, i):
    t0, t1 = g["trip_off"][i], g["trip_off"][i + 1]
    u0, u1 = g["u_off"][i], g["u_off"][i + 1]
    s0, s1 = g["stress_off"][i], g["stress_off"][i + 1]
    return (int(g["num_nodes"][i]), float(g["a_coeff"][i]), g["rows"][t0:t1], g["cols"][t0:t1], g["vals"][t0:t1],
            g["curden"][u0:u1], float(g["t"][i]), g["stress"][s0:s1])


def test_restatement_matches_reference_binary_goldens_bit_exact(oracle, binary_golden):
    g = binary_golden
    n = len(g["t"])
    assert n >= 400
    for i in range(n):
        N, a, rows, cols, vals, u, t, want = _case(g, i)
        got, mx, amx = oracle.line(N, a, rows, cols, vals, u, [t])
        # same libm, same op order, same (shim) DCT  =>  identical bits
        assert np.array_equal(got[0], want), g["tag"][i]
        assert mx[0] == want.max() and amx[0] == int(np.argmax(want))


def test_double_precision_port_within_1e13(oracle, binary_golden):
    g = binary_golden
    for i in range(0, len(g["t"]), 7):
        N, a, rows, cols, vals, u, t, want = _case(g, i)
        got, _, _ = oracle.line(N, a, rows, cols, vals, u, [t], use_double=True)
        scale = np.abs(want).max()
        assert np.abs(got[0] - want).max() <= 1e-13 * scale


def test_live_reference_binary(oracle, binary_golden):
    if not oracle.ref_binary_available():
        pytest.skip("oracle/_ref not built here")
    g = binary_golden
    for i in list(range(0, len(g["t"]), 40))[:8]:
        N, a, rows, cols, vals, u, t, want = _case(g, i)
        live = oracle.reference_binary_line(N, a, rows, cols, vals, u, t)
        assert np.array_equal(live, want)


@pytest.mark.parametrize("n", [1, 2, 3, 5, 8, 12, 17, 64, 100, 129])
def test_r2r_against_direct_cosine_sums(oracle, n):
    rng = np.random.default_rng(n)
    x = rng.normal(size=n)
    j = np.arange(n)
    k = j[:, None]
    want10 = 2.0 * (np.cos(np.pi * (j[None, :] + 0.5) * k / n) * x[None, :]).sum(1)
    want01 = x[0] + 2.0 * (np.cos(np.pi * j[None, 1:] * (k + 0.5) / n) * x[None, 1:]).sum(1)
    assert np.allclose(oracle.r2r(x, 5), want10, rtol=0, atol=1e-13 * n)
    assert np.allclose(oracle.r2r(x, 4), want01, rtol=0, atol=1e-13 * n)
    assert np.allclose(oracle.r2r(x, 5, use_double=True), want10, rtol=0, atol=1e-12 * n)
    # REDFT01(REDFT10(x)) = 2n x
    assert np.allclose(oracle.r2r(oracle.r2r(x, 5), 4), 2 * n * x, rtol=0, atol=1e-12 * n)


def test_closed_forms(oracle):
    # N = 1: lambda_0 == 0  =>  sigma = t * r_0  
    st, mx, amx = oracle.line(1, 2e-9, [0], [0], [-8e-14], [1.25e10], [1e6, 6.38e8])
    assert st[0, 0] == 1e6 * (-8e-14 * 1.25e10) and st[1, 0] == 6.38e8 * (-8e-14 * 1.25e10)
    # steady state (t -> inf): zero-mean, piecewise linear with per-node step -(bcoef/acoef) j_s
    N, a, b = 200, 5e-9, 8e-14
    rows, cols, vals = [0, 80, 80, N - 1], [0, 0, 1, 1], [b, -b, b, -b]
    u = np.array([1e10, -3e10])
    st, _, _ = oracle.line(N, a, rows, cols, vals, u, [1e15])
    s = st[0]
    assert abs(s.mean()) <= 1e-9 * np.abs(s).max()
    d = np.diff(s)
    assert np.allclose(d[:80], -(b / a) * u[0], rtol=1e-6)
    assert np.allclose(d[80:], -(b / a) * u[1], rtol=1e-6)
    # small t: sigma ~= t * B u  (at the via rows; first order in a t)
    st, _, _ = oracle.line(N, a, rows, cols, vals, u, [1.0])
    r = oracle.spmv(N, 2, rows, cols, vals, u)
    assert np.allclose(st[0][[0, 80, N - 1]], r[[0, 80, N - 1]], rtol=1e-6)


def test_cross_solver_backward_euler(oracle):
    """Independent check from the reference's numerical path (simulation/simulation.py:217-277):
    x' = G x + B u, x(0) = 0 with G = a*tridiag(1,-2,1), Neumann ends."""
    N, a, b, t_end = 40, 2e-9, 8e-14, 2.0e8
    rows, cols, vals = [0, 15, 15, N - 1], [0, 0, 1, 1], [b, -b, b, -b]
    u = np.array([2e10, -1e10])
    r = oracle.spmv(N, 2, rows, cols, vals, u)
    G = np.diag(np.full(N, -2.0)) + np.diag(np.ones(N - 1), 1) + np.diag(np.ones(N - 1), -1)
    G[0, 0] = G[-1, -1] = -1.0
    G *= a
    steps = 20000
    h = t_end / steps
    M = np.linalg.inv(np.eye(N) - h * G)
    x = np.zeros(N)
    for _ in range(steps):
        x = M @ (x + h * r)
    st, _, _ = oracle.line(N, a, rows, cols, vals, u, [t_end])
    assert np.abs(st[0] - x).max() <= 2e-4 * np.abs(x).max()


def test_batch_matches_single_lines(oracle, binary_golden):
    g = binary_golden
    idx = [i for i in range(len(g["t"])) if g["t"][i] == 6.38e8][:40]
    Ns = g["num_nodes"][idx]
    a = g["a_coeff"][idx]
    rows = np.concatenate([g["rows"][g["trip_off"][i]:g["trip_off"][i + 1]] for i in idx])
    cols = np.concatenate([g["cols"][g["trip_off"][i]:g["trip_off"][i + 1]] for i in idx])
    vals = np.concatenate([g["vals"][g["trip_off"][i]:g["trip_off"][i + 1]] for i in idx])
    u = np.concatenate([g["curden"][g["u_off"][i]:g["u_off"][i + 1]] for i in idx])
    toff = np.concatenate([[0], np.cumsum([g["trip_off"][i + 1] - g["trip_off"][i] for i in idx])])
    uoff = np.concatenate([[0], np.cumsum([g["u_off"][i + 1] - g["u_off"][i] for i in idx])])
    ts = np.array([1e6, 6.38e8])
    stress, mx, amx = oracle.batch(Ns, a, toff, rows, cols, vals, uoff, u, ts, nthreads=3)
    off = 0
    for k, i in enumerate(idx):
        N = int(Ns[k])
        blk = stress[off * 2:(off + N) * 2].reshape(2, N)
        want = g["stress"][g["stress_off"][i]:g["stress_off"][i + 1]]
        assert np.array_equal(blk[1], want)
        assert mx[k, 1] == want.max() and amx[k, 1] == int(np.argmax(want))
        off += N

"""Where does the 1e-9 parity bound hold, and which side is right where the reference cancels?

oracle/mp_truth.py evaluates the discretised solution without the reference's exp(x) - 1.0 cancellation (mpmath expm1).
The tests tabulate, for short simulated times, the norm-wise distance of the reference arithmetic (CPU oracle, pinned
bit for bit on the reference binary) and of the GPU path from that truth, and the GPU-vs-reference distance that the
parity bound is about (SURVEY.md F10, §7 hard part 1)."""
import numpy as np
import pytest

from gpu_common import synth_line

TS_SHORT = np.array([1e3, 1e4, 1e5, 1e6])


def _r_of(oracle, li):
    r = oracle.spmv(li.num_nodes, len(li.curden), li.rows, li.cols, li.vals, li.curden)
    idx = np.nonzero(r)[0]
    return idx, r[idx]


def _nodes(li, rng):
    N = li.num_nodes
    base = [0, N - 1, N // 2] + [int(x) for x in np.unique(li.rows)][:4]
    return np.unique(np.clip(np.concatenate([base, rng.integers(0, N, 4)]), 0, N - 1))


def test_truth_agrees_with_the_reference_arithmetic_where_nothing_cancels(oracle):
    """At the grid-analysis timestamps (t >= 1e6 s) the reference formula and the exact evaluation agree well inside
    the 1e-9 parity bound (the residual, up to ~1e-10 at t = 1e6 s, is the reference's own exp(x) - 1.0 cancellation)."""
    from oracle import mp_truth
    rng = np.random.default_rng(12)
    for N in (7, 199, 1024):
        li = synth_line(rng, N, 3)
        ts = np.array([1e6, 3e7, 6.38e8])
        want, _, _ = oracle.line(li.num_nodes, li.a_coeff, li.rows, li.cols, li.vals, li.curden, ts)
        nodes = _nodes(li, rng)
        idx, val = _r_of(oracle, li)
        tr = mp_truth.truth_nodes(N, li.a_coeff, idx, val, ts, nodes)
        scale = np.abs(want).max(axis=1)
        err = np.abs(want[:, nodes] - tr).max(axis=1) / scale
        print(f"N = {N}: reference arithmetic vs truth at t = {ts.tolist()}: {err.tolist()}")
        assert np.all(err < 1e-9) and err[-1] < 5e-12, (N, err)


def test_reference_formula_loses_digits_at_short_times(oracle):
    """The cancellation is real: with a small diffusivity and t = 1e3 s the reference arithmetic is orders of magnitude
    further from the truth than at t = 1e6 s (this is what the GPU path has to reproduce, not to fix)."""
    from oracle import mp_truth
    rng = np.random.default_rng(13)
    li = synth_line(rng, 300, 4, a=3.3586e-21 / (40e-6) ** 2)
    want, _, _ = oracle.line(li.num_nodes, li.a_coeff, li.rows, li.cols, li.vals, li.curden, TS_SHORT)
    nodes = _nodes(li, rng)
    idx, val = _r_of(oracle, li)
    tr = mp_truth.truth_nodes(li.num_nodes, li.a_coeff, idx, val, TS_SHORT, nodes)
    err = np.abs(want[:, nodes] - tr).max(axis=1) / np.abs(want).max(axis=1)
    print("reference arithmetic vs truth, N = 300, dx = 40 um:", dict(zip(TS_SHORT.tolist(), err.tolist())))
    assert err[0] > 100 * max(err[-1], 1e-16)


@pytest.mark.gpu
def test_gpu_and_reference_against_truth_in_the_cancellation_regime(oracle):
    """Table of norm-wise errors for t in {1e3, 1e4, 1e5, 1e6} s at N in {199, 4096, 65536}: reference arithmetic vs
    truth, GPU vs truth, GPU vs reference.  Asserted: the GPU is never further from the truth than the reference
    arithmetic by more than rounding, and GPU-vs-reference meets 1e-9 for every entry of the table — the regime where a
    1e-9 comparison would fail does not begin before the reference itself has lost all digits."""
    from oracle import mp_truth
    from proton_b200.api import Context, LineBatch
    rng = np.random.default_rng(14)
    ctx = Context(0)
    rows = []
    try:
        for N, dx in ((199, 12e-6), (199, 40e-6), (4096, 5e-6), (4096, 40e-6), (65536, 1e-6), (65536, 12e-6)):
            li = synth_line(rng, N, 4, a=3.3586e-21 / dx ** 2)
            batch = LineBatch.from_lines([li])
            got = ctx.analyze_host(batch, TS_SHORT).line(0)
            want, _, _ = oracle.line(li.num_nodes, li.a_coeff, li.rows, li.cols, li.vals, li.curden, TS_SHORT)
            nodes = _nodes(li, rng)
            idx, val = _r_of(oracle, li)
            tr = mp_truth.truth_nodes(N, li.a_coeff, idx, val, TS_SHORT, nodes)
            scale = np.abs(want).max(axis=1)
            e_ref = np.abs(want[:, nodes] - tr).max(axis=1) / scale
            e_gpu = np.abs(got[:, nodes] - tr).max(axis=1) / scale
            d = np.abs(got - want).max(axis=1) / scale
            for j, t in enumerate(TS_SHORT):
                rows.append((N, dx, t, e_ref[j], e_gpu[j], d[j]))
    finally:
        ctx.close()
    print("\n     N   dx[um]      t[s]   ref-vs-truth   gpu-vs-truth   gpu-vs-ref")
    for N, dx, t, er, eg, d in rows:
        print(f"{N:6d} {dx * 1e6:8.1f} {t:9.0e} {er:14.3e} {eg:14.3e} {d:12.3e}")
    for N, dx, t, er, eg, d in rows:
        assert d <= 1e-9, (N, dx, t, d)
        assert eg <= 2.0 * er + 1e-12, (N, dx, t, er, eg)

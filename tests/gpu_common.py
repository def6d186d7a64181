"""Shared helpers for the GPU parity tests (CUDA path through the C ABI vs the CPU oracle)."""
import numpy as np

from proton_b200.api import LineBatch


def synth_line(rng, N, S=None, a=None, bcoef=None):
    """A line shaped like what matrix_formulation emits: +b at node 0, (-b, +b) at via rows, -b at N-1."""
    bcoef = 8.29e-14 * rng.uniform(0.5, 2.0) if bcoef is None else bcoef
    a = 3.3586e-21 / (rng.uniform(1, 12) * 1e-6) ** 2 if a is None else a
    if S is None:
        S = 1 + rng.poisson(4)
    if N == 1:
        rows, cols, vals, S = [0], [0], [-bcoef], 1
    else:
        S = max(1, min(S, N - 1))
        if N == 2:
            S = 1
        vias = np.sort(rng.choice(np.arange(1, N - 1), size=S - 1, replace=False)) if S > 1 else []
        rows, cols, vals = [0], [0], [bcoef]
        for s, v in enumerate(vias, start=1):
            rows += [int(v), int(v)]
            cols += [s - 1, s]
            vals += [-bcoef, bcoef]
        rows.append(N - 1), cols.append(S - 1), vals.append(-bcoef)
    u = np.float32(rng.normal(0, 1e11, S)).astype(np.float64)

    class L:
        pass
    li = L()
    li.num_nodes, li.a_coeff = int(N), float(a)
    li.rows, li.cols, li.vals, li.curden = np.array(rows, np.int32), np.array(cols, np.int32), np.array(vals), u
    li.name = f"synth_N{N}"
    return li


def synth_batch(rng, sizes, **kw):
    return LineBatch.from_lines([synth_line(rng, int(n), **kw) for n in sizes])


def oracle_batch(oracle, batch: LineBatch, ts, want_stress=True):
    return oracle.batch(batch.num_nodes, batch.a_coeff, batch.trip_off, batch.trip_row, batch.trip_col,
                        batch.trip_val, batch.u_off, batch.curden, ts, want_stress=want_stress)


class Escapes:
    """How often a comparison with the oracle needed the near-tie escape: two different DCT algorithms cannot agree in
    the last bit (SURVEY.md §7 hard part 2), so an arg-max may move between two nodes whose stresses differ by less
    than the tolerance, and a flag may flip when the maximum sits within the tolerance of the threshold.  Every use of
    the escape is counted here; the tests assert the counts against committed bounds (bit-exact agreement of max /
    argmax / flag is asserted separately against the CPU twin of the GPU arithmetic, tests/test_twin.py)."""

    def __init__(self):
        self.rows = self.argmax = self.flag = 0

    def __repr__(self):
        return f"{self.argmax} arg-max and {self.flag} flag escapes in {self.rows} (line, t) rows"


def compare(batch: LineBatch, ts, got, want_stress, want_mx, want_amx, tol=1e-9, critical=None, escapes: Escapes = None):
    """Norm-wise parity per (line, t): max|d| <= tol * max|sigma| (SURVEY.md §7 hard part 1-iii); max within
    tol; arg-max equal unless the runner-up is within tol of the maximum; flag equal unless the maximum sits
    within tol of the threshold (both escapes are counted in `escapes`).  Returns the worst norm-wise error seen."""
    T = len(ts)
    off = batch.node_off
    worst = 0.0
    esc = escapes if escapes is not None else Escapes()
    for l in range(batch.num_lines):
        N = int(batch.num_nodes[l])
        g = got.stress[off[l] * T:off[l + 1] * T].reshape(T, N)
        w = want_stress[off[l] * T:off[l + 1] * T].reshape(T, N)
        scale = np.abs(w).max(axis=1)
        err = np.abs(g - w).max(axis=1)
        ok = err <= tol * scale + 1e-300
        assert ok.all(), f"line {l} (N={N}): norm-wise error {(err / np.maximum(scale, 1e-300)).max():.3e}"
        worst = max(worst, float((err / np.maximum(scale, 1e-300)).max()))
        assert np.all(np.abs(got.max_stress[l] - want_mx[l]) <= tol * scale + 1e-300)
        esc.rows += T
        for j in np.nonzero(got.argmax[l] != want_amx[l])[0]:
            esc.argmax += 1
            assert abs(w[j, got.argmax[l, j]] - w[j, want_amx[l, j]]) <= 2 * tol * scale[j], (l, j)
        if critical is not None:
            for j in np.nonzero(got.flag[l].astype(bool) != (want_mx[l] > critical))[0]:
                esc.flag += 1
                assert abs(want_mx[l, j] - critical) <= 2 * tol * scale[j], (l, j)
    return worst


def self_consistent(batch: LineBatch, ts, got, critical):
    """Bit-exact internal consistency: the fused reductions equal what a consumer computes from the field."""
    T = len(ts)
    off = batch.node_off
    for l in range(batch.num_lines):
        N = int(batch.num_nodes[l])
        g = got.stress[off[l] * T:off[l + 1] * T].reshape(T, N)
        assert np.array_equal(got.max_stress[l], g.max(axis=1))
        assert np.array_equal(got.argmax[l], g.argmax(axis=1))
        assert np.array_equal(got.flag[l].astype(bool), g.max(axis=1) > critical)

"""Bit-exact statement of parity (SURVEY.md §7 hard part 2): the stress field the GPU writes — and therefore the
per-line max, arg-max and void-nucleation flag taken from it — equals, bit for bit, what oracle/em_twin.c computes on
the CPU by repeating the kernels' own arithmetic (same butterflies, same order, same fused multiply-adds), from the same
tables.  The norm-wise 1e-9 comparison with the reference's algorithm lives in tests/test_gpu_parity.py; the two
together say: the GPU result is a deterministic, reproducible function of its inputs, and that function is within
1e-9 of the reference."""
import numpy as np
import pytest

from gpu_common import synth_batch

pytestmark = pytest.mark.gpu

TS = np.array([1e3, 1e6, 3e6, 2e7, 1e8, 6.38e8, 5e8])      # odd count: the last transform carries one timestamp


def _twin_of(oracle, plan, batch, l, kind, ts):
    N = int(batch.num_nodes[l])
    lam, W = plan.debug_table(l, "lambda"), plan.debug_table(l, "W")
    table = plan.debug_table(l, "costab" if kind == "dense" else "pre")
    return oracle.twin_line(kind, N, float(batch.a_coeff[l]), lam, W, table, ts)


@pytest.mark.parametrize("kind,sizes", [("dense", [1, 2, 3, 5, 8, 16, 31, 32, 33, 47, 63, 65, 80, 88]),
                                        ("pow2", [64, 128, 256, 512, 1024, 2048, 4096, 8192])])
def test_gpu_field_and_reductions_equal_the_cpu_twin_bit_for_bit(oracle, kind, sizes):
    from proton_b200.api import Context
    rng = np.random.default_rng(5 if kind == "dense" else 6)
    batch = synth_batch(rng, sizes)
    critical = 40e6
    ctx = Context(0)
    try:
        plan = ctx.pack(batch)
        got = plan.eval(TS, critical=critical)
        for l in range(batch.num_lines):
            want = _twin_of(oracle, plan, batch, l, kind, TS)
            g = got.line(l)
            same = g.view(np.uint64) == want.view(np.uint64)
            assert same.all(), (f"{kind} N={batch.num_nodes[l]}: {np.count_nonzero(~same)} of {same.size} values differ, "
                                f"worst {np.abs(g - want).max() / np.abs(want).max():.3e}")
            assert np.array_equal(got.max_stress[l], want.max(axis=1))
            assert np.array_equal(got.argmax[l], want.argmax(axis=1))
            assert np.array_equal(got.flag[l].astype(bool), want.max(axis=1) > critical)
        plan.close()
    finally:
        ctx.close()


def _class_of(n):
    """(kind, lr, rows) of a non-power-of-two line above the dense limit, from the library's own classify()."""
    from proton_b200._lib import class_of
    kind, lr, rows = class_of(n)
    assert kind in ("fft_blue", "fft_mix"), (n, kind)
    return ("mix", lr, rows) if kind == "fft_mix" else ("blue", 0, 0)


# one length just below the capacity of every mixed class (rows x 2^lr >= 2N - 1), plus lengths that stay on 2^p
MIX_CLASSES = [(7, 5), (7, 6), (8, 5), (8, 6), (8, 7), (8, 9), (8, 11), (8, 14), (9, 5), (9, 6), (9, 9), (9, 10), (9, 11),
               (9, 12), (9, 14), (9, 16)]
MIX_SIZES = [(rows << lr) // 2 - 3 for lr, rows in MIX_CLASSES]


@pytest.mark.parametrize("sizes", [[100, 129, 200, 255, 400, 511, 1000, 2047],                # P = 2^p
                                   [300, 330, 360, 600, 700, 1100, 1300, 1500, 2500, 3000], MIX_SIZES])
def test_bluestein_classes_equal_the_cpu_twin_bit_for_bit(oracle, sizes):
    """In-CTA Bluestein with power-of-two and with mixed-radix lengths (5 to 16 rows of 2^k points)."""
    from proton_b200.api import Context
    rng = np.random.default_rng(8)
    batch = synth_batch(rng, sizes)
    ctx = Context(0)
    try:
        plan = ctx.pack(batch)
        got = plan.eval(TS, critical=40e6)
        kinds = set()
        for l, n in enumerate(sizes):
            kind, lr, rows = _class_of(n)
            kinds.add((kind, lr, rows))
            t = {k: plan.debug_table(l, k) for k in ("lambda", "W", "pre", "post", "hhat")}
            twl = plan.debug_table(l, "twl") if kind == "mix" else None
            want = oracle.twin_bluestein(n, float(batch.a_coeff[l]), t["lambda"], t["W"], t["pre"], t["post"], t["hhat"], TS,
                                         twl=twl, lr=lr, rows=rows)
            g = got.line(l)
            same = g.view(np.uint64) == want.view(np.uint64)
            assert same.all(), (f"N={n} {kind} lr={lr} rows={rows}: {np.count_nonzero(~same)} of {same.size} values differ, "
                                f"worst {np.abs(g - want).max() / np.abs(want).max():.3e}")
            assert np.array_equal(got.max_stress[l], want.max(axis=1)) and np.array_equal(got.argmax[l], want.argmax(axis=1))
        assert len(kinds) == 1 or len(kinds) >= 6          # the other lists go through the mixed-radix classes
        if sizes is MIX_SIZES:
            assert len(kinds) == len(MIX_SIZES), kinds
        plan.close()
    finally:
        ctx.close()


def test_long_lines_equal_the_cpu_twin_bit_for_bit(oracle):
    """The three-kernel long path (P = P1 x 4096 through the L2-resident scratch): the headline workload's lines."""
    from proton_b200.api import Context
    rng = np.random.default_rng(9)
    sizes = [4097, 5000, 6145, 9000, 12001, 19087, 25564, 33000, 65535]
    batch = synth_batch(rng, sizes)
    ts = TS[:3]
    ctx = Context(0)
    try:
        plan = ctx.pack(batch)
        got = plan.eval(ts, critical=40e6)
        for l, n in enumerate(sizes):
            t = {k: plan.debug_table(l, k) for k in ("lambda", "W", "pre", "post", "hhat", "twl")}
            p1 = len(t["hhat"]) // (2 * 4096)
            want = oracle.twin_bluestein(n, float(batch.a_coeff[l]), t["lambda"], t["W"], t["pre"], t["post"], t["hhat"], ts,
                                         twl=t["twl"], lr=12, rows=p1, long_path=True)
            g = got.line(l)
            same = g.view(np.uint64) == want.view(np.uint64)
            assert same.all(), (f"N={n} P1={p1}: {np.count_nonzero(~same)} of {same.size} values differ, "
                                f"worst {np.abs(g - want).max() / np.abs(want).max():.3e}")
            assert np.array_equal(got.max_stress[l], want.max(axis=1)) and np.array_equal(got.argmax[l], want.argmax(axis=1))
        plan.close()
    finally:
        ctx.close()


def test_long_power_of_two_lines_equal_the_cpu_twin_bit_for_bit(oracle):
    """N = 16384 ... 131072: scatter kernel, inverse-only row pass, column pass without chirps — the last class."""
    from proton_b200.api import Context
    rng = np.random.default_rng(10)
    sizes = [16384, 32768, 65536, 131072]
    batch = synth_batch(rng, sizes)
    ts = TS[:3]
    ctx = Context(0)
    try:
        plan = ctx.pack(batch)
        got = plan.eval(ts, critical=40e6)
        for l, n in enumerate(sizes):
            t = {k: plan.debug_table(l, k) for k in ("lambda", "W", "pre", "twl")}
            want = oracle.twin_long_pow2(n, float(batch.a_coeff[l]), t["lambda"], t["W"], t["pre"], t["twl"], ts)
            g = got.line(l)
            same = g.view(np.uint64) == want.view(np.uint64)
            assert same.all(), (f"N={n}: {np.count_nonzero(~same)} of {same.size} values differ, "
                                f"worst {np.abs(g - want).max() / np.abs(want).max():.3e}")
            assert np.array_equal(got.max_stress[l], want.max(axis=1)) and np.array_equal(got.argmax[l], want.argmax(axis=1))
        plan.close()
    finally:
        ctx.close()

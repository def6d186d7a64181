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

"""GPU test of proton_em_eval_line_noalloc, the entry point for callers that may not allocate device memory
(for example a host program whose own pool owns all of it, or a plug-in invoked where a cudaMalloc would stall the
GPU).  The entry point keeps its working memory in a fixed array inside the library, shared by all calls, so the
test makes two calls in a row — the second with a different, shorter line, so anything left over from the first
would show — and compares each with the regular context/plan path called through the same C ABI.

Collected last on purpose (file name): if it fails, `pytest -x` has already run every other test.
"""
import ctypes

import numpy as np
import pytest

from gpu_common import synth_line

pytestmark = pytest.mark.gpu

TOL = 1e-9   # norm-wise per timestamp, the tolerance of the other parity tests


def _call_noalloc(lib, torch, li, ts, critical):
    dev = torch.device("cuda:0")
    tt = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    rows, cols, vals = tt(li.rows, np.int32), tt(li.cols, np.int32), tt(li.vals, np.float64)
    u, d_ts = tt(li.curden, np.float64), tt(ts, np.float64)
    N, T = li.num_nodes, len(ts)
    stress = torch.full((T, N), float("nan"), dtype=torch.float64, device=dev)
    mx = torch.empty(T, dtype=torch.float64, device=dev)
    amx = torch.empty(T, dtype=torch.int32, device=dev)
    flag = torch.empty(T, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    rc = lib.proton_em_eval_line_noalloc(0, None, N, float(li.a_coeff), len(li.rows), rows.data_ptr(), cols.data_ptr(),
                                         vals.data_ptr(), len(li.curden), u.data_ptr(), d_ts.data_ptr(), T,
                                         float(critical), stress.data_ptr(), mx.data_ptr(), amx.data_ptr(),
                                         flag.data_ptr())
    return rc, stress.cpu().numpy(), mx.cpu().numpy(), amx.cpu().numpy(), flag.cpu().numpy()


def test_noalloc_entry_point_matches_the_regular_path_twice_in_a_row():
    import torch
    from proton_b200 import _lib
    from proton_b200.api import Context, LineBatch
    lib = _lib.load()
    rng = np.random.default_rng(77)
    first = synth_line(rng, 1500, 6)
    second = synth_line(rng, 333, 3)          # shorter: stale weights of the first call must not leak in
    ts1 = np.array([1e6, 3e7, 6.38e8])
    ts2 = np.array([2e6, 6.38e8])
    critical = 1e7
    ctx = Context(0)
    try:
        for li, ts in ((first, ts1), (second, ts2)):
            rc, stress, mx, amx, flag = _call_noalloc(lib, torch, li, ts, critical)
            assert rc == 0, lib.proton_em_last_error().decode()
            want = ctx.analyze_host(LineBatch.from_lines([li]), ts, critical=critical)
            w = want.line(0)
            scale = np.abs(w).max(axis=1)
            assert np.all(np.abs(stress - w).max(axis=1) <= TOL * scale)
            assert np.all(np.abs(mx - want.max_stress[0]) <= TOL * scale)
            assert np.array_equal(mx, stress.max(axis=1)) and np.array_equal(amx, stress.argmax(axis=1))
            assert np.array_equal(flag.astype(bool), mx > critical)
    finally:
        ctx.close()


def test_noalloc_entry_point_rejects_what_does_not_fit():
    from proton_b200 import _lib
    lib = _lib.load()
    rc = lib.proton_em_eval_line_noalloc(0, None, 4097, 1e-9, 0, None, None, None, 0, None, None, 1, 0.0, None, None,
                                         None, None)
    assert rc == _lib.EINVAL and b"4096" in lib.proton_em_last_error()
    rc = lib.proton_em_eval_line_noalloc(0, None, 10, 1e-9, 5000, None, None, None, 1, None, None, 1, 0.0, None, None,
                                         None, None)
    assert rc == _lib.EINVAL and b"4096" in lib.proton_em_last_error()

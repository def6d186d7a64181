"""GPU test of proton_em_eval_line_noalloc, the entry point for callers that may not allocate device memory
(for example a host program whose own pool owns all of it, or a plug-in invoked where a cudaMalloc would stall the
GPU).  The entry point keeps its working memory in a fixed array inside the library, shared by all calls, so the
test makes two calls in a row — the second with a different, shorter line, so anything left over from the first
would show — and compares each with the regular context/plan path called through the same C ABI and with the CPU
oracle.  The second call runs on a stream of the caller's own, and the free device memory reported by the driver
must be the same before and after it: that is the property the entry point exists for.

Collected last on purpose (file name): if it fails, `pytest -x` has already run every other test.
"""
import ctypes

import numpy as np
import pytest

from gpu_common import synth_line

pytestmark = pytest.mark.gpu

TOL = 1e-9   # norm-wise per timestamp, the tolerance of the other parity tests


def _call_noalloc(lib, torch, li, ts, critical, stream=None, check_free_memory=False):
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
    free_before = torch.cuda.mem_get_info(dev)[0]
    rc = lib.proton_em_eval_line_noalloc(0, ctypes.c_void_p(stream.cuda_stream) if stream is not None else None, N,
                                         float(li.a_coeff), len(li.rows), rows.data_ptr(), cols.data_ptr(),
                                         vals.data_ptr(), len(li.curden), u.data_ptr(), d_ts.data_ptr(), T,
                                         float(critical), stress.data_ptr(), mx.data_ptr(), amx.data_ptr(),
                                         flag.data_ptr())
    free_after = torch.cuda.mem_get_info(dev)[0]
    if check_free_memory:        # (the first call on a device may load the module lazily; the second may not take a byte)
        assert free_after == free_before, f"free device memory changed by {free_before - free_after} bytes across the call"
    return rc, stress.cpu().numpy(), mx.cpu().numpy(), amx.cpu().numpy(), flag.cpu().numpy()


def test_noalloc_entry_point_matches_the_regular_path_twice_in_a_row(oracle):
    import torch
    from proton_b200 import _lib
    from proton_b200.api import Context, LineBatch
    from gpu_common import oracle_batch
    lib = _lib.load()
    rng = np.random.default_rng(77)
    first = synth_line(rng, 1500, 6)
    second = synth_line(rng, 333, 3)          # shorter: stale weights of the first call must not leak in
    ts1 = np.array([1e6, 3e7, 6.38e8])
    ts2 = np.array([2e6, 6.38e8])
    critical = 1e7
    ctx = Context(0)
    own_stream = torch.cuda.Stream(device=torch.device("cuda:0"))
    with torch.cuda.stream(own_stream):       # first use of a stream may make the driver set things up: do it now
        torch.zeros(8, device="cuda:0").sum().item()
    try:
        for call, (li, ts) in enumerate(((first, ts1), (second, ts2))):
            rc, stress, mx, amx, flag = _call_noalloc(lib, torch, li, ts, critical,
                                                      stream=own_stream if call == 1 else None,
                                                      check_free_memory=call == 1)
            assert rc == 0, lib.proton_em_last_error().decode()
            batch = LineBatch.from_lines([li])
            want = ctx.analyze_host(batch, ts, critical=critical)
            ws, wm, wa = oracle_batch(oracle, batch, ts)
            for w, w_mx in ((want.line(0), want.max_stress[0]), (ws.reshape(len(ts), li.num_nodes), wm[0])):
                scale = np.abs(w).max(axis=1)
                assert np.all(np.abs(stress - w).max(axis=1) <= TOL * scale)
                assert np.all(np.abs(mx - w_mx) <= TOL * scale)
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

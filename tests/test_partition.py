"""Multi-GPU host logic on CPU: the segment-balanced split and the gather of per-line reductions (world_size 2,
gloo) — the only exchange step of the path."""
import os
import socket

import numpy as np
import pytest

from proton_b200.workloads import partition_lines, synthetic_ragged


def test_partition_is_a_balanced_cover():
    rng = np.random.default_rng(0)
    for world in (1, 2, 4, 8):
        for n in (3, 50, 5000, 300000):
            nn = np.exp(rng.uniform(0, np.log(20000), n)).astype(np.int64) + 1
            parts = partition_lines(nn, world)
            allidx = np.sort(np.concatenate(parts))
            assert np.array_equal(allidx, np.arange(n))
            loads = np.array([nn[p].sum() for p in parts], dtype=float)
            if n >= 50 * world:
                assert loads.max() / loads.mean() < 1.05
            for p in parts:
                assert np.all(np.diff(p) > 0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, q):
    """What bench.py does with N > 1, on CPU: every rank evaluates its shard (here: with the oracle), the three per-line
    reductions are all-gathered, and rank 0 checks lines owned by the other rank against the oracle through the same
    bench.check_gathered that runs on the GPUs."""
    import sys
    from pathlib import Path
    import torch
    import torch.distributed as dist
    sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
    import bench
    from oracle import restatement as R
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    batch = synthetic_ragged(200, seed=5, nmin=1, nmax=300)
    parts = partition_lines(batch.num_nodes, world)
    mine = batch.subset(parts[rank])
    ts = np.array([1e6, 5e7, 6.38e8])
    T = len(ts)
    _, mx, amx = R.batch(mine.num_nodes, mine.a_coeff, mine.trip_off, mine.trip_row, mine.trip_col, mine.trip_val,
                         mine.u_off, mine.curden, ts, want_stress=False)
    flag = (mx > bench.CRITICAL).astype(np.uint8)
    Lmax = max(len(p) for p in parts)
    recv = []
    for arr, dt in ((mx, torch.float64), (amx, torch.int32), (flag, torch.uint8)):
        send = torch.zeros(Lmax * T, dtype=dt)
        send[: arr.size] = torch.from_numpy(np.ascontiguousarray(arr)).reshape(-1)
        out = torch.empty(world * Lmax * T, dtype=dt)
        dist.all_gather_into_tensor(out, send)
        recv.append(out.numpy())
    ok = True
    if rank == 0:
        good = bench.check_gathered(batch, parts, ts, recv, Lmax, T, per_rank=5)
        ok = good["ok"] and good["lines_checked"] == 5 * (world - 1)
        # a shard -> line mapping that is off by one line must be caught
        shifted = [np.roll(r.reshape(world, Lmax, T), 1, axis=1).reshape(-1) for r in recv]
        ok = ok and not bench.check_gathered(batch, parts, ts, shifted, Lmax, T, per_rank=5)["ok"]
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_gather_of_reductions_world_size_2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]

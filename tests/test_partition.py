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
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    batch = synthetic_ragged(200, seed=5, nmin=1, nmax=300)
    parts = partition_lines(batch.num_nodes, world)
    mine = parts[rank]
    T = 3
    # stand-in for the per-line reductions of this rank's shard: a function of the global line id
    local = torch.tensor(np.outer(mine + 1, np.arange(1, T + 1)).astype(np.float64))
    Lmax = max(len(p) for p in parts)
    send = torch.zeros(Lmax * T, dtype=torch.float64)
    send[: local.numel()] = local.reshape(-1)
    recv = torch.empty(world * Lmax * T, dtype=torch.float64)
    dist.all_gather_into_tensor(recv, send)
    full = np.zeros((batch.num_lines, T))
    for r in range(world):
        blk = recv[r * Lmax * T:(r + 1) * Lmax * T].numpy().reshape(Lmax, T)
        full[parts[r]] = blk[: len(parts[r])]
    ok = np.array_equal(full, np.outer(np.arange(batch.num_lines) + 1, np.arange(1, T + 1)))
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

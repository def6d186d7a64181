"""Workload builders for bench.py and the tests (synthetic data only — no network, no datasets).

* ibmpg1: the real line population of benchmarks/ibmpg1.spice as parsed by the reference's own SPICE parser
  (fixture tests/golden/ibmpg1_lines.npz, made by tools/gen_golden.py with DC currents from tools/dc_solve.py),
  discretised with the whole-grid driver's policy dx = 0.99 * shortest segment (PROTON_cli.py:1005-1012).
* ibmpg2-shaped: benchmarks/ibmpg2.spice is a missing blob, so BASELINE config 2 is run on ibmpg1's
  line population replicated `replicas` times with seeded jitter of segment lengths and currents (ibmpg2 has about
  7x the resistors of ibmpg1).
* synthetic ragged batches for BASELINE configs 3-5, generated vectorised.
"""
from __future__ import annotations

from pathlib import Path
from typing import List, Tuple

import numpy as np

from . import formulation as F
from .api import LineBatch

GOLDEN = Path(__file__).resolve().parents[1] / "tests" / "golden" / "ibmpg1_lines.npz"
TECH, TEMP, WIDTH = "CuDD", 378.0, 0.1
KAPPA_B = None


def ibmpg1_segments() -> List[Tuple[str, list]]:
    z = np.load(GOLDEN)
    names = [str(n) for n in z["names"]]
    texts = bytes(z["csv_blob"].tobytes()).decode().split("\x00")
    out = []
    for name, text in zip(names, texts):
        segs = []
        for line in text.splitlines():
            if line and line[0] == "R":
                c = line.split(",")
                segs.append((c[1], c[2], float(c[3]), float(c[4])))
        out.append((name, segs))
    return out


def ibmpg1_lines(max_nodes: int = 0) -> list:
    lines = []
    for name, segs in ibmpg1_segments():
        li = F.formulate_segments(segs, sp_step=F.analyze_policy_step(segs), technology=TECH, temperature=TEMP,
                                  givenWidth=WIDTH, name=name)
        if max_nodes and li.num_nodes > max_nodes:
            continue
        lines.append(li)
    return lines


def ibmpg2_shaped_lines(replicas: int = 7, seed: int = 2, max_nodes: int = 0) -> list:
    rng = np.random.default_rng(seed)
    base = ibmpg1_segments()
    lines = []
    for r in range(replicas):
        for name, segs in base:
            if r == 0:
                s2 = segs
            else:
                scale = rng.uniform(0.8, 1.25)
                cur = rng.lognormal(0.0, 0.5)
                s2 = [(a, b, float(max(1, round(length * scale))), i * cur) for (a, b, length, i) in segs]
            li = F.formulate_segments(s2, sp_step=F.analyze_policy_step(s2), technology=TECH, temperature=TEMP,
                                      givenWidth=WIDTH, name=f"{name}#{r}")
            if max_nodes and li.num_nodes > max_nodes:
                continue
            lines.append(li)
    return lines


_TABLE_CACHE = {}


def ibmpg1_table() -> F.SegmentTable:
    """ibmpg1's line population as a flat segment table (parsed once per process)."""
    if "ibmpg1" not in _TABLE_CACHE:
        _TABLE_CACHE["ibmpg1"] = F.SegmentTable.from_lines(ibmpg1_segments())
    return _TABLE_CACHE["ibmpg1"]


def ibmpg1_batch() -> LineBatch:
    """`LineBatch.from_lines(ibmpg1_lines())` through the vectorised discretiser (bit-identical, tests/test_formulation.py)."""
    return F.formulate_table(ibmpg1_table(), technology=TECH, temperature=TEMP, givenWidth=WIDTH)


def ibmpg2_shaped_batch(replicas: int = 7, seed: int = 2) -> LineBatch:
    """`LineBatch.from_lines(ibmpg2_shaped_lines(replicas, seed))`, vectorised: the same seeded jitter (one length scale
    and one current factor per replicated line, drawn in the same order) applied to the flat segment table."""
    rng = np.random.default_rng(seed)
    base = ibmpg1_table()
    L, K = base.num_lines, len(base.length_um)
    counts = np.diff(base.seg_off)
    lengths, currents = [base.length_um], [base.current]
    for r in range(1, replicas):
        draws = np.array([(rng.uniform(0.8, 1.25), rng.lognormal(0.0, 0.5)) for _ in range(L)])
        scale, cur = np.repeat(draws[:, 0], counts), np.repeat(draws[:, 1], counts)
        lengths.append(np.maximum(1.0, np.rint(base.length_um * scale)))
        currents.append(base.current * cur)
    seg_off = np.concatenate([[0]] + [base.seg_off[1:] + r * K for r in range(replicas)])
    names = [f"{n}#{r}" for r in range(replicas) for n in base.names]
    tab = F.SegmentTable(names, seg_off, np.concatenate(lengths), np.concatenate(currents),
                         np.tile(base.lcoord, (replicas, 1)), np.tile(base.rcoord, (replicas, 1)),
                         np.tile(base.named, replicas))
    return F.formulate_table(tab, technology=TECH, temperature=TEMP, givenWidth=WIDTH)


def synthetic_ragged(num_lines: int, seed: int = 0, nmin: int = 16, nmax: int = 4096, dist: str = "loguniform",
                     sizes=None) -> LineBatch:
    """N log-uniform in [nmin, nmax]; S = min(N-1, 1 + Poisson(4)); vias at sorted distinct
    interior nodes; j = float32(N(0, 1e11)) A/m^2; bcoef = 8.29e-14; a = 3.3586e-21 / dx^2, dx ~ U(1, 12) um."""
    rng = np.random.default_rng(seed)
    L = int(num_lines)
    if sizes is not None:
        N = np.asarray(sizes, dtype=np.int64)
        L = len(N)
    elif dist == "loguniform":
        N = np.floor(np.exp(rng.uniform(np.log(nmin), np.log(nmax + 1), L))).astype(np.int64)
    else:
        N = rng.integers(nmin, nmax + 1, L)
    N = np.clip(N, 1, None)
    S = np.minimum(np.maximum(N - 1, 1), 1 + rng.poisson(4.0, L)).astype(np.int64)
    S[N <= 2] = 1
    bcoef = 8.29e-14
    dx = rng.uniform(1.0, 12.0, L) * 1e-6
    a = 3.3586e-21 / (dx * dx)
    u_off = np.zeros(L + 1, dtype=np.int64)
    np.cumsum(S, out=u_off[1:])
    curden = rng.normal(0.0, 1e11, int(u_off[-1])).astype(np.float32).astype(np.float64)
    # triplets per line: N == 1 -> one entry (0, S-1, -b);  else 2*S entries
    ntrip = np.where(N == 1, 1, 2 * S)
    trip_off = np.zeros(L + 1, dtype=np.int64)
    np.cumsum(ntrip, out=trip_off[1:])
    rows = np.zeros(int(trip_off[-1]), dtype=np.int32)
    cols = np.zeros(int(trip_off[-1]), dtype=np.int32)
    vals = np.zeros(int(trip_off[-1]), dtype=np.float64)
    line_of_seg = np.repeat(np.arange(L), S)
    seg_in_line = np.arange(int(u_off[-1])) - u_off[line_of_seg]
    Ns, Ss = N[line_of_seg], S[line_of_seg]
    # start node of segment s: 0 for s = 0, else the via; distinct increasing interior nodes
    frac = rng.uniform(0.0, 1.0, int(u_off[-1]))
    order = np.lexsort((frac, line_of_seg))
    frac_sorted = frac[order]                      # sorted within each line
    span = np.maximum(Ns - 2 - (Ss - 1), 0)
    via = 1 + np.floor(span * frac_sorted).astype(np.int64) + (seg_in_line - 1)
    start = np.where(seg_in_line == 0, 0, via)
    end = np.empty_like(start)
    end[:-1] = start[1:]
    last = seg_in_line == Ss - 1
    end[last] = Ns[last] - 1
    multi = Ns > 1
    base = trip_off[line_of_seg] + 2 * seg_in_line
    # segment s contributes (+b at its start node, col s) and (-b at its end node, col s); rows sorted by
    # (row, col) come out as: (0,0,+b), (via_1,0,-b), (via_1,1,+b), ..., (N-1,S-1,-b)
    m = multi
    rows[base[m]] = start[m]
    cols[base[m]] = seg_in_line[m]
    vals[base[m]] = bcoef
    rows[base[m] + 1] = end[m]
    cols[base[m] + 1] = seg_in_line[m]
    vals[base[m] + 1] = -bcoef
    one = ~multi
    rows[trip_off[line_of_seg[one]]] = 0
    cols[trip_off[line_of_seg[one]]] = 0
    vals[trip_off[line_of_seg[one]]] = -bcoef
    return LineBatch(N.astype(np.int32), a, trip_off, rows, cols, vals, u_off, curden)


def config4_batch(target_nodes: int = 10_000_000, seed: int = 4) -> LineBatch:
    """BASELINE config 4: a 10 M-node power grid with the ibmpg length mix — the ibmpg2-shaped
    population, and further seeded draws from it (with replacement) until the node count is reached."""
    base = ibmpg2_shaped_batch(replicas=7, seed=2)
    rng = np.random.default_rng(seed)
    nn = base.num_nodes.astype(np.int64)
    extra = []
    total = int(nn.sum())
    while total < target_nodes:
        i = int(rng.integers(0, base.num_lines))
        if total + nn[i] > target_nodes + 30000:
            continue
        extra.append(i)
        total += int(nn[i])
    idx = np.concatenate([np.arange(base.num_lines), np.array(extra, dtype=np.int64)])
    return base.subset(idx)


def kernel_class(n: int) -> str:
    """Which kernel family evaluates a line of n nodes (classify() of csrc/proton_em.cu, asked from the library)."""
    from ._lib import class_of
    return class_of(n)[0]


def partition_lines(num_nodes: np.ndarray, world: int, T: int = 1) -> List[np.ndarray]:
    """Segment-balanced static split over `world` GPUs (BASELINE north_star): longest-processing-
    time-first on the cost N * (1 + 0.12 log2 N); returns, per rank, the sorted line indices it owns."""
    n = np.asarray(num_nodes, dtype=np.float64)
    cost = n * (1.0 + 0.12 * np.log2(np.maximum(n, 2.0)))
    order = np.argsort(-cost, kind="stable")
    loads = np.zeros(world)
    owner = np.empty(len(n), dtype=np.int64)
    if len(n) > 200000:
        # large batches: deal the sorted lines round-robin in snake order (LPT within 1e-3 for many small items)
        k = np.arange(len(n))
        rnd, pos = k // world, k % world
        owner[order] = np.where(rnd % 2 == 0, pos, world - 1 - pos)
    else:
        import heapq
        heap = [(0.0, r) for r in range(world)]
        heapq.heapify(heap)
        for i in order:
            load, r = heapq.heappop(heap)
            owner[i] = r
            heapq.heappush(heap, (load + cost[i], r))
    return [np.nonzero(owner == r)[0] for r in range(world)]

"""tools/dc_solve.py — fixture tooling (not on the product path).

DC operating point of an IBM power-grid SPICE deck, so that the reference's own SPICE parser
(spice_function_cli.py:94-115) can be fed a `DC_analysis_<bench>.txt`.  The reference normally gets
this file from bin/circuit_simulation (needs libblas.so.3, absent here) or from
benchmarks/DC_analyses/ (missing blobs, .MISSING_LARGE_BLOBS) — SURVEY.md F7, §8f-3.

Method: nodes joined by 0 V sources are merged (union-find), nodes tied to ground through a source
are Dirichlet nodes, the remaining conductance Laplacian is SPD and solved with SuperLU
(scipy.sparse.linalg.spsolve).  Resistor current = (V_left - V_right) / R, left/right as written in
the deck.  Deterministic; residual is checked against 1e-9 * |rhs|.
"""
from __future__ import annotations

import sys

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


def solve_deck(spice_path: str):
    res, vsrc, isrc = [], [], []
    with open(spice_path) as f:
        for line in f:
            w = line.split()
            if len(w) < 4:
                continue
            c = w[0][0]
            if c in "Rr":
                res.append((w[0], w[1], w[2], float(w[3])))
            elif c in "Vv":
                vsrc.append((w[1], w[2], float(w[3])))
            elif c in "Ii":
                isrc.append((w[1], w[2], float(w[3])))
    names = {}

    def nid(n):
        if n not in names:
            names[n] = len(names)
        return names[n]

    gnd = nid("0")
    for _, a, b, _ in res:
        nid(a), nid(b)
    for a, b, _ in vsrc:
        nid(a), nid(b)
    for a, b, _ in isrc:
        nid(a), nid(b)
    n = len(names)
    parent = np.arange(n)

    def find(x):
        while parent[x] != x:
            parent[x] = parent[parent[x]]
            x = parent[x]
        return x

    fixed = {}
    for a, b, v in vsrc:
        ia, ib = names[a], names[b]
        if ia != gnd and ib != gnd:
            if v != 0.0:
                raise ValueError("non-zero floating voltage source is not supported")
            ra, rb = find(ia), find(ib)
            if ra != rb:
                parent[ra] = rb
    for a, b, v in vsrc:
        ia, ib = names[a], names[b]
        if ib == gnd and ia != gnd:
            fixed[find(ia)] = v
        elif ia == gnd and ib != gnd:
            fixed[find(ib)] = -v
    root = np.array([find(i) for i in range(n)])
    volt_known = np.full(n, np.nan)
    volt_known[find(gnd)] = 0.0
    for r, v in fixed.items():
        volt_known[r] = v
    free_roots = np.unique(root[np.isnan(volt_known[root])])
    idx = -np.ones(n, dtype=np.int64)
    idx[free_roots] = np.arange(len(free_roots))
    m = len(free_roots)
    rhs = np.zeros(m)
    ri, ci, vi = [], [], []
    for _, a, b, r in res:
        ra, rb = root[names[a]], root[names[b]]
        if ra == rb:
            continue
        g = 1.0 / r
        ia, ib = idx[ra], idx[rb]
        if ia >= 0:
            ri.append(ia), ci.append(ia), vi.append(g)
            if ib >= 0:
                ri.append(ia), ci.append(ib), vi.append(-g)
            else:
                rhs[ia] += g * volt_known[rb]
        if ib >= 0:
            ri.append(ib), ci.append(ib), vi.append(g)
            if ia >= 0:
                ri.append(ib), ci.append(ia), vi.append(-g)
            else:
                rhs[ib] += g * volt_known[ra]
    for a, b, i in isrc:           # current i flows from node a through the source into node b
        ia, ib = idx[root[names[a]]], idx[root[names[b]]]
        if ia >= 0:
            rhs[ia] -= i
        if ib >= 0:
            rhs[ib] += i
    G = sp.csc_matrix((vi, (ri, ci)), shape=(m, m))
    x = spla.spsolve(G, rhs)
    resid = np.abs(G @ x - rhs).max()
    if not resid <= 1e-9 * max(1.0, np.abs(rhs).max()):
        raise RuntimeError(f"DC solve residual too large: {resid}")
    volt = volt_known[root].copy()
    free_mask = idx[root] >= 0
    volt[free_mask] = x[idx[root[free_mask]]]
    currents = {}
    for name, a, b, r in res:
        currents[name] = float((volt[names[a]] - volt[names[b]]) / r)
    return currents, resid


def write_dc_file(currents, path: str) -> None:
    """One line per element in a layout the reference parser accepts: it keeps lines that contain the
    word '|Current' and reads name = words[3] + words[4][:-1], amps = float(words[5])."""
    with open(path, "w") as f:
        for name, amps in currents.items():
            f.write(f"dc op |Current {name[0]} {name[1:]}: {amps!r}\n")


if __name__ == "__main__":
    cur, resid = solve_deck(sys.argv[1])
    write_dc_file(cur, sys.argv[2])
    print(f"{len(cur)} resistor currents, residual {resid:.3e}")

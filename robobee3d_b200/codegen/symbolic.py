"""Symbolic analysis of the UprightMPC2 KKT system (host-side, offline).

Plays the role OSQP-Python's codegen plays for the reference
(template_controllers.py:197-202 -> template/uprightmpc2/workspace.c): given the QP
sparsity it picks an elimination order for the quasi-definite KKT matrix
[[P+sigma I, A'],[A, -1/rho]] and derives the fill pattern of L.  The order is
our own (greedy minimum-degree with a structure-aware tie-break), not the
reference's AMD permutation (workspace.c:1347).
"""
import numpy as np


def umpc_pattern(N=3, ny=6, nu=3):
    """(rows, cols, kind) of A's structural non-zeros in CSC order, following
    initConstraint (template_controllers.py:28-63).  kind: ('c', +-1) constant,
    ('dt',), ('T',), ('s', i), ('b', i, j)."""
    nx, nc = N * (2 * ny + nu), 2 * N * ny + N
    ent = {}
    n1, n2, nc1, nc2 = N * ny, 2 * N * ny, N * ny, 2 * N * ny
    for k in range(N):
        for i in range(ny):
            ent[(k * ny + i, k * ny + i)] = ("c", -1.0)
            ent[(k * ny + i, n1 + k * ny + i)] = ("dt",)
            if k > 0:
                ent[(k * ny + i, (k - 1) * ny + i)] = ("c", 1.0)
            ent[(nc1 + k * ny + i, n1 + k * ny + i)] = ("c", -1.0)
            if k > 0:
                ent[(nc1 + k * ny + i, n1 + (k - 1) * ny + i)] = ("c", 1.0)
        for i in range(3):
            ent[(nc1 + k * ny + i, n2 + k * nu)] = ("s", i)
            ent[(nc1 + k * ny + 3 + i, n2 + k * nu + 1)] = ("b", i, 0)
            ent[(nc1 + k * ny + 3 + i, n2 + k * nu + 2)] = ("b", i, 1)
        if k > 1:
            for i in range(3):
                ent[(nc1 + k * ny + i, (k - 2) * ny + 3 + i)] = ("T",)
        ent[(nc2 + k, n2 + 3 * k)] = ("c", 1.0)
    keys = sorted(ent.keys(), key=lambda rc: (rc[1], rc[0]))
    return nx, nc, [(r, c, ent[(r, c)]) for r, c in keys]


def kkt_adjacency(nx, nc, entries):
    adj = [set() for _ in range(nx + nc)]
    for r, c, _ in entries:
        adj[c].add(nx + r)
        adj[nx + r].add(c)
    return adj


def min_degree_order(adj, seed=None):
    """Greedy minimum-degree elimination; returns (perm, fill_cols) where
    fill_cols[k] = sorted list of later-eliminated neighbours of perm[k]
    (the below-diagonal pattern of column k of L, in ORIGINAL indices)."""
    n = len(adj)
    adj = [set(a) for a in adj]
    alive = set(range(n))
    rng = np.random.default_rng(seed) if seed is not None else None
    perm, cols = [], []
    while alive:
        dmin = min(len(adj[v]) for v in alive)
        cand = sorted(v for v in alive if len(adj[v]) == dmin)
        v = cand[0] if rng is None else cand[rng.integers(len(cand))]
        nb = sorted(adj[v])
        perm.append(v)
        cols.append(nb)
        for a in nb:
            adj[a].discard(v)
            adj[a].update(x for x in nb if x != a)
        alive.discard(v)
    return perm, cols


def best_order(adj, tries=200):
    best = None
    for s in [None] + list(range(tries)):
        perm, cols = min_degree_order(adj, s)
        nnz = sum(len(c) for c in cols)
        if best is None or nnz < best[0]:
            best = (nnz, perm, cols)
    return best

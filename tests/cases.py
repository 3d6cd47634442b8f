"""Seeded synthetic graphs / features shared by the golden generator, the oracle tests and
the GPU parity tests. Everything here is numpy; no reference, oracle or product imports.

Mirrors the reference's own test inputs:
  * kernel_graph():      the fixed 20-node graph of tests/compute/test_kernel.py:338-351
  * generate_feature():  tests/compute/test_kernel.py:33-73 (np.random.seed(31), uniform(-1,1)),
                         with smaller D1..D4 so committed fixtures stay small
  * div/add/sub rescaling of operands: test_kernel.py:204-222
"""
import itertools

import numpy as np

D1, D2, D3, D4 = 2, 3, 2, 3
TARGETS = {"u": 0, "v": 1, "e": 2}
OPS = ["add", "sub", "mul", "div", "dot"]
REDUCERS = ["sum", "max", "min", "prod", "none"]


def kernel_graph():
    """(num_nodes, src, dst) of test_kernel.py:338-351 in edge-id order."""
    n = 20
    src = list(range(n))
    dst = list(range(n))
    for i in range(2, 18):
        src += [0, 1, i, i]
        dst += [i, i, 18, 19]
    src += [18, 18, 19, 19]
    dst += [0, 1, 0, 1]
    return n, np.array(src, np.int64), np.array(dst, np.int64)


def powerlaw_graph(n_src, n_dst, n_edges, alpha, seed):
    """SURVEY 8(d) generator: endpoints drawn with probability ~ (rank+1)^-alpha under a
    seeded permutation. Returns int64 (src, dst) in edge-id order."""
    rng = np.random.default_rng(seed)

    def draw(n):
        w = (np.arange(n, dtype=np.float64) + 1.0) ** (-alpha)
        cdf = np.cumsum(w)
        cdf /= cdf[-1]
        perm = rng.permutation(n)
        return perm[np.minimum(np.searchsorted(cdf, rng.random(n_edges)), n - 1)]

    return draw(n_src).astype(np.int64), draw(n_dst).astype(np.int64)


def generate_feature(nv, ne, broadcast="none", binary_op="none"):
    rs = np.random.RandomState(31)
    full = (D1, D2, D3, D4) if binary_op == "dot" else (D1, D2, D3)
    small = (D2, 1, D4) if binary_op == "dot" else (D2, 1)
    shp = {"u": full, "e": full, "v": full}
    if broadcast in shp:
        shp[broadcast] = small
    u = rs.uniform(-1, 1, (nv,) + shp["u"])
    e = rs.uniform(-1, 1, (ne,) + shp["e"])
    v = rs.uniform(-1, 1, (nv,) + shp["v"])
    return u.astype(np.float32), v.astype(np.float32), e.astype(np.float32)


def binary_case_inputs(nv, ne, lhs, rhs, op, broadcast):
    """Operand tensors for one test_all_binary_builtins case (test_kernel.py:201-222)."""
    hu, hv, he = generate_feature(nv, ne, broadcast, op)
    feats = {"u": hu, "v": hv, "e": he}
    if op == "div":
        feats[rhs] = (feats[rhs] + 3) / 2
    if op in ("add", "sub"):
        feats = {k: x / 2 for k, x in feats.items()}
    return feats[lhs].astype(np.float32), feats[rhs].astype(np.float32)


def binary_cases():
    """Every (lhs, rhs, op, reducer, broadcast) of test_kernel.py:353-362 (partial handled
    separately through id mappings)."""
    for lhs, rhs in itertools.product("uve", "uve"):
        if lhs == rhs:
            continue
        for op in OPS:
            for red in REDUCERS:
                for bc in ["none", lhs, rhs]:
                    yield lhs, rhs, op, red, bc


def case_name(lhs, rhs, op, red, bc):
    return "%s_%s_%s__%s__bc_%s" % (lhs, op, rhs, red, bc)


def grad_out_for(shape, seed=7):
    return np.random.RandomState(seed).uniform(-1, 1, shape).astype(np.float32)


# Mid-size cases: (name, n_src, n_dst, n_edges, alpha, seed)
MID_GRAPHS = [
    ("square", 300, 300, 6000, 0.6, 11),
    ("bipartite", 200, 350, 5000, 0.4, 12),   # UnitGraph-style num_src != num_dst
    ("sparse", 400, 400, 500, 0.0, 13),       # many zero-degree rows
]

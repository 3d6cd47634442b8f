"""Seeded graphs for the partition tests (shared with tests/golden/make_golden_partition.py)."""
import numpy as np

import cases


def halo_case():
    n = 300
    src, dst = cases.powerlaw_graph(n, n, 6000, 0.6, 11)
    part = np.random.RandomState(3).randint(0, 3, n).astype(np.int64)
    return n, src, dst, part


def sbm(n, e, blocks, p_in, seed):
    rs = np.random.RandomState(seed)
    bsz = n // blocks
    sb = rs.randint(0, blocks, e)
    db = np.where(rs.rand(e) < p_in, sb, rs.randint(0, blocks, e))
    return sb * bsz + rs.randint(0, bsz, e), db * bsz + rs.randint(0, bsz, e)


def grid(side):
    idx = np.arange(side * side).reshape(side, side)
    return (np.concatenate([idx[:, :-1].ravel(), idx[:-1].ravel()]),
            np.concatenate([idx[:, 1:].ravel(), idx[1:].ravel()]))


def quality_cases():
    """name -> (num_nodes, src, dst, k)."""
    out = {}
    s, d = sbm(4000, 40000, 8, 0.85, 1)
    out["sbm8_k8"] = (4000, s, d, 8)
    s, d = sbm(20000, 200000, 64, 0.85, 1)
    out["sbm64_k8"] = (20000, s, d, 8)
    s, d = sbm(20000, 200000, 64, 0.85, 1)
    out["sbm64_k3"] = (20000, s, d, 3)
    s, d = grid(100)
    out["grid100_k4"] = (10000, s, d, 4)
    s, d = cases.powerlaw_graph(5000, 5000, 60000, 0.6, 9)
    out["powerlaw_k4"] = (5000, s, d, 4)
    return out

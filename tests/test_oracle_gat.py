"""Pins the oracle's fused-GAT restatement to the builtin composition the reference itself uses
as its check (examples/pytorch/gat/fused_gat_unit_test.py:36-52), run through the pinned oracle
ops: u_add_v -> leaky_relu -> exp -> copy_e+sum -> e_div_v -> u_mul_e+sum (no max-subtraction)."""
import numpy as np

import cases
from oracle import oracle as O


def composition(g, ft, el, er, slope):
    n_dst, ne = g.num_dst, g.num_edges
    e = O.binary_op_reduce("none", "add", g, 0, 1, el, er, ne)
    e = np.where(e > 0, e, e * np.float32(slope)).astype(np.float32)
    ex = np.exp(e).astype(np.float32)
    s = O.copy_reduce("sum", g, 2, ex, n_dst)
    a = O.binary_op_reduce("none", "div", g, 2, 1, ex, s, ne)
    ret = O.binary_op_reduce("sum", "mul", g, 0, 2, ft, a, n_dst)
    return s, ex, ret


def test_fused_gat_oracle_matches_builtin_composition():
    for (name, ns, nd, ne, alpha, seed) in cases.MID_GRAPHS[:2]:
        src, dst = cases.powerlaw_graph(ns, nd, ne, alpha, seed)
        if ns != nd:
            continue
        g = O.Graph(ns, nd, src, dst)
        rs = np.random.RandomState(seed)
        for H, F in ((4, 8), (1, 16), (8, 1), (3, 5)):
            ft = rs.uniform(-1, 1, (ns, H, F)).astype(np.float32)
            el = rs.uniform(-1, 1, (ns, H, 1)).astype(np.float32)
            er = rs.uniform(-1, 1, (nd, H, 1)).astype(np.float32)
            s, ex, ret = O.fused_gat_forward(g, ft, el, er, 0.2)
            s2, ex2, ret2 = composition(g, ft, el, er, 0.2)
            np.testing.assert_allclose(ex, ex2, rtol=2e-6)
            np.testing.assert_allclose(s, s2, rtol=1e-5)
            np.testing.assert_allclose(ret, ret2, rtol=1e-4, atol=1e-5)


def test_fused_gat_oracle_backward_matches_autograd_of_composition():
    import torch
    name, ns, nd, ne, alpha, seed = cases.MID_GRAPHS[0]
    src, dst = cases.powerlaw_graph(ns, nd, ne, alpha, seed)
    g = O.Graph(ns, nd, src, dst)
    rs = np.random.RandomState(3)
    H, F = 4, 8
    ft = rs.uniform(-1, 1, (ns, H, F)).astype(np.float32)
    el = rs.uniform(-1, 1, (ns, H, 1)).astype(np.float32)
    er = rs.uniform(-1, 1, (nd, H, 1)).astype(np.float32)
    gout = rs.uniform(-1, 1, (nd, H, F)).astype(np.float32)
    s, ex, ret = O.fused_gat_forward(g, ft, el, er, 0.2)
    gft, gel, ger = O.fused_gat_backward(g, ft, el, er, s, ex, ret, gout, 0.2)
    # dense float64 autograd of the same composition
    tft, tel, ter = [torch.tensor(a, dtype=torch.float64, requires_grad=True) for a in (ft, el, er)]
    S, D = torch.tensor(src), torch.tensor(dst)
    e = torch.nn.functional.leaky_relu(tel[S] + ter[D], 0.2).exp()
    ssum = torch.zeros((nd, H, 1), dtype=torch.float64).index_add(0, D, e)
    out = torch.zeros((nd, H, F), dtype=torch.float64).index_add(0, D, e / ssum[D] * tft[S])
    out.backward(torch.tensor(gout, dtype=torch.float64))
    np.testing.assert_allclose(gft, tft.grad.numpy(), rtol=1e-3, atol=1e-4)
    np.testing.assert_allclose(gel, tel.grad.numpy(), rtol=1e-3, atol=1e-4)
    np.testing.assert_allclose(ger, ter.grad.numpy(), rtol=1e-3, atol=1e-4)

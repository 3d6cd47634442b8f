"""The reference's own check for this path, tests/compute/test_kernel.py:196-362
(test_all_binary_builtins / test_copy_src_reduce / test_copy_edge_reduce): a builtin
(message, reduce) pair through the kernels must equal the same message written as a UDF, for every
(lhs, rhs) target pair, operator, reducer and broadcast pattern -- values and gradients. Here the
builtin pair is ONE kernel call (Graph.update_all fast path) and the UDF is a torch expression on
gathered endpoint features reduced with copy_e (the reference reduces UDF messages by degree
bucketing; the reduction semantics are the same)."""
import itertools

import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
FIELD = {"u": "src", "v": "dst", "e": "data"}


def _udf(lhs, rhs, op):
    def message(edges):
        a = getattr(edges, FIELD[lhs])["x" + lhs]
        b = getattr(edges, FIELD[rhs])["x" + rhs]
        while a.dim() < b.dim():      # batched broadcast, e.g. (E, 3, 1) op (E, 2, 3, 2):
            a = a.unsqueeze(1)        # test_kernel.py:236-241 does the same in its UDF
        while b.dim() < a.dim():
            b = b.unsqueeze(1)
        if op == "add":
            return {"m": a + b}
        if op == "sub":
            return {"m": a - b}
        if op == "mul":
            return {"m": a * b}
        if op == "div":
            return {"m": a / b}
        return {"m": (a * b).sum(-1, keepdim=True)}  # dot
    return message


@pytest.mark.parametrize("lhs,rhs", [p for p in itertools.product("uve", "uve") if p[0] != p[1]])
@pytest.mark.parametrize("op", ["add", "sub", "mul", "div", "dot"])
def test_binary_builtin_equals_udf(lhs, rhs, op):
    import dgl_hack_b200.function as fn
    from dgl_hack_b200.graph import Graph
    n, src, dst = cases.kernel_graph()
    ne = src.shape[0]
    for red, bc in itertools.product(["sum", "max", "min", "mean"], ["none", lhs, rhs]):
        a, b = cases.binary_case_inputs(n, ne, lhs, rhs, op, bc)
        outs = []
        for udf in (False, True):
            g = Graph(src, dst, num_nodes=n, device=DEV)
            ta = torch.from_numpy(a).to(DEV).requires_grad_(True)
            tb = torch.from_numpy(b).to(DEV).requires_grad_(True)
            for t, x in ((lhs, ta), (rhs, tb)):
                (g.edata if t == "e" else g.ndata)["x" + t] = x
            builtin = getattr(fn, "%s_%s_%s" % (lhs, op, rhs))("x" + lhs, "x" + rhs, "m")
            g.update_all(_udf(lhs, rhs, op) if udf else builtin, getattr(fn, red)("m", "r"))
            r = g.ndata["r"]
            if op == "dot" and not udf:
                r = r.unsqueeze(-1)      # the builtin drops the reduced axis (binary_reduce.cc)
            gout = torch.from_numpy(cases.grad_out_for(tuple(r.shape), 3)).to(DEV)
            has = torch.isfinite(r) & (r.abs() < 1e30)   # zero-degree rows hold +-FLT_MAX for max/min
            (torch.where(has, r, torch.zeros_like(r)) * gout).sum().backward()
            outs.append((r.detach(), ta.grad, tb.grad))
        for x, y, what in zip(outs[0], outs[1], ("out", "grad_lhs", "grad_rhs")):
            torch.testing.assert_close(x, y, rtol=2e-5, atol=2e-5,
                                       msg=lambda m: "%s %s_%s_%s %s bc=%s: %s" % (what, lhs, op, rhs, red, bc, m))


@pytest.mark.parametrize("target", ["u", "e"])
def test_copy_builtin_equals_udf(target):
    import dgl_hack_b200.function as fn
    from dgl_hack_b200.graph import Graph
    n, src, dst = cases.kernel_graph()
    hu, hv, he = cases.generate_feature(n, src.shape[0])
    x = hu if target == "u" else he
    for red in ("sum", "max", "min", "mean", "prod"):
        outs = []
        for udf in (False, True):
            g = Graph(src, dst, num_nodes=n, device=DEV)
            t = torch.from_numpy(x).to(DEV).requires_grad_(True)
            (g.edata if target == "e" else g.ndata)["x"] = t
            msg = (lambda edges: {"m": (edges.data if target == "e" else edges.src)["x"] * 1.0}) \
                if udf else (fn.copy_e if target == "e" else fn.copy_u)("x", "m")
            g.update_all(msg, getattr(fn, red)("m", "r"))
            r = g.ndata["r"]
            gout = torch.from_numpy(cases.grad_out_for(tuple(r.shape), 4)).to(DEV)
            (r * gout).sum().backward()
            outs.append((r.detach(), t.grad))
        torch.testing.assert_close(outs[0][0], outs[1][0], rtol=1e-6, atol=1e-6)
        torch.testing.assert_close(outs[0][1], outs[1][1], rtol=1e-5, atol=1e-5)

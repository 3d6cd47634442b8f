"""Host-side checks (no GPU compute): builtin function descriptors carry the reference's names /
target codes (python/dgl/function/message.py, reducer.py, base.py:7-21), and the nn modules expose
the reference modules' parameter names and shapes (read from the golden fixture keys)."""
import os

import numpy as np
import torch

import nn_cases

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "nn.npz")


def test_builtin_descriptors():
    import dgl_hack_b200.function as fn
    assert (fn.TargetCode.SRC, fn.TargetCode.DST, fn.TargetCode.EDGE) == (0, 1, 2)
    names = [a + "_" + op + "_" + b for a, b in ("uv", "ue", "vu", "ve", "eu", "ev")
             for op in ("add", "sub", "mul", "div", "dot")]
    assert len(names) == 30
    for n in names:
        m = getattr(fn, n)("a", "b", "o")
        assert m.name == n and (m.lhs_field, m.rhs_field, m.out_field) == ("a", "b", "o")
        assert fn.TargetCode.CODE2STR[m.lhs] == n[0] and fn.TargetCode.CODE2STR[m.rhs] == n[-1]
    assert fn.src_mul_edge("h", "w", "m").name == "u_mul_e"
    assert fn.copy_src("h", "m").target == 0 and fn.copy_edge("w", "m").target == 2
    assert fn.copy_u("h", "m").name == "copy_u" and fn.copy_e("h", "m").name == "copy_e"
    for r in ("sum", "max", "min", "mean", "prod"):
        f = getattr(fn, r)("m", "h")
        assert f.name == r and (f.msg_field, f.out_field) == ("m", "h")


def test_module_parameters_match_reference_names():
    import dgl_hack_b200.nn as nn_b200
    gold = np.load(GOLD)
    for idx, (name, cls, kwargs, _, _, _) in enumerate(nn_cases.CASES):
        mod = nn_cases.make_module(nn_b200, cls, kwargs)
        want = {k[len(name) + len("/gparam/"):]: gold[k].shape for k in gold.files
                if k.startswith(name + "/gparam/")}
        got = {n: tuple(p.shape) for n, p in mod.named_parameters()}
        assert got == want, name


def test_invalid_arguments():
    import pytest
    import dgl_hack_b200.nn as nn_b200
    with pytest.raises(Exception):
        nn_b200.GraphConv(4, 4, norm="left")
    with pytest.raises(NotImplementedError):
        nn_b200.SAGEConv(4, 4, "lstm")
    with pytest.raises(KeyError):
        nn_b200.SAGEConv(4, 4, "median")
    with pytest.raises(ValueError):
        nn_b200.RelGraphConv(10, 10, 3, "bdd", num_bases=3)
    with pytest.raises(ValueError):
        nn_b200.RelGraphConv(10, 10, 3, "lowrank")

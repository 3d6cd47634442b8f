"""Partial message passing on the GPU against the unmodified reference (tests/golden/partial.npz):
Graph.send_and_recv / Graph.pull through the id-mapped kernels, edge_softmax over an edge subset,
values and gradients; plus reference-style output maps (-1 for non-receivers, shorter than the
node count) at the kernel boundary."""
import os

import numpy as np
import pytest
import torch

import partial_cases as pc

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "partial.npz")
DEV = "cuda:0"


def make_graph(n, src, dst):
    from dgl_hack_b200.graph import Graph
    return Graph(src, dst, num_nodes=n, device=DEV)


def test_partial_message_passing_matches_reference():
    import dgl_hack_b200.function as fn
    from dgl_hack_b200.nn import edge_softmax
    gold = np.load(GOLD)
    res = pc.run(fn, make_graph, edge_softmax, DEV)
    assert sorted(res) == sorted(gold.files)
    for k in gold.files:
        scale = max(1.0, float(np.abs(gold[k]).max())) if np.abs(gold[k]).max() < 1e30 else 1.0
        np.testing.assert_allclose(res[k], gold[k], rtol=1e-5, atol=2e-6 * scale, err_msg=k)


def test_reference_style_out_map_is_safe():
    """_build_idx_map (runtime/scheduler.py) gives old->new ids with -1 for nodes that receive
    nothing and stops at max(recv) + 1: rows without edges must neither read nor write through it."""
    from dgl_hack_b200 import GraphIndex, kernel as K
    n = 1000
    src = np.array([5, 6, 7, 900], np.int64)
    dst = np.array([3, 3, 10, 10], np.int64)
    g = GraphIndex.from_coo(n, n, src, dst, DEV)
    x = torch.arange(n * 4, dtype=torch.float32, device=DEV).reshape(n, 4)
    out = torch.full((2, 4), 7.0, device=DEV)
    o2n = torch.full((11,), -1, dtype=torch.int32, device=DEV)   # length max(recv) + 1 only
    o2n[3], o2n[10] = 0, 1
    guard = torch.zeros(4096, device=DEV)                        # something to corrupt, if it would
    for red in ("sum", "max"):
        K.copy_reduce(red, g, 0, x, out, None, o2n)
        want0 = x[5] + x[6] if red == "sum" else torch.maximum(x[5], x[6])
        want1 = x[7] + x[900] if red == "sum" else torch.maximum(x[7], x[900])
        assert torch.equal(out[0], want0) and torch.equal(out[1], want1)
    gin = torch.empty_like(x)
    K.backward_copy_reduce("sum", g, 0, x, out, torch.ones(2, 4, device=DEV), gin, None, o2n)
    want = torch.zeros_like(x)
    want[[5, 6, 7, 900]] = 1
    assert torch.equal(gin, want)
    torch.cuda.synchronize()
    assert float(guard.abs().sum()) == 0

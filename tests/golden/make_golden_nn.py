"""Golden outputs of the UNMODIFIED reference's torch modules (dgl.nn.pytorch.{GraphConv, SAGEConv,
GATConv, RelGraphConv}, edge_softmax) and of the 2-layer Cora-shaped GCN (BASELINE configs[0]),
run on the reference's CPU path. Build container only:

    make -C oracle/ref_build -j8
    OMP_NUM_THREADS=1 DGLBACKEND=pytorch DGL_LIBRARY_PATH=oracle/_ref \
        PYTHONPATH=/root/reference/python python tests/golden/make_golden_nn.py

Writes tests/golden/nn.npz. Inputs and parameters are regenerated from tests/nn_cases.py seeds by
the tests; only the reference's outputs and gradients are stored.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import nn_cases  # noqa: E402

# the fork's GATConv.forward calls th.cuda.synchronize() for its timing prints (gatconv.py:146);
# that is not part of the arithmetic and has no meaning on the CPU build.
torch.cuda.synchronize = lambda *a, **k: None

import dgl  # noqa: E402
import dgl.nn.pytorch as refnn  # noqa: E402
from dgl.nn.pytorch.softmax import edge_softmax  # noqa: E402


def make_graph(ns, nd, src, dst, bip):
    if bip:
        return dgl.bipartite((src, dst), card=(ns, nd))
    g = dgl.DGLGraph()
    g.add_nodes(ns)
    g.add_edges(src, dst)
    return g


def main():
    store = {}
    for idx, case in enumerate(nn_cases.CASES):
        res = nn_cases.run_case(idx, refnn, make_graph, "cpu")
        for k, v in res.items():
            store["%s/%s" % (case[0], k)] = v
        print(case[0], res["out"].shape, float(np.abs(res["out"]).mean()))
    for idx in range(len(nn_cases.SOFTMAX_CASES)):
        graph, z, go = nn_cases.softmax_inputs(idx)
        g = make_graph(*nn_cases.graph_arrays(graph))
        zt = torch.from_numpy(z).requires_grad_(True)
        a = edge_softmax(g, zt)
        a.backward(torch.from_numpy(go))
        store["softmax%d/out" % idx] = a.detach().numpy()
        store["softmax%d/gz" % idx] = zt.grad.numpy()
    res = nn_cases.run_gcn_cora(refnn, make_graph, "cpu")
    for k, v in res.items():
        store["gcn_cora/" + k] = v
    print("gcn_cora loss", res["loss"])
    np.savez_compressed(os.path.join(HERE, "nn.npz"), **store)
    print("wrote %d arrays" % len(store))


if __name__ == "__main__":
    main()

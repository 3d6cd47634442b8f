"""Golden outputs of the UNMODIFIED reference for partial message passing: DGLGraph.send_and_recv,
DGLGraph.pull (id-mapped kernel calls, runtime/scheduler.py schedule_snr / schedule_pull) and
edge_softmax(graph, logits, eids). Build container only. Writes partial.npz."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import partial_cases as pc  # noqa: E402

import dgl  # noqa: E402
import dgl.function as fn  # noqa: E402
from dgl.nn.pytorch.softmax import edge_softmax  # noqa: E402


def make_graph(n, src, dst):
    g = dgl.DGLGraph()
    g.add_nodes(n)
    g.add_edges(src, dst)
    return g


if __name__ == "__main__":
    res = pc.run(fn, make_graph, edge_softmax, "cpu")
    np.savez_compressed(os.path.join(HERE, "partial.npz"), **res)
    print({k: (v.shape, float(np.abs(v).sum())) for k, v in res.items() if k.endswith("/out")})

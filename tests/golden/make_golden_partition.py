"""Golden outputs of the UNMODIFIED reference for the partition row:
`dgl.transform.partition_graph_with_halo(g, node_part, 1)` on the seeded "square" graph with a
seeded 3-way assignment, and the edge cuts METIS (`_CAPI_DGLMetisPartition` on
`to_bidirected(g)`) reaches on the seeded quality graphs of tests/test_partition_kway.py.
Build container only (same environment line as make_golden.py). Writes partition.npz."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402
import partition_cases as pc  # noqa: E402

import dgl  # noqa: E402
from dgl import transform as T, utils  # noqa: E402


def ref_graph(n, src, dst):
    g = dgl.DGLGraph()
    g.add_nodes(n)
    g.add_edges(src, dst)
    g.readonly()
    return g


def main():
    store = {}
    n, src, dst, part = pc.halo_case()
    g = ref_graph(n, src, dst)
    subs = T.partition_graph_with_halo(g, part, 1)
    for p, sg in subs.items():
        s, d = sg.all_edges(order="eid")
        store["halo/%d/induced_nodes" % p] = sg.parent_nid.numpy()
        store["halo/%d/induced_edges" % p] = sg.parent_eid.numpy()
        store["halo/%d/inner_node" % p] = sg.ndata["inner_node"].numpy()
        store["halo/%d/inner_edge" % p] = sg.edata["inner_edge"].numpy()
        store["halo/%d/src" % p] = s.numpy()
        store["halo/%d/dst" % p] = d.numpy()
    for name, (n, src, dst, k) in pc.quality_cases().items():
        g = ref_graph(n, src, dst)
        sym = T.to_bidirected(g, readonly=True)
        mp = utils.toindex(T._CAPI_DGLMetisPartition(sym._graph, k)).tonumpy()
        store["metis/%s/cut" % name] = np.int64((mp[src] != mp[dst]).sum())
        store["metis/%s/sizes" % name] = np.bincount(mp, minlength=k)
        print(name, int(store["metis/%s/cut" % name]), store["metis/%s/sizes" % name].tolist())
    np.savez_compressed(os.path.join(HERE, "partition.npz"), **store)


if __name__ == "__main__":
    main()

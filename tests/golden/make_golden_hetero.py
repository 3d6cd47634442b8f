"""Golden outputs of the UNMODIFIED reference's dgl.nn.pytorch.HeteroGraphConv over GraphConv
sub-modules (python/dgl/nn/pytorch/hetero.py:120-171, conv/graphconv.py:128-176) on the small
heterograph of tests/hetero_cases.py, CPU path. Build container only:

    make -C oracle/ref_build -j8
    OMP_NUM_THREADS=1 DGLBACKEND=pytorch DGL_LIBRARY_PATH=oracle/_ref \
        PYTHONPATH=/root/reference/python python tests/golden/make_golden_hetero.py

Writes tests/golden/hetero.npz (outputs and gradients only; inputs are regenerated from seeds).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import hetero_cases as H  # noqa: E402

import dgl  # noqa: E402
import dgl.nn.pytorch as refnn  # noqa: E402


def make_graph(edge_dict, num_nodes):
    return dgl.heterograph({k: (list(s), list(d)) for k, (s, d) in edge_dict.items()},
                           num_nodes_dict=num_nodes)


def main():
    store = {}
    for idx, case in enumerate(H.CASES):
        res, _ = H.run_case(idx, make_graph, refnn.GraphConv, refnn.HeteroGraphConv, torch, "cpu")
        for k, v in res.items():
            store["%s/%s" % (case[0], k)] = v
        print(case[0], sorted(k for k in res if k.startswith("out/")))
    np.savez_compressed(os.path.join(HERE, "hetero.npz"), **store)
    print("wrote %d arrays" % len(store))


if __name__ == "__main__":
    main()

"""Offline graph partitioning: the reference's tools/partition.py (:8-53) for this library.

    python -m dgl_hack_b200.partition_tool --data graph.npz -k 8 -m metis -o parts/

`graph.npz` holds `src`, `dst` (int64 COO, edge id = position), `num_nodes` and any number of
node-feature arrays with `num_nodes` rows. The reference loads / saves DGL's binary graph format
(dgl.data.utils.load_graphs / save_graphs, src/graph/graph_serialize.cc); that container is outside
this path, so parts are written as plain `.npz` files with the same content a saved part carries:
the 1-hop halo subgraph (`induced_nodes`, `induced_edges`, local `src` / `dst`), the `inner_node` /
`inner_edge` masks, `part_id` of every induced node and the node features copied from the parent
(`copy_from_parent`). `Partition.from_files` (partition.py) turns the files into a rank's
halo-exchange plan without touching the whole graph again.
"""
import argparse
import os

import numpy as np
import torch

from .partition import metis_partition_assignment, partition_graph_with_halo


def partition_to_files(data, num_parts, method, output, seed=0, balance="nodes"):
    src, dst, n = data["src"], data["dst"], int(data["num_nodes"])
    if method == "metis":
        node_part, cut = metis_partition_assignment(n, src, dst, num_parts, balance=balance,
                                                    seed=seed, return_cut=True)
    elif method == "random":
        node_part = torch.from_numpy(np.random.RandomState(seed).choice(num_parts, n))
    else:
        raise Exception("unknown partitioning method: " + method)
    parts = partition_graph_with_halo(n, src, dst, node_part, num_parts)
    feats = {k: v for k, v in data.items() if k not in ("src", "dst", "num_nodes")
             and getattr(v, "shape", ())[:1] == (n,)}
    os.makedirs(output, exist_ok=True)
    tot_inner = 0
    for pid, part in parts.items():
        inner_nodes = int(part["inner_node"].sum())
        inner_edges = int(part["inner_edge"].sum())
        print("part {} has {} nodes and {} edges. {} nodes and {} edges are inside the partition"
              .format(pid, part["induced_nodes"].numel(), part["induced_edges"].numel(),
                      inner_nodes, inner_edges))
        tot_inner += inner_edges
        arrays = {k: v.numpy() for k, v in part.items()}
        ids = arrays["induced_nodes"]
        for k, v in feats.items():
            arrays["feat/" + k] = np.asarray(v)[ids]
        np.savez(os.path.join(output, "%d.npz" % pid), num_parts=num_parts, num_nodes=n, **arrays)
    ne = int(np.asarray(src).shape[0])
    print("there are {} edges in the graph and {} edge cuts for {} partitions.".format(
        ne, ne - tot_inner, len(parts)))
    return node_part


def main():
    ap = argparse.ArgumentParser(description="Partition a graph")
    ap.add_argument("--data", required=True, help="input graph (.npz with src, dst, num_nodes)")
    ap.add_argument("-k", "--num-parts", required=True, type=int)
    ap.add_argument("--num-hops", type=int, default=1, help="only 1 hop of HALO nodes is supported")
    ap.add_argument("-m", "--method", required=True, help="random, metis")
    ap.add_argument("-o", "--output", required=True)
    ap.add_argument("--balance", default="nodes", choices=["nodes", "in_edges"])
    args = ap.parse_args()
    if args.num_hops != 1:
        raise Exception("the halo exchange of this library uses exactly one hop")
    partition_to_files(dict(np.load(args.data)), args.num_parts, args.method, args.output,
                       balance=args.balance)


if __name__ == "__main__":
    main()

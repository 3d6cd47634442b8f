"""Golden outputs of the UNMODIFIED reference for the mini-batch path: dgl.in_subgraph and
dgl.to_block on the seeded "square" graph (deterministic functions; sample_neighbors is random and
is tested on invariants). Build container only. Writes sampling.npz."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402

import dgl  # noqa: E402


def main():
    n = 300
    src, dst = cases.powerlaw_graph(n, n, 6000, 0.6, 11)
    g = dgl.graph((torch.from_numpy(src), torch.from_numpy(dst)), card=n)
    seeds = torch.from_numpy(np.random.RandomState(4).permutation(n)[:40].astype(np.int64))
    sub = dgl.in_subgraph(g, seeds)
    s, d = sub.all_edges(order="eid")
    store = {"seeds": seeds.numpy(), "in_sub/src": s.numpy(), "in_sub/dst": d.numpy(),
             "in_sub/eid": sub.edata[dgl.EID].numpy()}
    block = dgl.to_block(sub, seeds)
    bs, bd = block.all_edges(order="eid")
    store.update({"block/src_nid": block.srcdata[dgl.NID].numpy(),
                  "block/dst_nid": block.dstdata[dgl.NID].numpy(),
                  "block/src": bs.numpy(), "block/dst": bd.numpy()})
    np.savez_compressed(os.path.join(HERE, "sampling.npz"), **store)
    print({k: v.shape for k, v in store.items()})


if __name__ == "__main__":
    main()

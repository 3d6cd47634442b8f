"""Heterograph facade + batched per-relation message passing (BASELINE config 4).

The reference's heterograph is a set of UnitGraphs, one per canonical edge type, and
`dgl.nn.pytorch.HeteroGraphConv` (python/dgl/nn/pytorch/hetero.py:120-171) loops over them: R
module calls, R kernel launches that each write every destination row (R = 133 on the AM graph).
Here the relations additionally live in ONE typed graph over the concatenated node id space
(node type t occupies [offset[t], offset[t] + n_t)), so that

  * `relation_copy_reduce / relation_binary_reduce` run a builtin for ALL relations in one kernel
    launch on the relation-segment graph (GraphIndex.relation_segments): row s of the result is
    the reduction over relation seg_type[s] into node seg_dst[s]; and
  * `nn.HeteroGraphConv` of GraphConv modules with sum aggregation is one fused typed kernel
    (backend.rgcn_layer1: gather, per-relation weight, cross-relation sum).

Only what the kernels need of DGLHeteroGraph is mirrored (python/dgl/heterograph.py:
canonical_etypes, ntypes, number_of_nodes(ntype), g[stype, etype, dtype]).
"""
import torch

from . import backend as F
from .graph import Graph, bipartite
from .graph_index import GraphIndex


class HeteroGraph(object):
    """`dgl.heterograph(data_dict, num_nodes_dict)`: data_dict maps (stype, etype, dtype) to
    (src, dst) id arrays local to their node types."""

    def __init__(self, data_dict, num_nodes_dict, device="cuda:0"):
        self.device = torch.device(device)
        self.canonical_etypes = list(data_dict.keys())
        self.ntypes = sorted(num_nodes_dict.keys())
        self._num_nodes = {k: int(v) for k, v in num_nodes_dict.items()}
        self._edges = {}
        for key, (src, dst) in data_dict.items():
            src = torch.as_tensor(src, dtype=torch.int64).to(self.device)
            dst = torch.as_tensor(dst, dtype=torch.int64).to(self.device)
            self._edges[key] = (src, dst)
        self._rel_graphs = {}
        self._unified = None
        off, self.offsets = 0, {}
        for t in self.ntypes:
            self.offsets[t] = off
            off += self._num_nodes[t]
        self.total_nodes = off

    @property
    def etypes(self):
        return [e for _, e, _ in self.canonical_etypes]

    @property
    def srctypes(self):
        return self.ntypes

    @property
    def dsttypes(self):
        return self.ntypes

    def number_of_nodes(self, ntype):
        return self._num_nodes[ntype]

    def number_of_edges(self, key):
        return int(self._edges[key][0].numel())

    def __getitem__(self, key):
        """The relation's unit graph as a (bipartite) Graph, built on first use."""
        if key not in self._rel_graphs:
            stype, _, dtype = key
            src, dst = self._edges[key]
            if stype == dtype:
                g = Graph(src, dst, num_nodes=self._num_nodes[stype], device=self.device)
            else:
                g = bipartite(src, dst, self._num_nodes[stype], self._num_nodes[dtype],
                              device=self.device)
            self._rel_graphs[key] = g
        return self._rel_graphs[key]

    def unified(self):
        """(typed GraphIndex over the concatenated id space, edge offsets per relation): edge ids
        are relation-major in canonical_etypes order, relation id = position in that list."""
        if self._unified is None:
            srcs, dsts, tys, eoff, off = [], [], [], {}, 0
            for r, key in enumerate(self.canonical_etypes):
                stype, _, dtype = key
                s, d = self._edges[key]
                srcs.append(s + self.offsets[stype])
                dsts.append(d + self.offsets[dtype])
                tys.append(torch.full_like(s, r))
                eoff[key] = off
                off += int(s.numel())
            cat = lambda xs: torch.cat(xs) if xs else torch.empty(0, dtype=torch.int64,  # noqa: E731
                                                                  device=self.device)
            g = GraphIndex.from_coo_with_type(self.total_nodes, cat(srcs), cat(dsts), cat(tys),
                                              max(len(self.canonical_etypes), 1), self.device)
            self._unified = (g, eoff)
        return self._unified

    def cat_nodes(self, feats):
        """Node features of every type stacked in the unified id order (types without an entry
        contribute zero rows)."""
        some = next(iter(feats.values()))
        parts = [feats[t] if t in feats else
                 some.new_zeros((self._num_nodes[t],) + tuple(some.shape[1:])) for t in self.ntypes]
        return torch.cat(parts, 0)

    def split_nodes(self, x):
        return {t: x[self.offsets[t]: self.offsets[t] + self._num_nodes[t]] for t in self.ntypes}


def relation_copy_reduce(reducer, hg, feats):
    """`copy_src + reducer` over EVERY relation in one launch. feats: {ntype: [n_t, ...]} (same
    trailing shape). Returns (out [n_seg, ...], seg_node (unified id), seg_rel): out[s] = what the
    reference's `update_all(copy_src, reducer)` on relation seg_rel[s]'s unit graph leaves in
    destination row seg_node[s]."""
    g, _ = hg.unified()
    sg, seg_dst, seg_type = g.relation_segments()
    out = F.copy_reduce(reducer, sg, 0, hg.cat_nodes(feats), sg.num_dst)
    return out, seg_dst, seg_type


def relation_binary_reduce(reducer, op, hg, feats, edge_feats):
    """`u_<op>_e + reducer` over every relation in one launch; edge_feats: {canonical etype:
    [E_r, ...]} (stacked in relation-major edge id order)."""
    g, _ = hg.unified()
    sg, seg_dst, seg_type = g.relation_segments()
    ef = torch.cat([edge_feats[k] for k in hg.canonical_etypes], 0)
    out = F.binary_reduce(reducer, op, sg, 0, 2, hg.cat_nodes(feats), ef, sg.num_dst)
    return out, seg_dst, seg_type

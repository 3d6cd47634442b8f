"""`dgl.nn.pytorch.HeteroGraphConv` (python/dgl/nn/pytorch/hetero.py:14-171) over the B200 kernels.

Same constructor / forward contract (mods keyed by edge type name, cross-relation aggregate
'sum' | 'max' | 'min' | 'mean' | 'stack' or a callable, relations without edges or without an
input are skipped). Two execution paths with the same results:

  * generic: one sub-module call per relation on that relation's unit graph -- what the reference
    does (R launches; any sub-module);
  * fused (all participating sub-modules are GraphConv with a weight, no activation, equal
    in_feats <= out_feats with in * out <= 1024, aggregate == 'sum'): ONE typed kernel each way
    for all relations -- sum_r D_in,r^-1/2 A_r D_out,r^-1/2 X W_r is the RGCN layer-1 contraction
    (backend.rgcn_layer1) on the unified typed graph with the per-edge norm
    outdeg_r(u)^-1/2 * indeg_r(v)^-1/2; the biases of the relations reaching a node type are
    added afterwards. B200_HETERO_FUSED=0 forces the generic path.
"""
import os

import torch as th
from torch import nn

from .. import backend as B
from .._lib import DGLError
from .conv import GraphConv

__all__ = ["HeteroGraphConv"]


def _aggregate_fn(agg):
    """hetero.py:173-214 get_aggregate_fn."""
    if agg == "stack":
        return lambda xs, ntype: th.stack(xs, dim=1) if xs else None
    table = {"sum": lambda t: t.sum(0), "max": lambda t: t.max(0)[0],
             "min": lambda t: t.min(0)[0], "mean": lambda t: t.mean(0)}
    if agg not in table:
        raise DGLError('Invalid cross type aggregator. Must be one of "sum", "max", "min", "mean" '
                       'or "stack". But got "%s"' % agg)
    red = table[agg]
    return lambda xs, ntype: red(th.stack(xs, dim=0)) if xs else None


class HeteroGraphConv(nn.Module):
    def __init__(self, mods, aggregate="sum"):
        super(HeteroGraphConv, self).__init__()
        self.mods = nn.ModuleDict(mods)
        self._agg_name = aggregate if isinstance(aggregate, str) else None
        self.agg_fn = _aggregate_fn(aggregate) if isinstance(aggregate, str) else aggregate
        self.last_path = None   # "fused" | "generic" (introspection for tests / benches)

    # ---- fused path
    def _active(self, g, inputs):
        return [k for k in g.canonical_etypes
                if g.number_of_edges(k) > 0 and k[0] in inputs]

    def _fusable(self, g, inputs, mod_args, mod_kwargs):
        if os.environ.get("B200_HETERO_FUSED", "1") == "0" or self._agg_name != "sum":
            return False
        if isinstance(inputs, tuple) or mod_args or mod_kwargs:
            return False
        active = self._active(g, inputs)
        if not active:
            return False
        shapes = set()
        for _, etype, _ in active:
            m = self.mods[etype]
            if type(m) is not GraphConv or m.weight is None or m._activation is not None:
                return False
            shapes.add((m._in_feats, m._out_feats))
        if len(shapes) != 1:
            return False
        fin, fout = next(iter(shapes))
        ok_x = all(x.dim() == 2 and x.shape[1] == fin and x.is_cuda and x.dtype == th.float32
                   for x in inputs.values())
        return ok_x and fin <= fout and fin * fout <= 1024

    def _edge_norm(self, g, active):
        """Per unified edge id: the product of GraphConv's two degree scalings on the edge's OWN
        relation graph (graphconv.py:128-133, 165-172; degrees clamped to >= 1). Relations that do
        not take part get 0 (their edges contribute nothing). Cached per (graph, norms)."""
        key = tuple((k, self.mods[k[1]]._norm) for k in active)
        cache = g.__dict__.setdefault("_hetero_norm", {})
        if key not in cache:
            gi, eoff = g.unified()
            norm = th.zeros(gi.number_of_edges(), device=g.device)
            for k in active:
                src, dst = g._edges[k]
                kind = self.mods[k[1]]._norm
                w = th.ones(src.numel(), device=g.device)
                if kind != "none":
                    ind = th.bincount(dst, minlength=g.number_of_nodes(k[2])).float().clamp(min=1)
                    if kind == "both":
                        outd = th.bincount(src, minlength=g.number_of_nodes(k[0])).float().clamp(min=1)
                        w = th.pow(outd, -0.5)[src] * th.pow(ind, -0.5)[dst]
                    else:
                        w = (1.0 / ind)[dst]
                norm[eoff[k]: eoff[k] + src.numel()] = w
            cache.clear()
            cache[key] = norm.view(-1, 1)
        return cache[key]

    def _forward_fused(self, g, inputs):
        active = self._active(g, inputs)
        gi, _ = g.unified()
        some = self.mods[active[0][1]]
        zero_w = some.weight.new_zeros(some.weight.shape)
        live = set(active)
        weight = th.stack([self.mods[k[1]].weight if k in live else zero_w
                           for k in g.canonical_etypes])
        out = B.rgcn_layer1(gi, g.cat_nodes(inputs), weight, self._edge_norm(g, active))
        parts = g.split_nodes(out)
        rsts = {}
        for nty in g.ntypes:
            rels = [k for k in active if k[2] == nty]
            if not rels:
                continue
            h = parts[nty]
            biases = [self.mods[k[1]].bias for k in rels if self.mods[k[1]].bias is not None]
            if biases:
                h = h + th.stack(biases).sum(0)
            rsts[nty] = h
        return rsts

    # ---- generic path (hetero.py:120-171)
    def forward(self, g, inputs, mod_args=None, mod_kwargs=None):
        mod_args, mod_kwargs = mod_args or {}, mod_kwargs or {}
        if self._fusable(g, inputs, mod_args, mod_kwargs):
            self.last_path = "fused"
            return self._forward_fused(g, inputs)
        self.last_path = "generic"
        outputs = {nty: [] for nty in g.dsttypes}
        pair = isinstance(inputs, tuple)
        src_inputs, dst_inputs = inputs if pair else (inputs, inputs)
        for stype, etype, dtype in g.canonical_etypes:
            key = (stype, etype, dtype)
            if g.number_of_edges(key) == 0 or stype not in src_inputs:
                continue
            if pair and dtype not in dst_inputs:
                continue
            feat = (src_inputs[stype], dst_inputs[dtype]) if pair else src_inputs[stype]
            outputs[dtype].append(self.mods[etype](g[key], feat, *mod_args.get(etype, ()),
                                                   **mod_kwargs.get(etype, {})))
        rsts = {}
        for nty, alist in outputs.items():
            if alist:
                rsts[nty] = self.agg_fn(alist, nty)
        return rsts

"""edge_softmax (python/dgl/nn/pytorch/softmax.py:16-110).

a[e] = exp(z[e] - max_{e' into dst(e)} z[e']) / sum_{e' into dst(e)} exp(z[e'] - max), per
trailing feature column. Two forms:

* the reference's 4-kernel form (`copy_e+max`, `e_sub_v`, `copy_e+sum`, `e_div_v`, softmax.py:69-78;
  backward `copy_e+sum` + `e_mul_v`, :104-110) through `backend.copy_reduce / binary_reduce`;
* the default: ONE kernel each way (`kernel.edge_softmax_forward / _backward`, csrc/softmax.cu)
  that walks every destination row once -- max, exp, sum and the division happen in registers.
  Sums run in the same (source, edge id) order as the reference's single-threaded walk.

`B200_EDGE_SOFTMAX=builtin` selects the 4-kernel form.
"""
import os

import torch as th

from .. import backend as F
from .. import kernel as K
from ..function import TargetCode

__all__ = ["edge_softmax"]


def _gidx(graph):
    return graph.gidx if hasattr(graph, "gidx") else graph


class EdgeSoftmaxBuiltin(th.autograd.Function):
    @staticmethod
    def forward(ctx, gidx, score):
        n_nodes, n_edges = gidx.num_dst, gidx.number_of_edges()
        ctx.backward_cache = n_nodes, n_edges, gidx
        smax = F.copy_reduce("max", gidx, TargetCode.EDGE, score, n_nodes)
        out = F.binary_reduce("none", "sub", gidx, TargetCode.EDGE, TargetCode.DST, score, smax,
                              n_edges)
        out = th.exp(out)
        out_sum = F.copy_reduce("sum", gidx, TargetCode.EDGE, out, n_nodes)
        out = F.binary_reduce("none", "div", gidx, TargetCode.EDGE, TargetCode.DST, out, out_sum,
                              n_edges)
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        n_nodes, n_edges, gidx = ctx.backward_cache
        out, = ctx.saved_tensors
        sds = out * grad_out
        accum = F.copy_reduce("sum", gidx, TargetCode.EDGE, sds, n_nodes)
        out = F.binary_reduce("none", "mul", gidx, TargetCode.EDGE, TargetCode.DST, out, accum,
                              n_edges)
        return None, sds - out


class EdgeSoftmaxFused(th.autograd.Function):
    @staticmethod
    def forward(ctx, gidx, score):
        score = score.contiguous()
        out = th.empty_like(score)
        K.edge_softmax_forward(gidx, score, out)
        ctx.gidx = gidx
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        out, = ctx.saved_tensors
        grad_score = th.empty_like(out)
        K.edge_softmax_backward(ctx.gidx, out, grad_out.contiguous(), grad_score)
        return None, grad_score


def edge_softmax(graph, logits, eids=None):
    """graph: dgl_hack_b200.Graph or GraphIndex; logits [E, *, 1] (any trailing shape) in edge-id
    order; with `eids` the softmax runs over that edge subgraph (softmax.py:58-59) and logits holds one
    row per selected edge."""
    if eids is not None:
        # softmax.py:58-59: `g.edge_subgraph(eids)`; row i of logits belongs to edge eids[i]
        gidx, _ = graph.edge_subindex(eids)
    else:
        gidx = _gidx(graph)
    if logits.shape[0] != gidx.number_of_edges():
        raise ValueError("logits must hold one row per edge")
    if os.environ.get("B200_EDGE_SOFTMAX", "fused") == "builtin":
        return EdgeSoftmaxBuiltin.apply(gidx, logits)
    return EdgeSoftmaxFused.apply(gidx, logits)

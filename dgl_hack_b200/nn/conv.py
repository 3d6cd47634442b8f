"""Graph convolution modules (python/dgl/nn/pytorch/conv/*.py) over `Graph.update_all`.

Same constructor arguments, parameter names / shapes / initialisers and forward arithmetic as the
reference modules, so results match the reference CPU path on identical inputs and state dicts
(tests/test_gpu_nn.py against fixtures produced by the unmodified reference). Dense projections
are torch (cuBLAS) matmuls exactly as in the reference; everything that touches the graph goes
through the B200 kernels.
"""
import os

import torch as th
from torch import nn
from torch.nn import functional as tF
from torch.nn import init

from .. import backend as B
from .. import dense
from .. import function as fn
from .._lib import DGLError
from .softmax import edge_softmax

__all__ = ["GraphConv", "SAGEConv", "GATConv", "FusedGATConv", "RelGraphConv"]


class Identity(nn.Module):
    def forward(self, x):
        return x


def _bcast_rows(vec, like):
    return vec.reshape(vec.shape + (1,) * (like.dim() - 1))


def _matmul(x, w):
    """`th.matmul(x, w)`; with B200_DENSE=bf16x9 a 2-D product runs as the emulated-fp32
    tensor-core GEMM of `dgl_hack_b200.dense` (opt-in)."""
    if x.dim() == 2 and w.dim() == 2 and x.is_cuda and x.dtype == th.float32 and dense.enabled():
        return dense.matmul(x, w)
    return th.matmul(x, w)


class GraphConv(nn.Module):
    """graphconv.py:14-188. h' = act(D_in^-1/2 A D_out^-1/2 h W + b) for norm='both';
    'right' divides by the in-degree, 'none' does not normalise. W is applied before the
    aggregation when in_feats > out_feats (graphconv.py:144-152)."""

    def __init__(self, in_feats, out_feats, norm="both", weight=True, bias=True, activation=None):
        super(GraphConv, self).__init__()
        if norm not in ("none", "both", "right"):
            raise DGLError('Invalid norm value. Must be either "none", "both" or "right".'
                           ' But got "{}".'.format(norm))
        self._in_feats, self._out_feats, self._norm = in_feats, out_feats, norm
        if weight:
            self.weight = nn.Parameter(th.Tensor(in_feats, out_feats))
        else:
            self.register_parameter("weight", None)
        if bias:
            self.bias = nn.Parameter(th.Tensor(out_feats))
        else:
            self.register_parameter("bias", None)
        self.reset_parameters()
        self._activation = activation

    def reset_parameters(self):
        if self.weight is not None:
            init.xavier_uniform_(self.weight)
        if self.bias is not None:
            init.zeros_(self.bias)

    def _norms(self, graph):
        """(source scale, destination scale) of graphconv.py:128-133, 165-172 as float vectors,
        cached on the graph index (they depend on the structure only)."""
        cache = graph.gidx.__dict__.setdefault("_graphconv_norm", {})
        if self._norm not in cache:
            ind = graph.in_degrees().float().clamp(min=1)
            if self._norm == "both":
                outd = graph.out_degrees().float().clamp(min=1)
                cache[self._norm] = (th.pow(outd, -0.5), th.pow(ind, -0.5))
            else:
                cache[self._norm] = (None, 1.0 / ind)
        return cache[self._norm]

    def forward(self, graph, feat, weight=None):
        if weight is not None:
            if self.weight is not None:
                raise DGLError("External weight is provided while at the same time the module has"
                               " defined its own weight parameter. Please create the module with"
                               " flag weight=False.")
        else:
            weight = self.weight
        mult_first = self._in_feats > self._out_feats
        if os.environ.get("B200_GRAPHCONV_FUSED", "1") == "0":
            return self._forward_unfused(graph, feat, weight, mult_first)
        # Same arithmetic as the reference, with `feat * norm_src` and `rst * norm_dst` applied
        # inside the aggregation kernel (one pass over the features instead of three).
        cs, rs = self._norms(graph) if self._norm != "none" else (None, None)
        pad = 0
        if mult_first and weight is not None:
            # aggregate at a width that is a multiple of 4 floats: 16-byte aligned rows take the
            # vector path of the kernel (47 classes: 4.7 -> 2.6 ms at products scale); the zero
            # columns are dropped again after the aggregation
            pad = (-self._out_feats) % 4 if feat.dim() == 2 else 0
            w = tF.pad(weight, (0, pad)) if pad else weight
            # the reference scales the rows before the product (graphconv.py:133, 148); scaling
            # inside the aggregation is the same product per element, one rounding later
            feat = _matmul(feat, w)
        rst = B.scaled_copy_sum(graph.gidx, feat, cs, rs)
        if pad:
            rst = rst[..., :self._out_feats]
        if not mult_first and weight is not None:
            rst = _matmul(rst, weight)
        if self.bias is not None:
            rst = rst + self.bias
        if self._activation is not None:
            rst = self._activation(rst)
        return rst

    def _forward_unfused(self, graph, feat, weight, mult_first):
        """graphconv.py:128-176 step by step (B200_GRAPHCONV_FUSED=0)."""
        graph = graph.local_var()
        if self._norm == "both":
            degs = graph.out_degrees().to(feat.device).float().clamp(min=1)
            feat = feat * _bcast_rows(th.pow(degs, -0.5), feat)
        if mult_first and weight is not None:
            feat = th.matmul(feat, weight)
        graph.srcdata["h"] = feat
        graph.update_all(fn.copy_src(src="h", out="m"), fn.sum(msg="m", out="h"))
        rst = graph.dstdata["h"]
        if not mult_first and weight is not None:
            rst = th.matmul(rst, weight)
        if self._norm != "none":
            degs = graph.in_degrees().to(feat.device).float().clamp(min=1)
            norm = th.pow(degs, -0.5) if self._norm == "both" else 1.0 / degs
            rst = rst * _bcast_rows(norm, feat)
        if self.bias is not None:
            rst = rst + self.bias
        if self._activation is not None:
            rst = self._activation(rst)
        return rst

    def extra_repr(self):
        return "in={}, out={}, normalization={}".format(self._in_feats, self._out_feats, self._norm)


class SAGEConv(nn.Module):
    """sageconv.py:12-161; aggregator_type in mean / gcn / pool. 'lstm' reduces through a UDF
    over degree buckets (sageconv.py:85-96), which is outside the builtin kernel path.

    `project_first` (extension, off by default = the reference's order of operations): for the
    linear aggregators (mean, gcn) apply fc_neigh's weight BEFORE the aggregation, the way
    GraphConv does when in_feats > out_feats (graphconv.py:144-152). mean(X) W == mean(X W)
    up to fp32 rounding, and the SpMM then gathers out_feats instead of in_feats floats per edge
    (602 -> 16 on the Reddit-shaped config). "auto" enables it when in_feats > out_feats."""

    def __init__(self, in_feats, out_feats, aggregator_type, feat_drop=0., bias=True, norm=None,
                 activation=None, project_first=False):
        super(SAGEConv, self).__init__()
        if isinstance(in_feats, tuple):
            self._in_src_feats, self._in_dst_feats = in_feats
        else:
            self._in_src_feats = self._in_dst_feats = in_feats
        self._out_feats = out_feats
        self._aggre_type = aggregator_type
        self.norm = norm
        self.feat_drop = nn.Dropout(feat_drop)
        self.activation = activation
        if aggregator_type not in ("mean", "gcn", "pool"):
            if aggregator_type == "lstm":
                raise NotImplementedError("the lstm aggregator is a degree-bucketed UDF reducer, "
                                          "outside the builtin kernel path")
            raise KeyError("Aggregator type {} not recognized.".format(aggregator_type))
        if aggregator_type == "pool":
            self.fc_pool = nn.Linear(self._in_src_feats, self._in_src_feats)
        if aggregator_type != "gcn":
            self.fc_self = nn.Linear(self._in_dst_feats, out_feats, bias=bias)
        self.fc_neigh = nn.Linear(self._in_src_feats, out_feats, bias=bias)
        if project_first == "auto":
            project_first = self._in_src_feats > out_feats
        self._project_first = bool(project_first) and aggregator_type in ("mean", "gcn")
        self.reset_parameters()

    def reset_parameters(self):
        gain = nn.init.calculate_gain("relu")
        if self._aggre_type == "pool":
            nn.init.xavier_uniform_(self.fc_pool.weight, gain=gain)
        if self._aggre_type != "gcn":
            nn.init.xavier_uniform_(self.fc_self.weight, gain=gain)
        nn.init.xavier_uniform_(self.fc_neigh.weight, gain=gain)

    def forward(self, graph, feat):
        graph = graph.local_var()
        if isinstance(feat, tuple):
            feat_src, feat_dst = self.feat_drop(feat[0]), self.feat_drop(feat[1])
        else:
            feat_src = feat_dst = self.feat_drop(feat)
        h_self = feat_dst
        if self._project_first:
            return self._forward_project_first(graph, feat, feat_src, feat_dst)
        if self._aggre_type == "mean":
            graph.srcdata["h"] = feat_src
            graph.update_all(fn.copy_src("h", "m"), fn.mean("m", "neigh"))
            h_neigh = graph.dstdata["neigh"]
        elif self._aggre_type == "gcn":
            if isinstance(feat, tuple) and feat[0].shape[1:] != feat[1].shape[1:]:
                raise DGLError("gcn aggregator needs source and destination features of one shape")
            graph.srcdata["h"] = feat_src
            graph.update_all(fn.copy_src("h", "m"), fn.sum("m", "neigh"))
            degs = graph.in_degrees().to(feat_dst)
            h_neigh = (graph.dstdata["neigh"] + feat_dst) / (degs.unsqueeze(-1) + 1)
        else:  # pool
            graph.srcdata["h"] = tF.relu(self.fc_pool(feat_src))
            graph.update_all(fn.copy_src("h", "m"), fn.max("m", "neigh"))
            h_neigh = graph.dstdata["neigh"]
        if self._aggre_type == "gcn":
            rst = self.fc_neigh(h_neigh)
        else:
            rst = self.fc_self(h_self) + self.fc_neigh(h_neigh)
        if self.activation is not None:
            rst = self.activation(rst)
        if self.norm is not None:
            rst = self.norm(rst)
        return rst

    def _forward_project_first(self, graph, feat, feat_src, feat_dst):
        w, b = self.fc_neigh.weight, self.fc_neigh.bias
        graph.srcdata["h"] = tF.linear(feat_src, w)
        if self._aggre_type == "mean":
            graph.update_all(fn.copy_src("h", "m"), fn.mean("m", "neigh"))
            rst = self.fc_self(feat_dst) + graph.dstdata["neigh"]
        else:  # gcn
            if isinstance(feat, tuple) and feat[0].shape[1:] != feat[1].shape[1:]:
                raise DGLError("gcn aggregator needs source and destination features of one shape")
            graph.update_all(fn.copy_src("h", "m"), fn.sum("m", "neigh"))
            degs = graph.in_degrees().to(feat_dst)
            h_dst = graph.srcdata["h"] if feat_dst is feat_src else tF.linear(feat_dst, w)
            rst = (graph.dstdata["neigh"] + h_dst[:graph.number_of_dst_nodes()]) \
                / (degs.unsqueeze(-1) + 1)
        if b is not None:
            rst = rst + b
        if self.activation is not None:
            rst = self.activation(rst)
        if self.norm is not None:
            rst = self.norm(rst)
        return rst


class GATConv(nn.Module):
    """gatconv.py:14-171 (the fork's timing prints and cuda synchronisations are not part of the
    arithmetic and are left out). Attention = u_add_v -> leaky_relu -> edge_softmax -> u_mul_e+sum.
    With `fused=True` (what the fork's FusedGATConv does, fusedGatConv.py:151) the attention runs
    as the fused GAT kernel; it needs attn_drop == 0 and, like the reference's fused op, applies no
    max-subtraction. Head widths the fused kernel does not take (backend.fused_gat_supported) run
    the builtin composition on the same kernels library."""

    def __init__(self, in_feats, out_feats, num_heads, feat_drop=0., attn_drop=0.,
                 negative_slope=0.2, residual=False, activation=None, fused=False):
        super(GATConv, self).__init__()
        self._num_heads = num_heads
        if isinstance(in_feats, tuple):
            self._in_src_feats, self._in_dst_feats = in_feats
        else:
            self._in_src_feats = self._in_dst_feats = in_feats
        self._out_feats = out_feats
        if isinstance(in_feats, tuple):
            self.fc_src = nn.Linear(self._in_src_feats, out_feats * num_heads, bias=False)
            self.fc_dst = nn.Linear(self._in_dst_feats, out_feats * num_heads, bias=False)
        else:
            self.fc = nn.Linear(self._in_src_feats, out_feats * num_heads, bias=False)
        self.attn_l = nn.Parameter(th.FloatTensor(size=(1, num_heads, out_feats)))
        self.attn_r = nn.Parameter(th.FloatTensor(size=(1, num_heads, out_feats)))
        self.feat_drop = nn.Dropout(feat_drop)
        self.attn_drop = nn.Dropout(attn_drop)
        self.negative_slope = negative_slope
        self.leaky_relu = nn.LeakyReLU(negative_slope)
        if residual:
            if self._in_dst_feats != out_feats:
                self.res_fc = nn.Linear(self._in_dst_feats, num_heads * out_feats, bias=False)
            else:
                self.res_fc = Identity()
        else:
            self.register_buffer("res_fc", None)
        self.reset_parameters()
        self.activation = activation
        if fused and attn_drop != 0.:
            raise DGLError("the fused attention kernel has no attention dropout")
        self._fused = fused

    def reset_parameters(self):
        gain = nn.init.calculate_gain("relu")
        if hasattr(self, "fc"):
            nn.init.xavier_normal_(self.fc.weight, gain=gain)
        else:
            nn.init.xavier_normal_(self.fc_src.weight, gain=gain)
            nn.init.xavier_normal_(self.fc_dst.weight, gain=gain)
        nn.init.xavier_normal_(self.attn_l, gain=gain)
        nn.init.xavier_normal_(self.attn_r, gain=gain)
        if isinstance(self.res_fc, nn.Linear):
            nn.init.xavier_normal_(self.res_fc.weight, gain=gain)

    def forward(self, graph, feat):
        graph = graph.local_var()
        if isinstance(feat, tuple):
            h_src, h_dst = self.feat_drop(feat[0]), self.feat_drop(feat[1])
            feat_src = self.fc_src(h_src).view(-1, self._num_heads, self._out_feats)
            feat_dst = self.fc_dst(h_dst).view(-1, self._num_heads, self._out_feats)
        else:
            h_src = h_dst = self.feat_drop(feat)
            feat_src = feat_dst = self.fc(h_src).view(-1, self._num_heads, self._out_feats)
        # el = (feat_src * attn_l).sum(-1), er = (feat_dst * attn_r).sum(-1), gatconv.py:143-144:
        # one streaming kernel each way instead of torch's mul + reduce passes
        if feat_src is feat_dst and self._num_heads * self._out_feats <= 1024:
            el, er = B.gat_attn_proj(feat_src, self.attn_l, self.attn_r)
        elif self._num_heads * self._out_feats <= 1024:
            el, _ = B.gat_attn_proj(feat_src, self.attn_l, None)
            _, er = B.gat_attn_proj(feat_dst, None, self.attn_r)
        else:
            el = (feat_src * self.attn_l).sum(dim=-1).unsqueeze(-1)
            er = (feat_dst * self.attn_r).sum(dim=-1).unsqueeze(-1)
        if self._fused and B.fused_gat_supported(self._num_heads, self._out_feats):
            rst = B.fused_gat(graph.gidx, feat_src, el, er, self.negative_slope)
        else:
            graph.srcdata.update({"ft": feat_src, "el": el})
            graph.dstdata.update({"er": er})
            graph.apply_edges(fn.u_add_v("el", "er", "e"))
            e = self.leaky_relu(graph.edata.pop("e"))
            graph.edata["a"] = self.attn_drop(edge_softmax(graph, e))
            graph.update_all(fn.u_mul_e("ft", "a", "m"), fn.sum("m", "ft"))
            rst = graph.dstdata["ft"]
        if self.res_fc is not None:
            resval = self.res_fc(h_dst).view(h_dst.shape[0], -1, self._out_feats)
            rst = rst + resval
        if self.activation:
            rst = self.activation(rst)
        return rst


class FusedGATConv(GATConv):
    """fusedGatConv.py:14-167: GATConv whose attention is the fused kernel (no attention dropout
    in the arithmetic, no max-subtraction)."""

    def __init__(self, in_feats, out_feats, num_heads, feat_drop=0., attn_drop=0.,
                 negative_slope=0.2, residual=False, activation=None):
        super(FusedGATConv, self).__init__(in_feats, out_feats, num_heads, feat_drop, 0.,
                                           negative_slope, residual, activation, fused=True)


class RelGraphConv(nn.Module):
    """relgraphconv.py:13-193. h'_v = act(sum_{(u->v, r)} norm_e * h_u W_r + b [+ h_v W_loop]),
    W_r from the 'basis' or 'bdd' regulariser.

    The reference evaluates a message UDF that materialises W[type_e] per edge ([E, in, out]) and
    reduces the [E, out] messages with copy_e+sum. Here the basis form with dense input and
    in*out <= 1024 runs as ONE typed-adjacency kernel (backend.rgcn_layer1: gather, per-relation
    contraction and reduction fused, no [E, .] temporaries); the typed adjacency is built on the
    device once per (graph, etypes) and cached on the graph. Other cases (bdd, wide layers,
    integer node ids) follow the reference's message form through `Graph.update_all`.
    Reference quirk not reproduced: for integer `x`, utils.bmm_maybe_select (nn/pytorch/utils.py:
    84-88) indexes the flattened weight with `index * out_feat + x` where `index * num_nodes + x`
    is meant; here W[type_e, x_u, :] is used."""

    def __init__(self, in_feat, out_feat, num_rels, regularizer="basis", num_bases=None,
                 bias=True, activation=None, self_loop=False, dropout=0.0):
        super(RelGraphConv, self).__init__()
        self.in_feat, self.out_feat, self.num_rels = in_feat, out_feat, num_rels
        self.regularizer = regularizer
        self.num_bases = num_bases
        if self.num_bases is None or self.num_bases > self.num_rels or self.num_bases <= 0:
            self.num_bases = self.num_rels
        self.bias, self.activation, self.self_loop = bias, activation, self_loop
        gain = nn.init.calculate_gain("relu")
        if regularizer == "basis":
            self.weight = nn.Parameter(th.Tensor(self.num_bases, self.in_feat, self.out_feat))
            if self.num_bases < self.num_rels:
                self.w_comp = nn.Parameter(th.Tensor(self.num_rels, self.num_bases))
            nn.init.xavier_uniform_(self.weight, gain=gain)
            if self.num_bases < self.num_rels:
                nn.init.xavier_uniform_(self.w_comp, gain=gain)
            self.message_func = self.basis_message_func
        elif regularizer == "bdd":
            if in_feat % self.num_bases != 0 or out_feat % self.num_bases != 0:
                raise ValueError("Feature size must be a multiplier of num_bases (%d)."
                                 % self.num_bases)
            self.submat_in = in_feat // self.num_bases
            self.submat_out = out_feat // self.num_bases
            self.weight = nn.Parameter(th.Tensor(
                self.num_rels, self.num_bases * self.submat_in * self.submat_out))
            nn.init.xavier_uniform_(self.weight, gain=gain)
            self.message_func = self.bdd_message_func
        else:
            raise ValueError("Regularizer must be either 'basis' or 'bdd'")
        if self.bias:
            self.h_bias = nn.Parameter(th.Tensor(out_feat))
            nn.init.zeros_(self.h_bias)
        if self.self_loop:
            self.loop_weight = nn.Parameter(th.Tensor(in_feat, out_feat))
            nn.init.xavier_uniform_(self.loop_weight, gain=gain)
        self.dropout = nn.Dropout(dropout)

    def _relation_weight(self):
        if self.num_bases < self.num_rels:
            weight = self.weight.view(self.num_bases, self.in_feat * self.out_feat)
            return th.matmul(self.w_comp, weight).view(self.num_rels, self.in_feat, self.out_feat)
        return self.weight

    def basis_message_func(self, edges):
        weight = self._relation_weight()
        h, etype = edges.src["h"], edges.data["type"]
        if h.dtype == th.int64 and h.dim() == 1:
            msg = weight.reshape(-1, self.out_feat).index_select(0, etype * weight.shape[1] + h)
        else:
            msg = th.bmm(h.unsqueeze(1), weight.index_select(0, etype)).squeeze(1)
        if "norm" in edges.data:
            msg = msg * edges.data["norm"]
        return {"msg": msg}

    def bdd_message_func(self, edges):
        h = edges.src["h"]
        if h.dtype == th.int64 and h.dim() == 1:
            raise TypeError("Block decomposition does not allow integer ID feature.")
        weight = self.weight.index_select(0, edges.data["type"]).view(
            -1, self.submat_in, self.submat_out)
        msg = th.bmm(h.reshape(-1, 1, self.submat_in), weight).view(-1, self.out_feat)
        if "norm" in edges.data:
            msg = msg * edges.data["norm"]
        return {"msg": msg}

    def _typed_index(self, g, etypes):
        """GraphIndex with the typed adjacency for `etypes`, cached on the graph."""
        # the cache holds the tensor itself: its address cannot be handed to another tensor while
        # the entry lives, so (identity, version) identifies the contents
        cache = g.gidx.__dict__.setdefault("_typed_cache", {})
        if cache.get("etypes") is not etypes or cache.get("version") != etypes._version or \
                cache.get("num_rels") != self.num_rels:
            g.gidx.set_edge_types(etypes, self.num_rels)
            cache.update(etypes=etypes, version=etypes._version, num_rels=self.num_rels)
        return g.gidx

    def _fused_ok(self, x):
        return (self.regularizer == "basis" and x.is_floating_point() and x.dim() == 2
                and self.in_feat <= 128 and self.out_feat <= 128
                and self.in_feat * self.out_feat <= 1024)

    def _tc_ok(self, x):
        """Wide layers (beyond the fused SIMT kernel), or B200_RGCN=tc: per-relation contraction
        on the tensor cores (backend.rgcn_layer1_tc)."""
        if not (self.regularizer == "basis" and x.is_floating_point() and x.dim() == 2):
            return False
        want = os.environ.get("B200_RGCN", "") == "tc" or not self._fused_ok(x)
        return want and os.environ.get("B200_RGCN", "") != "udf" and \
            self.in_feat % 4 == 0 and self.in_feat <= 256 and self.out_feat <= 256 and \
            x.is_cuda and x.dtype == th.float32

    def forward(self, g, x, etypes, norm=None):
        assert g.is_homograph(), \
            "not a homograph; convert it with to_homo and pass in the edge type as argument"
        if self.self_loop:
            if x.dtype == th.int64 and x.dim() == 1:
                loop_message = self.loop_weight.index_select(0, x)
            else:
                loop_message = th.matmul(x, self.loop_weight)
        if self._tc_ok(x):
            gidx = self._typed_index(g, etypes)
            node_repr = B.rgcn_layer1_tc(gidx, x, self._relation_weight(), norm)
        elif self._fused_ok(x):
            gidx = self._typed_index(g, etypes)
            if norm is None:
                norm = x.new_ones((g.number_of_edges(), 1))
            node_repr = B.rgcn_layer1(gidx, x, self._relation_weight(), norm)
        else:
            with g.local_scope():
                g.ndata["h"] = x
                g.edata["type"] = etypes
                if norm is not None:
                    g.edata["norm"] = norm
                g.update_all(self.message_func, fn.sum(msg="msg", out="h"))
                node_repr = g.ndata["h"]
        if self.bias:
            node_repr = node_repr + self.h_bias
        if self.self_loop:
            node_repr = node_repr + loop_message
        if self.activation:
            node_repr = self.activation(node_repr)
        return self.dropout(node_repr)

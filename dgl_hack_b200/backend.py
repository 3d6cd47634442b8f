"""PyTorch-side kernel dispatch: autograd wrappers around dgl_hack_b200.kernel.

Same public surface as the reference's torch backend for this path
(python/dgl/backend/pytorch/tensor.py:286-381 BinaryReduce / binary_reduce, :519-569 CopyReduce /
copy_reduce, :572-601 _reduce_grad; interface python/dgl/backend/backend.py:1189-1300):

    binary_reduce(reducer, binary_op, graph, lhs, rhs, lhs_data, rhs_data, out_size,
                  lhs_map=(None, None), rhs_map=(None, None), out_map=(None, None))
    copy_reduce(reducer, graph, target, in_data, out_size, in_map=(None, None),
                out_map=(None, None))

`graph` is a dgl_hack_b200.GraphIndex; maps are (forward_map, backward_map) pairs of int32 CUDA
tensors. "mean" is handled here like in the reference (the C kernels reject it): sum, then divide
by the clamped in-degree; gradients are first divided by the same degrees. When there are no id
mappings the degree comes straight from the in-CSR and copy+mean runs as ONE fused kernel.
All tensors must live on the graph's CUDA device: there is no CPU fallback.
"""
import torch as th

from . import kernel as K
from ._lib import DGLError

SRC, DST, EDGE = 0, 1, 2  # python/dgl/function/base.py TargetCode


def _degrees_for_mean(graph, target, n_in, in_map, out_map, out_rows, like):
    """In-degree per output row as a float tensor (clamped to >= 1), shaped for broadcasting
    against an output of `like.dim()` dims. With mappings this is the reference's second
    copy_reduce over a ones-vector (tensor.py:313-327, 529-540)."""
    if in_map is None and out_map is None and out_rows == graph.num_dst:
        return graph.mean_divisor(like.dtype).reshape((out_rows,) + (1,) * (like.dim() - 1))
    else:
        ones = like.new_ones((n_in,))
        degs = like.new_empty((out_rows,))
        K.copy_reduce("sum", graph, target, ones, degs, in_map, out_map)
    return degs.reshape((out_rows,) + (1,) * (like.dim() - 1)).clamp(min=1)


def _reduce_grad(grad, shape):
    """Sum an unreduced gradient (output feature shape) back onto the operand's own feature shape
    (tensor.py:572-601)."""
    shape = tuple(shape)
    g_feat, in_feat = tuple(grad.shape[1:]), shape[1:]
    if g_feat == in_feat:
        return grad
    padded = (1,) * (len(g_feat) - len(in_feat)) + in_feat
    dims = tuple(i + 1 for i, (a, b) in enumerate(zip(g_feat, padded)) if a != b)
    return grad.sum(dim=dims, keepdim=True).view(shape)


def _edge_scalar_grad(kreducer, binary_op, graph, tgt, other_tgt, data, other, grad_out, maps,
                      feat_shape):
    """Gradient of an EDGE operand that broadcasts along the last feature axis under mul + sum
    (GAT's u_mul_e with ft [N,H,F] x a [E,H,1]): the reference materialises the unreduced
    gradient [E,H,F] and lets torch sum it over F (tensor.py:345-360, _reduce_grad); the same
    values are one dot product per (edge, head) of the other operand's row with grad_out[dst], so
    the [E,H,F] temporary (E*H*F*4 bytes written and read back) never exists. Returns None when
    the shapes are not of that form."""
    if kreducer != "sum" or binary_op != "mul" or tgt != EDGE or other_tgt not in (SRC, DST):
        return None
    if any(m is not None for pair in maps for m in pair):
        return None
    ds, os_ = tuple(data.shape[1:]), tuple(other.shape[1:])
    if (os_ != tuple(feat_shape) or len(ds) != len(os_) or len(ds) == 0 or ds[:-1] != os_[:-1]
            or ds[-1] != 1 or os_[-1] == 1):
        return None
    out = grad_out.new_empty((data.shape[0],) + ds[:-1])
    K.binary_op_reduce("none", "dot", graph, other_tgt, DST, other, grad_out, out)
    return out.unsqueeze(-1)


class BinaryReduce(th.autograd.Function):
    @staticmethod
    def forward(ctx, reducer, binary_op, graph, lhs, rhs, lhs_data, rhs_data, out_data,
                out_size, lhs_map, rhs_map, out_map):
        feat_shape = K.infer_binary_feature_shape(binary_op, lhs_data, rhs_data)
        kreducer = "sum" if reducer == "mean" else reducer
        K.binary_op_reduce(kreducer, binary_op, graph, lhs, rhs, lhs_data, rhs_data, out_data,
                           lhs_map[0], rhs_map[0], out_map[0])
        degs = None
        if reducer == "mean":
            # degree over the operand that is NOT the destination (tensor.py:313-321)
            if lhs != DST:
                tgt, n, imap = lhs, lhs_data.shape[0], lhs_map[0]
            else:
                tgt, n, imap = rhs, rhs_data.shape[0], rhs_map[0]
            degs = _degrees_for_mean(graph, tgt, n, imap, out_map[0], out_data.shape[0], out_data)
            out_data = out_data / degs
        ctx.backward_cache = (reducer, binary_op, graph, lhs, rhs, lhs_map, rhs_map, out_map,
                              feat_shape, degs)
        ctx.save_for_backward(lhs_data, rhs_data, out_data)
        return out_data

    @staticmethod
    def backward(ctx, grad_out):
        (reducer, binary_op, graph, lhs, rhs, lhs_map, rhs_map, out_map, feat_shape,
         degs) = ctx.backward_cache
        lhs_data, rhs_data, out_data = ctx.saved_tensors
        kreducer = "sum" if reducer == "mean" else reducer
        if reducer == "mean":
            grad_out = grad_out / degs
        grad_out = grad_out.contiguous()
        grad_lhs = grad_rhs = None
        maps = (lhs_map, rhs_map, out_map)
        if ctx.needs_input_grad[5]:
            grad_lhs = _edge_scalar_grad(kreducer, binary_op, graph, lhs, rhs, lhs_data, rhs_data,
                                         grad_out, maps, feat_shape)
        if ctx.needs_input_grad[5] and grad_lhs is None:
            grad_lhs = grad_out.new_empty((lhs_data.shape[0],) + feat_shape)
            K.backward_lhs_binary_op_reduce(kreducer, binary_op, graph, lhs, rhs, lhs_data,
                                            rhs_data, out_data, grad_out, grad_lhs, lhs_map[1],
                                            rhs_map[1], out_map[1])
            grad_lhs = _reduce_grad(grad_lhs, lhs_data.shape)
        if ctx.needs_input_grad[6]:
            grad_rhs = _edge_scalar_grad(kreducer, binary_op, graph, rhs, lhs, rhs_data, lhs_data,
                                         grad_out, maps, feat_shape)
        if ctx.needs_input_grad[6] and grad_rhs is None:
            grad_rhs = grad_out.new_empty((rhs_data.shape[0],) + feat_shape)
            K.backward_rhs_binary_op_reduce(kreducer, binary_op, graph, lhs, rhs, lhs_data,
                                            rhs_data, out_data, grad_out, grad_rhs, lhs_map[1],
                                            rhs_map[1], out_map[1])
            grad_rhs = _reduce_grad(grad_rhs, rhs_data.shape)
        return (None, None, None, None, None, grad_lhs, grad_rhs, None, None, None, None, None)


def binary_reduce(reducer, binary_op, graph, lhs, rhs, lhs_data, rhs_data, out_size,
                  lhs_map=(None, None), rhs_map=(None, None), out_map=(None, None)):
    lhs_data, rhs_data = lhs_data.contiguous(), rhs_data.contiguous()
    feat_shape = K.infer_binary_feature_shape(binary_op, lhs_data, rhs_data)
    out_shape = feat_shape[:-1] if binary_op == "dot" else feat_shape
    out_data = lhs_data.new_empty((out_size,) + out_shape)
    return BinaryReduce.apply(reducer, binary_op, graph, lhs, rhs, lhs_data, rhs_data, out_data,
                              out_size, lhs_map, rhs_map, out_map)


class CopyReduce(th.autograd.Function):
    @staticmethod
    def forward(ctx, reducer, graph, target, in_data, out_data, out_size, in_map, out_map):
        degs = None
        if reducer == "mean":
            degs = _degrees_for_mean(graph, target, in_data.shape[0], in_map[0], out_map[0],
                                     out_data.shape[0], out_data)
            if in_map[0] is None and out_map[0] is None and out_size == graph.num_dst:
                # one fused pass: sum in in-CSR order, then one IEEE division by the degree
                K.copy_reduce_mean(graph, target, in_data, out_data)
            else:
                K.copy_reduce("sum", graph, target, in_data, out_data, in_map[0], out_map[0])
                out_data = out_data / degs
        else:
            K.copy_reduce(reducer, graph, target, in_data, out_data, in_map[0], out_map[0])
        ctx.backward_cache = (reducer, graph, target, in_map, out_map, degs)
        ctx.save_for_backward(in_data, out_data)
        return out_data

    @staticmethod
    def backward(ctx, grad_out):
        reducer, graph, target, in_map, out_map, degs = ctx.backward_cache
        in_data, out_data = ctx.saved_tensors
        grad_in = None
        if reducer == "mean":
            grad_out = grad_out / degs
        grad_out = grad_out.contiguous()
        if ctx.needs_input_grad[3]:
            grad_in = grad_out.new_empty(in_data.shape)
            K.backward_copy_reduce("sum" if reducer == "mean" else reducer, graph, target,
                                   in_data, out_data, grad_out, grad_in, in_map[1], out_map[1])
        return None, None, None, grad_in, None, None, None, None


def copy_reduce(reducer, graph, target, in_data, out_size, in_map=(None, None),
                out_map=(None, None)):
    in_data = in_data.contiguous()
    out_data = in_data.new_empty((out_size,) + in_data.shape[1:])
    return CopyReduce.apply(reducer, graph, target, in_data, out_data, out_size, in_map, out_map)


class FusedGat(th.autograd.Function):
    """tensor.py:383-413 FusedGat (the fork's fused attention op), same saved state."""

    @staticmethod
    def forward(ctx, graph, feat_src, el, er, s, exp, ret, slope):
        ctx.backward_cache = graph, feat_src, el, er, s, exp, ret, slope
        K.fused_gat_kernel(graph, feat_src, el, er, s, exp, ret, slope)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        graph, feat_src, el, er, s, exp, ret, slope = ctx.backward_cache
        gradout = gradout.contiguous()
        grad_el, grad_er = th.empty_like(el), th.empty_like(er)
        grad_feat_src = th.empty_like(feat_src)
        K.backward_fused_gat(graph, feat_src, el, er, s, exp, ret, gradout, grad_feat_src,
                             grad_el, grad_er, slope)
        return None, grad_feat_src, grad_el, grad_er, None, None, None, None


class ScaledCopySum(th.autograd.Function):
    """out[v] = row_scale[v] * sum_{u->v} col_scale[u] * x[u]; the scales are constants of the
    graph (degree norms), only x is differentiated: grad_x[u] = col_scale[u] * sum_{u->v}
    row_scale[v] * grad_out[v] -- the same kernel on the transposed walk with the scales swapped."""

    @staticmethod
    def forward(ctx, graph, x, col_scale, row_scale):
        out = x.new_empty((graph.num_dst,) + tuple(x.shape[1:]))
        K.scaled_copy_sum(graph, False, x, col_scale, row_scale, out)
        ctx.backward_cache = graph, col_scale, row_scale
        return out

    @staticmethod
    def backward(ctx, grad_out):
        graph, col_scale, row_scale = ctx.backward_cache
        grad_out = grad_out.contiguous()
        grad_x = grad_out.new_empty((graph.num_src,) + tuple(grad_out.shape[1:]))
        K.scaled_copy_sum(graph, True, grad_out, row_scale, col_scale, grad_x)
        return None, grad_x, None, None


def scaled_copy_sum(graph, x, col_scale=None, row_scale=None):
    """update_all(copy_src, sum) with per-source and per-destination scalars applied inside the
    kernel (GraphConv's norm='both' / 'right'); scales are float vectors of length num_src /
    num_dst that need no gradient."""
    cs = col_scale.detach().contiguous().reshape(-1) if col_scale is not None else None
    rs = row_scale.detach().contiguous().reshape(-1) if row_scale is not None else None
    return ScaledCopySum.apply(graph, x.contiguous(), cs, rs)


class GatAttnProj(th.autograd.Function):
    """el / er of gatconv.py:143-144 as one streaming kernel each way (csrc/attn.cu)."""

    @staticmethod
    def forward(ctx, feat, attn_l, attn_r):
        n, h = feat.shape[0], feat.shape[1]
        el = feat.new_empty((n, h, 1)) if attn_l is not None else None
        er = feat.new_empty((n, h, 1)) if attn_r is not None else None
        K.gat_attn_proj(feat, attn_l, attn_r, el, er)
        ctx.save_for_backward(feat, attn_l, attn_r)
        return el, er

    @staticmethod
    def backward(ctx, grad_el, grad_er):
        feat, attn_l, attn_r = ctx.saved_tensors
        grad_feat = th.empty_like(feat)
        gl = grad_el.contiguous() if attn_l is not None else None
        gr = grad_er.contiguous() if attn_r is not None else None
        gal = th.empty_like(attn_l) if attn_l is not None else None
        gar = th.empty_like(attn_r) if attn_r is not None else None
        K.gat_attn_proj_backward(feat, attn_l, attn_r, gl, gr, grad_feat, gal, gar)
        return grad_feat, gal, gar


def gat_attn_proj(feat, attn_l=None, attn_r=None):
    """(el, er) = ((feat * attn_l).sum(-1, keepdim), (feat * attn_r).sum(-1, keepdim)) for
    feat [N, H, F] and attention vectors [1, H, F]; either vector may be None."""
    al = attn_l.contiguous() if attn_l is not None else None
    ar = attn_r.contiguous() if attn_r is not None else None
    return GatAttnProj.apply(feat.contiguous(), al, ar)


def fused_gat_supported(num_heads, out_feats):
    """Shapes the fused attention kernels take (csrc/gat.cu gat_plan): per-head width = VEC (4, 2
    or 1 floats) x a power of two, at most 1024 floats per row."""
    v = 4
    while v > 1 and out_feats % v:
        v >>= 1
    lph = out_feats // v
    return (lph & (lph - 1)) == 0 and (lph <= 32 or lph % 32 == 0) and \
        (num_heads * out_feats // v + 31) // 32 <= 8 and num_heads <= 1024


def fused_gat(graph, feat_src, el, er, slope):
    """tensor.py:415-420 fused_gat; `graph` is a GraphIndex on feat_src's device."""
    feat_src, el, er = feat_src.contiguous(), el.contiguous(), er.contiguous()
    exp = el.new_empty([graph.number_of_edges()] + list(el.size()[1:]))
    s = th.empty_like(er)
    ret = feat_src.new_empty((graph.num_dst,) + tuple(feat_src.shape[1:]))
    return FusedGat.apply(graph, feat_src, el, er, s, exp, ret, slope)


class RgcnFirstLayer(th.autograd.Function):
    """tensor.py:440-460 RgcnFirstLayer: one-hot input, weight [R, N, out] acts as an embedding."""

    @staticmethod
    def forward(ctx, graph, weight, norm, ret):
        ctx.backward_cache = graph, weight.size(), norm
        K.rgcn_layer0(graph, weight, norm, ret)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        graph, weight_size, norm = ctx.backward_cache
        grad_weight = th.zeros(weight_size, dtype=norm.dtype, device=norm.device)
        K.rgcn_layer0_backward(graph, gradout.contiguous(), norm, grad_weight)
        return None, grad_weight, None, None


def rgcn_layer0(graph, weight, norm):
    """tensor.py:462-465 rgcn_layer0; `graph` is a typed GraphIndex."""
    weight, norm = weight.contiguous(), norm.contiguous()
    ret = weight.new_empty((graph.num_dst, weight.size(2)))
    return RgcnFirstLayer.apply(graph, weight, norm, ret)


class RgcnSecondLayer(th.autograd.Function):
    """tensor.py:467-490 RgcnSecondLayer. The reference's backward computes grad_weight and then
    returns None for it (tensor.py:490), so the relation weights never train; here it is
    returned."""

    @staticmethod
    def forward(ctx, graph, x, weight, norm, ret):
        ctx.backward_cache = graph, weight, norm, x
        K.rgcn_layer1(graph, x, weight, norm, ret)
        return ret

    @staticmethod
    def backward(ctx, gradout):
        graph, weight, norm, x = ctx.backward_cache
        grad_x = th.empty_like(x)
        grad_weight = th.empty_like(weight)
        K.rgcn_layer1_backward(graph, x, weight, norm, gradout.contiguous(), grad_x, grad_weight)
        return None, grad_x, grad_weight, None, None


def rgcn_layer1(graph, x, weight, norm):
    """tensor.py:492-495 rgcn_layer1."""
    x, weight, norm = x.contiguous(), weight.contiguous(), norm.contiguous()
    ret = weight.new_empty((graph.num_dst, weight.size(2)))
    return RgcnSecondLayer.apply(graph, x, weight, norm, ret)


class CopyReduceBF16(th.autograd.Function):
    """copy_src + sum / mean on bf16 feature storage (fp32 accumulation): the explicit
    reduced-precision tier (no reference counterpart; its kernels are fp32-only)."""

    @staticmethod
    def forward(ctx, graph, x, mean):
        out = x.new_empty((graph.num_dst, x.shape[1]))
        K.copy_sum_bf16(graph, False, x, None, None, mean, out)
        ctx.backward_cache = (graph, mean)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        graph, mean = ctx.backward_cache
        grad_out = grad_out.contiguous()
        grad_x = grad_out.new_empty((graph.num_src, grad_out.shape[1]))
        # d mean / d x[u] = sum_{u -> v} grad_out[v] / max(deg v, 1): a per-gathered-node scale
        cs = (1.0 / graph.mean_divisor(th.float32)) if mean else None
        K.copy_sum_bf16(graph, True, grad_out, cs, None, False, grad_x)
        return None, grad_x, None


def to_bf16_storage(x):
    """[N, D] float tensor -> bf16 feature storage [N, ceil8(D)] (zero padded so that every row
    is a whole number of 16-byte vectors). Slice `[:, :D]` of a result to drop the padding."""
    pad = (-x.shape[1]) % 8
    x = x.to(th.bfloat16)
    return th.nn.functional.pad(x, (0, pad)).contiguous() if pad else x.contiguous()


def copy_reduce_bf16(reducer, graph, x):
    """`copy_src + sum|mean` on bf16 storage (see to_bf16_storage); returns bf16 [num_dst, D]."""
    if reducer not in ("sum", "mean"):
        raise DGLError("the bf16 tier implements the sum and mean reducers")
    if x.dtype != th.bfloat16 or x.dim() != 2 or x.shape[1] % 8:
        raise DGLError("copy_reduce_bf16 takes bf16 [N, D] storage with D % 8 == 0 "
                       "(backend.to_bf16_storage)")
    return CopyReduceBF16.apply(graph, x.contiguous(), reducer == "mean")


class GroupedMatMul(th.autograd.Function):
    """c[rows of group g] = a[rows of group g] @ w[g]: the per-relation dense contraction on the
    tensor cores (csrc/gemm.cu, grouped mode). grad_a runs on the same kernel (w[g] read
    transposed); grad_w[g] = a_g^T @ g_g reduces over the rows of a group (library GEMM per
    non-empty group)."""

    @staticmethod
    def forward(ctx, a, w, group_ptr):
        a, w = a.contiguous(), w.contiguous()
        c = a.new_empty((a.shape[0], w.shape[2]))
        K.grouped_dense_tf32x3(a, w, True, group_ptr, c)
        ctx.group_ptr = group_ptr
        ctx.save_for_backward(a, w)
        return c

    @staticmethod
    def backward(ctx, g):
        a, w = ctx.saved_tensors
        gp = ctx.group_ptr
        g = g.contiguous()
        ga = gw = None
        if ctx.needs_input_grad[0]:
            ga = th.empty_like(a)
            K.grouped_dense_tf32x3(g, w, False, gp, ga)
        if ctx.needs_input_grad[1]:
            gw = th.zeros_like(w)
            for r in range(w.shape[0]):
                if gp[r + 1] > gp[r]:
                    th.matmul(a[gp[r]:gp[r + 1]].t(), g[gp[r]:gp[r + 1]], out=gw[r])
        return ga, gw, None


def grouped_matmul_supported(x, weight):
    return (x.is_cuda and x.dtype == th.float32 and x.dim() == 2 and weight.dim() == 3
            and x.shape[1] % 4 == 0 and weight.shape[2] <= 256 and weight.shape[1] <= 256)


def rgcn_layer1_tc(graph, x, weight, norm=None):
    """ret[v] = sum_e norm_e * h[src_e] . W[type_e] with the per-relation contraction on the
    TENSOR CORES: (1) per-(relation, node) sums of norm_e * h[src_e] -- one SpMM launch on the
    relation-segment graph; (2) rows of relation r times W[r] -- grouped tcgen05 GEMM; (3) the
    relations' rows of a node added up -- one copy_e + sum launch. No [E, in, out] / [E, out]
    temporaries; widths up to 256 (the fused SIMT kernel rgcn_layer1 stops at in * out = 1024)."""
    sg, seg_dst, seg_type = graph.relation_segments()
    cache = sg.__dict__.setdefault("_tc_plan", {})
    if "group_ptr" not in cache:
        from .graph_index import GraphIndex
        counts = th.bincount(seg_type, minlength=weight.shape[0])
        cache["group_ptr"] = [0] + th.cumsum(counts, 0).tolist()
        cache["to_node"] = GraphIndex.from_coo(graph.num_dst, graph.num_dst, seg_dst, seg_dst,
                                               graph.device)
    n_seg = sg.num_dst
    if norm is None:
        s = copy_reduce("sum", sg, 0, x, n_seg)
    else:
        s = binary_reduce("sum", "mul", sg, 0, 2, x, norm.reshape(-1, 1), n_seg)
    y = GroupedMatMul.apply(s, weight, cache["group_ptr"])
    return copy_reduce("sum", cache["to_node"], 2, y, graph.num_dst)

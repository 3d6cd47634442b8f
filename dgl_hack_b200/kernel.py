"""Kernel wrappers with the names, argument order and meaning of python/dgl/kernel.py.

Each function binds one C-ABI entry point of include/dgl_b200.h, which replaces the PackedFunc of
the same name (`kernel._CAPI_DGLKernel*`, src/kernel/binary_reduce.cc). `G` is a
dgl_hack_b200.graph_index.GraphIndex; tensors are ndarray.NDArray views (or torch tensors);
optional `*_rows` mappings are None / NDArray / torch int32 tensors. Outputs are caller-allocated
and written in place, as in the reference.
"""
import ctypes

import torch

from . import _lib
from ._lib import LIB, check_call
from .ndarray import NDArray, null, zerocopy_to_dgl_ndarray


def _nd(x):
    return x if isinstance(x, NDArray) else zerocopy_to_dgl_ndarray(x)


def _map(x):
    """None -> nd.null() (python/dgl/kernel.py:138-148)."""
    return null() if x is None else _nd(x)


def _stream(G):
    return ctypes.c_void_p(torch.cuda.current_stream(G.device).cuda_stream)


def infer_binary_feature_shape(op, lhs, rhs):
    """kernel.py:8-26 -> kernel._CAPI_DGLKernelInferBinaryFeatureShape."""
    lhs, rhs = _nd(lhs), _nd(rhs)
    ndim = ctypes.c_int32()
    shape = (ctypes.c_int64 * 16)()
    check_call(LIB.DGLB200KernelInferBinaryFeatureShape(op.encode(), lhs.ptr, rhs.ptr,
                                                        ctypes.byref(ndim), shape))
    return tuple(int(shape[i]) for i in range(ndim.value))


def binary_op_reduce(reducer, op, G, A_target, B_target, A, B, out,
                     A_rows=None, B_rows=None, out_rows=None):
    """kernel.py:29-148 -> kernel._CAPI_DGLKernelBinaryOpReduce."""
    A, B, out = _nd(A), _nd(B), _nd(out)
    ar, br, orr = _map(A_rows), _map(B_rows), _map(out_rows)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelBinaryOpReduce(
            reducer.encode(), op.encode(), G.handle, int(A_target), int(B_target),
            A.ptr, B.ptr, out.ptr, ar.ptr, br.ptr, orr.ptr, _stream(G)))


def backward_lhs_binary_op_reduce(reducer, op, G, A_target, B_target, A, B, out, grad_out,
                                  grad_A, A_rows=None, B_rows=None, out_rows=None):
    """kernel.py:172-235 -> kernel._CAPI_DGLKernelBackwardLhsBinaryOpReduce."""
    A, B, out, grad_out, grad_A = _nd(A), _nd(B), _nd(out), _nd(grad_out), _nd(grad_A)
    ar, br, orr = _map(A_rows), _map(B_rows), _map(out_rows)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelBackwardLhsBinaryOpReduce(
            reducer.encode(), op.encode(), G.handle, int(A_target), int(B_target),
            ar.ptr, br.ptr, orr.ptr, A.ptr, B.ptr, out.ptr, grad_out.ptr, grad_A.ptr,
            _stream(G)))


def backward_rhs_binary_op_reduce(reducer, op, G, A_target, B_target, A, B, out, grad_out,
                                  grad_B, A_rows=None, B_rows=None, out_rows=None):
    """kernel.py:237-300 -> kernel._CAPI_DGLKernelBackwardRhsBinaryOpReduce."""
    A, B, out, grad_out, grad_B = _nd(A), _nd(B), _nd(out), _nd(grad_out), _nd(grad_B)
    ar, br, orr = _map(A_rows), _map(B_rows), _map(out_rows)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelBackwardRhsBinaryOpReduce(
            reducer.encode(), op.encode(), G.handle, int(A_target), int(B_target),
            ar.ptr, br.ptr, orr.ptr, A.ptr, B.ptr, out.ptr, grad_out.ptr, grad_B.ptr,
            _stream(G)))


def copy_reduce(reducer, G, target, X, out, X_rows=None, out_rows=None):
    """kernel.py:302-394 -> kernel._CAPI_DGLKernelCopyReduce."""
    X, out = _nd(X), _nd(out)
    xr, orr = _map(X_rows), _map(out_rows)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelCopyReduce(reducer.encode(), G.handle, int(target), X.ptr,
                                               out.ptr, xr.ptr, orr.ptr, _stream(G)))


def backward_copy_reduce(reducer, G, target, X, out, grad_out, grad_X, X_rows=None,
                         out_rows=None):
    """kernel.py:396-435 -> kernel._CAPI_DGLKernelBackwardCopyReduce."""
    X, out, grad_out, grad_X = _nd(X), _nd(out), _nd(grad_out), _nd(grad_X)
    xr, orr = _map(X_rows), _map(out_rows)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelBackwardCopyReduce(
            reducer.encode(), G.handle, int(target), X.ptr, out.ptr, grad_out.ptr, grad_X.ptr,
            xr.ptr, orr.ptr, _stream(G)))


def fused_gat_kernel(G, feat_src, el, er, s, exp, ret, slope):
    """python/dgl/kernel.py fused_gat_kernel -> kernel._CAPI_DGLFusedGatKernel."""
    t = [_nd(x) for x in (feat_src, el, er, s, exp, ret)]
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200FusedGatKernel(G.handle, *[x.ptr for x in t], float(slope),
                                             _stream(G)))


def backward_fused_gat(G, feat_src, el, er, s, exp, ret, grad_out, grad_feat_src, grad_el,
                       grad_er, slope):
    """python/dgl/kernel.py backward_fused_gat -> kernel._CAPI_DGLKernelBackwardFusedGat."""
    t = [_nd(x) for x in (feat_src, el, er, s, exp, ret, grad_out, grad_feat_src, grad_el,
                          grad_er)]
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelBackwardFusedGat(G.handle, *[x.ptr for x in t],
                                                     float(slope), _stream(G)))


def rgcn_layer0(G, weight, norm, ret):
    """python/dgl/kernel.py rgcn_layer0 -> kernel._CAPI_DGLRgcnLayer0."""
    weight, norm, ret = _nd(weight), _nd(norm), _nd(ret)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200RgcnLayer0(G.handle, weight.ptr, norm.ptr, ret.ptr, _stream(G)))


def rgcn_layer0_backward(G, grad_out, norm, grad_weight, accumulate=True):
    """python/dgl/kernel.py rgcn_layer0_backward -> kernel._CAPI_DGLRgcnLayer0Backward.
    accumulate=False reproduces the reference kernel's plain store (last edge of a (type, src)
    run wins, binary_reduce_impl.cu:1004); the default sums, which is the gradient."""
    grad_out, norm, grad_weight = _nd(grad_out), _nd(norm), _nd(grad_weight)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200RgcnLayer0Backward(G.handle, grad_out.ptr, norm.ptr, grad_weight.ptr,
                                                 int(bool(accumulate)), _stream(G)))


def rgcn_layer1(G, hidden, weight, norm, ret):
    """python/dgl/kernel.py rgcn_layer1 -> kernel._CAPI_DGLRgcnLayer1."""
    t = [_nd(x) for x in (hidden, weight, norm, ret)]
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200RgcnLayer1(G.handle, *[x.ptr for x in t], _stream(G)))


def rgcn_layer1_backward(G, hidden, weight, norm, grad_out, grad_hidden, grad_weight):
    """python/dgl/kernel.py rgcn_layer1_backward -> kernel._CAPI_DGLRgcnLayer1Backward."""
    t = [_nd(x) for x in (hidden, weight, norm, grad_out, grad_hidden, grad_weight)]
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200RgcnLayer1Backward(G.handle, *[x.ptr for x in t], _stream(G)))


def copy_reduce_mean(G, target, X, out):
    """Extension (no reference PackedFunc): copy + mean in one pass, see
    DGLB200KernelCopyReduceMean in include/dgl_b200.h."""
    X, out = _nd(X), _nd(out)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelCopyReduceMean(G.handle, int(target), X.ptr, out.ptr,
                                                   _stream(G)))


last_kernel = _lib.last_kernel
launch_count = _lib.launch_count


def edge_softmax_forward(G, score, out):
    """Extension: nn/pytorch/softmax.py EdgeSoftmax.forward as one kernel."""
    t = [_nd(x) for x in (score, out)]
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200EdgeSoftmaxForward(G.handle, t[0].ptr, t[1].ptr, _stream(G)))


def edge_softmax_backward(G, out, grad_out, grad_score):
    """Extension: nn/pytorch/softmax.py EdgeSoftmax.backward as one kernel."""
    t = [_nd(x) for x in (out, grad_out, grad_score)]
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200EdgeSoftmaxBackward(G.handle, t[0].ptr, t[1].ptr, t[2].ptr,
                                                  _stream(G)))


def _keep(*xs):
    """NDArray views (kept alive by the caller for the duration of the call) and their pointers;
    None stays NULL."""
    nds = [None if x is None else _nd(x) for x in xs]
    return nds, [None if t is None else t.ptr for t in nds]


def gat_attn_proj(feat, attn_l, attn_r, el, er):
    """Extension: el = (feat * attn_l).sum(-1), er = (feat * attn_r).sum(-1) (gatconv.py:143-144)
    in one pass; a None attention vector / output skips that projection."""
    keep, ptrs = _keep(feat, attn_l, attn_r, el, er)
    with torch.cuda.device(feat.device):
        s = ctypes.c_void_p(torch.cuda.current_stream(feat.device).cuda_stream)
        check_call(LIB.DGLB200GatAttnProj(*ptrs, s))
    del keep


def gat_attn_proj_backward(feat, attn_l, attn_r, grad_el, grad_er, grad_feat, grad_attn_l,
                           grad_attn_r):
    keep, ptrs = _keep(feat, attn_l, attn_r, grad_el, grad_er, grad_feat, grad_attn_l, grad_attn_r)
    with torch.cuda.device(feat.device):
        s = ctypes.c_void_p(torch.cuda.current_stream(feat.device).cuda_stream)
        check_call(LIB.DGLB200GatAttnProjBackward(*ptrs, s))
    del keep


def scaled_copy_sum(G, transpose, X, col_scale, row_scale, out, addend=None):
    """Extension: out[r] = row_scale[r] * sum over r's edges of col_scale[c] * X[c] on the in-CSR
    (transpose False) or the out-CSR (True); GraphConv's normalisation inside the SpMM. With
    `addend` (out's shape; may be `out` itself): out = row_scale * (addend + sum)."""
    if addend is not None:
        keep, ptrs = _keep(X, col_scale, row_scale, addend, out)
        with torch.cuda.device(G.device):
            check_call(LIB.DGLB200KernelScaledCopySumAdd(G.handle, int(bool(transpose)), *ptrs,
                                                         _stream(G)))
        del keep
        return
    keep, ptrs = _keep(X, col_scale, row_scale, out)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelScaledCopySum(G.handle, int(bool(transpose)), *ptrs,
                                                  _stream(G)))
    del keep


def peer_copy_sum(G, transpose, slot_bases, node_code, row_scale, out):
    """Extension: copy_src + sum over a feature matrix spread over several GPUs' memories
    (DGLB200KernelPeerCopySum): gathered node c reads row (code & (2^27-1)) of block
    slot_bases[code >> 27], code = node_code[c]; slot_bases are raw device pointers (own or
    peer-mapped buffers, row pitch = out's width)."""
    keep, ptrs = _keep(node_code, row_scale, out)
    k = len(slot_bases)
    bases = (ctypes.c_void_p * k)(*[int(b) for b in slot_bases])
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelPeerCopySum(G.handle, int(bool(transpose)), k, bases, *ptrs,
                                                _stream(G)))
    del keep


def dense_tf32x3(a, b, b_is_kn, bias, relu, c):
    """Extension: c = a @ (b if b_is_kn else b.T) (+ bias) (ReLU) on the tensor cores at fp32
    accuracy (DGLB200DenseTF32x3, csrc/gemm.cu: tcgen05 + TMEM + TMA, 3xTF32)."""
    keep, ptrs = _keep(a, b, bias, c)
    with torch.cuda.device(a.device):
        s = ctypes.c_void_p(torch.cuda.current_stream(a.device).cuda_stream)
        check_call(LIB.DGLB200DenseTF32x3(ptrs[0], ptrs[1], int(bool(b_is_kn)), ptrs[2],
                                          int(bool(relu)), ptrs[3], s))
    del keep


def copy_sum_bf16(G, transpose, X, col_scale, row_scale, mean, out):
    """Extension: copy_src + sum / mean over bf16 feature rows with fp32 accumulation
    (DGLB200KernelCopySumBF16); X / out are torch.bfloat16 [rows, D], D % 8 == 0."""
    keep, ptrs = _keep(X, col_scale, row_scale, out)
    with torch.cuda.device(G.device):
        check_call(LIB.DGLB200KernelCopySumBF16(G.handle, int(bool(transpose)), ptrs[0], ptrs[1],
                                                ptrs[2], int(bool(mean)), ptrs[3], _stream(G)))
    del keep


def grouped_dense_tf32x3(a, b, b_is_kn, group_ptr, c):
    """Extension: rows [group_ptr[g], group_ptr[g+1]) of a times b[g] (or b[g].T) -> c, on the
    tensor cores (DGLB200GroupedDenseTF32x3); group_ptr = python list of len(b) + 1 offsets."""
    keep, ptrs = _keep(a, b, c)
    gp = (ctypes.c_int64 * len(group_ptr))(*[int(v) for v in group_ptr])
    with torch.cuda.device(a.device):
        s = ctypes.c_void_p(torch.cuda.current_stream(a.device).cuda_stream)
        check_call(LIB.DGLB200GroupedDenseTF32x3(ptrs[0], ptrs[1], int(bool(b_is_kn)), gp,
                                                 len(group_ptr) - 1, ptrs[2], s))
    del keep


def dense_tf32x3_weight_grad(a, g, c):
    """Extension: c[K, N] = a[M, K].T @ g[M, N] on the tensor cores (DGLB200DenseTF32x3WeightGrad)."""
    keep, ptrs = _keep(a, g, c)
    with torch.cuda.device(a.device):
        s = ctypes.c_void_p(torch.cuda.current_stream(a.device).cuda_stream)
        check_call(LIB.DGLB200DenseTF32x3WeightGrad(ptrs[0], ptrs[1], ptrs[2], s))
    del keep

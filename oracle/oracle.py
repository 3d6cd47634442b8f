"""TEST INFRASTRUCTURE ONLY -- numpy front end of the CPU oracle (oracle/dgl_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.  Parity status: PINNED (see the header of
dgl_oracle.c).

The Python here restates the *host-side* logic that sits above the per-edge arithmetic in the
reference:
  * CalcBcastInfo            src/kernel/binary_reduce.cc:96-155
  * NeedSwitchOrder + swap   src/kernel/binary_reduce.cc:214-219, 313-317, 476-482, 574-580
  * BinaryOpReduce / CopyReduce / Backward* argument plumbing  binary_reduce.cc:295-336,
    452-500, 551-598, 628-647, 667-695
  * in/out CSR materialisation  src/graph/immutable_graph.cc:407-437
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libdgl_oracle.so")
_SRC = os.path.join(_HERE, "dgl_oracle.c")

T_SRC, T_DST, T_EDGE, T_NONE = 0, 1, 2, 3
OPS = {"add": 0, "sub": 1, "mul": 2, "div": 3, "dot": 4, "use_lhs": 5}
REDUCERS = {"sum": 0, "max": 1, "min": 2, "prod": 3, "none": 4}


def build(force=False):
    """Compile the C restatement with gcc (seconds)."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-std=c99",
                               "-fno-fast-math", "-ffp-contract=off", "-o", _SO, _SRC, "-lm"])
    return _SO


class _Bcast(ctypes.Structure):
    _fields_ = [("ndim", ctypes.c_int32),
                ("lhs_shape", ctypes.c_int64 * 8), ("lhs_stride", ctypes.c_int64 * 8),
                ("rhs_shape", ctypes.c_int64 * 8), ("rhs_stride", ctypes.c_int64 * 8),
                ("out_shape", ctypes.c_int64 * 8), ("out_stride", ctypes.c_int64 * 8),
                ("lhs_len", ctypes.c_int64), ("rhs_len", ctypes.c_int64),
                ("out_len", ctypes.c_int64), ("data_len", ctypes.c_int64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _i64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int64)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


# ------------------------------------------------------------------ index prep
def coo_to_csr(n_rows, row, col):
    """aten::COOToCSR (spmat_op_impl_coo.cc:230-283) -> (indptr, indices, data)."""
    row, col = _i64(row), _i64(col)
    nnz = row.shape[0]
    indptr = np.empty(n_rows + 1, np.int64)
    indices = np.empty(nnz, np.int64)
    data = np.empty(nnz, np.int64)
    lib().orc_coo_to_csr(ctypes.c_int64(n_rows), ctypes.c_int64(nnz), _p(row), _p(col),
                         _p(indptr), _p(indices), _p(data))
    return indptr, indices, data


def csr_transpose(n_rows, n_cols, indptr, indices, data=None):
    """aten::CSRTranspose (spmat_op_impl.cc:323-369)."""
    indptr, indices, data = _i64(indptr), _i64(indices), _i64(data)
    nnz = indices.shape[0]
    bp = np.empty(n_cols + 1, np.int64)
    bi = np.empty(nnz, np.int64)
    bx = np.empty(nnz, np.int64)
    lib().orc_csr_transpose(ctypes.c_int64(n_rows), ctypes.c_int64(n_cols), _p(indptr),
                            _p(indices), _p(data), _p(bp), _p(bi), _p(bx))
    return bp, bi, bx


def csr_to_coo_rows(indptr):
    indptr = _i64(indptr)
    n = indptr.shape[0] - 1
    row = np.empty(int(indptr[-1]), np.int64)
    lib().orc_csr_to_coo_rows(ctypes.c_int64(n), _p(indptr), _p(row))
    return row


class Graph(object):
    """In-CSR + out-CSR of a (num_src x num_dst) graph given as COO, materialised in the order
    the reference's kernel path does it: GraphIndex.get_immutable_gidx
    (python/dgl/graph_index.py:672-686) -> ImmutableGraph::AsNumBits / CopyTo
    (src/graph/immutable_graph.cc:548-559, 574-590) call GetInCSR() FIRST, which on a COO-only
    graph is coo.Transpose().ToCSR() (immutable_graph.cc:407-421: COOToCSR keyed by dst, stable
    in edge-id order), and THEN GetOutCSR(), which now finds in_csr_ and returns
    in_csr.Transpose() (immutable_graph.cc:424-437: within an out-row, dst ascending, parallel
    edges in edge-id order). UnitGraph does the same (src/graph/unit_graph.cc GetInCSR/GetOutCSR).
    Checked bit-exactly against _CAPI_DGLGraphGetAdj / _CAPI_DGLHeteroGetAdj outputs in
    tests/golden/*.npz."""

    def __init__(self, num_src, num_dst, src, dst):
        self.num_src, self.num_dst = int(num_src), int(num_dst)
        self.src, self.dst = _i64(src), _i64(dst)
        self.num_edges = self.src.shape[0]
        self.in_csr = coo_to_csr(self.num_dst, self.dst, self.src)
        self.out_csr = csr_transpose(self.num_dst, self.num_src, *self.in_csr)

    def in_degrees(self):
        return np.diff(self.in_csr[0])

    def out_degrees(self):
        return np.diff(self.out_csr[0])


# ------------------------------------------------------------------ broadcast info
def _stride(shape):
    st = [1] * len(shape)
    for i in range(len(shape) - 2, -1, -1):
        st[i] = st[i + 1] * shape[i + 1]
    return st


def calc_bcast_info(op, lhs_shape, rhs_shape):
    """CalcBcastInfo (binary_reduce.cc:96-155). Shapes include the leading row dim."""
    lhs_shape, rhs_shape = list(lhs_shape), list(rhs_shape)
    ln, rn = len(lhs_shape), len(rhs_shape)
    max_ndim = max(ln, rn) - 1
    accum, j = 0, 0
    ls, rs, os_, real = [], [], [], []
    if op == "dot":
        data_len = lhs_shape[-1]
        j = 1
        real.append(data_len)
    else:
        data_len = 1
    while j < max_ndim:
        dl = 1 if ln - 1 - j < 1 else lhs_shape[ln - 1 - j]
        dr = 1 if rn - 1 - j < 1 else rhs_shape[rn - 1 - j]
        if dl != dr:
            if dl != 1 and dr != 1:
                raise ValueError("Invalid broadcasting between feature shapes")
            if accum != 0:
                ls.append(accum); rs.append(accum); os_.append(accum)
                accum = 0
            ls.append(dl); rs.append(dr); os_.append(max(dl, dr))
        else:
            accum = dl if accum == 0 else accum * dl
        real.append(max(dl, dr))
        j += 1
    if accum != 0:
        ls.append(accum); rs.append(accum); os_.append(accum)
    real.reverse(); ls.reverse(); rs.reverse(); os_.reverse()
    return dict(data_len=data_len, real_out_shape=tuple(real), lhs_shape=ls, rhs_shape=rs,
                out_shape=os_, lhs_stride=_stride(ls), rhs_stride=_stride(rs),
                out_stride=_stride(os_))


def infer_binary_feature_shape(op, lhs_shape, rhs_shape):
    """InferBinaryFeatureShape (binary_reduce.cc:274-279)."""
    return calc_bcast_info(op, lhs_shape, rhs_shape)["real_out_shape"]


def _bcast_struct(op, lhs, rhs):
    b = _Bcast()
    if rhs is None:  # CopyReduce: BinaryReduceImpl with use_lhs, x_len = prod(out feat)
        d = int(np.prod(lhs.shape[1:], dtype=np.int64))
        shp = dict(data_len=1, lhs_shape=[d], rhs_shape=[d], out_shape=[d],
                   lhs_stride=[1], rhs_stride=[1], out_stride=[1])
    else:
        same = tuple(lhs.shape[1:]) == tuple(rhs.shape[1:])
        if same:  # non-broadcast path, binary_reduce.cc:325-334 / binary_reduce_impl.h:54-60
            feat = list(lhs.shape[1:])
            dl = feat[-1] if (op == "dot" and feat) else 1
            d = int(np.prod(feat, dtype=np.int64)) // dl
            shp = dict(data_len=dl, lhs_shape=[d], rhs_shape=[d], out_shape=[d],
                       lhs_stride=[1], rhs_stride=[1], out_stride=[1])
        else:
            shp = calc_bcast_info(op, lhs.shape, rhs.shape)
    n = len(shp["out_shape"])
    if n > 8:
        raise ValueError("Too many broadcasting dimensions.")
    b.ndim = n
    for i in range(n):
        b.lhs_shape[i] = shp["lhs_shape"][i]; b.lhs_stride[i] = shp["lhs_stride"][i]
        b.rhs_shape[i] = shp["rhs_shape"][i]; b.rhs_stride[i] = shp["rhs_stride"][i]
        b.out_shape[i] = shp["out_shape"][i]; b.out_stride[i] = shp["out_stride"][i]
    b.lhs_len = int(np.prod(shp["lhs_shape"], dtype=np.int64))
    b.rhs_len = int(np.prod(shp["rhs_shape"], dtype=np.int64))
    b.out_len = int(np.prod(shp["out_shape"], dtype=np.int64))
    b.data_len = shp["data_len"]
    return b


def _need_switch(op, lhs_tgt, rhs_tgt):
    if lhs_tgt == rhs_tgt:
        raise ValueError("Check failed: lhs != rhs")
    return op in ("add", "mul") and lhs_tgt > rhs_tgt


# ------------------------------------------------------------------ kernels
def out_feat_shape(op, lhs, rhs):
    """Feature shape of the output tensor as the torch backend allocates it
    (python/dgl/backend/pytorch/tensor.py:371-376): dot drops the reduced last dim."""
    feat = infer_binary_feature_shape(op, lhs.shape, rhs.shape)
    return feat[:-1] if op == "dot" else feat


def binary_op_reduce(reducer, op, g, lhs_tgt, rhs_tgt, lhs, rhs, out_size,
                     lhs_map=None, rhs_map=None, out_map=None):
    """BinaryOpReduce (binary_reduce.cc:295-336). Returns out [out_size, *feat]."""
    lhs, rhs = _f32(lhs), _f32(rhs)
    feat = out_feat_shape(op, lhs, rhs)
    if _need_switch(op, lhs_tgt, rhs_tgt):
        lhs_tgt, rhs_tgt, lhs, rhs = rhs_tgt, lhs_tgt, rhs, lhs
        lhs_map, rhs_map = rhs_map, lhs_map
    bc = _bcast_struct(op, lhs, rhs)
    out = np.empty((out_size,) + tuple(feat), np.float32)
    indptr, indices, eids = g.out_csr
    lm, rm, om = _i64(lhs_map), _i64(rhs_map), _i64(out_map)
    lib().orc_binary_reduce(REDUCERS[reducer], OPS[op], lhs_tgt, rhs_tgt,
                            ctypes.c_int64(g.num_src), _p(indptr), _p(indices), _p(eids),
                            _p(lhs), _p(rhs), _p(out), ctypes.c_int64(out_size),
                            _p(lm), _p(rm), _p(om), ctypes.byref(bc))
    return out


def copy_reduce(reducer, g, target, in_data, out_size, in_map=None, out_map=None):
    """CopyReduce (binary_reduce.cc:628-647)."""
    in_data = _f32(in_data)
    bc = _bcast_struct("use_lhs", in_data, None)
    out = np.empty((out_size,) + in_data.shape[1:], np.float32)
    indptr, indices, eids = g.out_csr
    im, om = _i64(in_map), _i64(out_map)
    lib().orc_binary_reduce(REDUCERS[reducer], OPS["use_lhs"], target, T_NONE,
                            ctypes.c_int64(g.num_src), _p(indptr), _p(indices), _p(eids),
                            _p(in_data), None, _p(out), ctypes.c_int64(out_size),
                            _p(im), None, _p(om), ctypes.byref(bc))
    return out


def _backward(mode, reducer, op, g, lhs_tgt, rhs_tgt, lhs, rhs, out, grad_out,
              lhs_map, rhs_map, out_map):
    lhs, rhs, out, grad_out = _f32(lhs), _f32(rhs), _f32(out), _f32(grad_out)
    if rhs is not None and _need_switch(op, lhs_tgt, rhs_tgt):
        # Backward{Lhs,Rhs}BinaryOpReduce swap operands AND the requested side
        # (binary_reduce.cc:476-482, 574-580)
        lhs_tgt, rhs_tgt, lhs, rhs = rhs_tgt, lhs_tgt, rhs, lhs
        lhs_map, rhs_map = rhs_map, lhs_map
        mode = 1 - mode
    bc = _bcast_struct(op, lhs, rhs)
    rows = lhs.shape[0] if mode == 0 else rhs.shape[0]
    if rhs is None:
        feat = tuple(lhs.shape[1:])
    else:
        feat = tuple(infer_binary_feature_shape(op, lhs.shape, rhs.shape))
    grad = np.empty((rows,) + feat, np.float32)
    indptr, indices, eids = g.in_csr
    lm, rm, om = _i64(lhs_map), _i64(rhs_map), _i64(out_map)
    lib().orc_backward_binary_reduce(
        mode, REDUCERS[reducer], OPS[op], lhs_tgt, rhs_tgt, ctypes.c_int64(g.num_dst),
        _p(indptr), _p(indices), _p(eids), _p(lhs), _p(rhs), _p(out), _p(grad_out), _p(grad),
        ctypes.c_int64(rows), _p(lm), _p(rm), _p(om), ctypes.byref(bc))
    return grad


def backward_lhs_binary_op_reduce(reducer, op, g, lhs_tgt, rhs_tgt, lhs, rhs, out, grad_out,
                                  lhs_map=None, rhs_map=None, out_map=None):
    """BackwardLhsBinaryOpReduce (binary_reduce.cc:452-500). Returns the UNREDUCED
    grad_lhs [lhs_rows, *out_feat] (tensor.py:351 then _reduce_grad's it)."""
    return _backward(0, reducer, op, g, lhs_tgt, rhs_tgt, lhs, rhs, out, grad_out,
                     lhs_map, rhs_map, out_map)


def backward_rhs_binary_op_reduce(reducer, op, g, lhs_tgt, rhs_tgt, lhs, rhs, out, grad_out,
                                  lhs_map=None, rhs_map=None, out_map=None):
    """BackwardRhsBinaryOpReduce (binary_reduce.cc:551-598)."""
    return _backward(1, reducer, op, g, lhs_tgt, rhs_tgt, lhs, rhs, out, grad_out,
                     lhs_map, rhs_map, out_map)


def backward_copy_reduce(reducer, g, target, in_data, out, grad_out, in_map=None,
                         out_map=None):
    """BackwardCopyReduce (binary_reduce.cc:667-695)."""
    return _backward(0, reducer, "use_lhs", g, target, T_NONE, in_data, None, out, grad_out,
                     in_map, None, out_map)


def reduce_grad(grad, shape):
    """_reduce_grad (python/dgl/backend/pytorch/tensor.py:572-601) in numpy."""
    grad_shape = grad.shape[1:]
    in_shape = tuple(shape[1:])
    if in_shape == grad_shape:
        return grad
    pad = (1,) * (len(grad_shape) - len(in_shape)) + in_shape
    axes = tuple(i + 1 for i, (a, b) in enumerate(zip(grad_shape, pad)) if a != b)
    return grad.sum(axis=axes, keepdims=True).reshape(shape)


# ------------------------------------------------------------------ fused GAT (hack ops)
def fused_gat_forward(g, ft, el, er, slope):
    """FusedGatKernelImpl (src/kernel/cuda/binary_reduce_impl.cu:51-113, 359-422).
    ft [N,H,F], el/er [N,H,1] -> (sum [N,H,1], exp [E,H,1], ret [N,H,F])."""
    ft, el, er = _f32(ft), _f32(el), _f32(er)
    H = int(np.prod(el.shape[1:]))
    F = int(np.prod(ft.shape[1:])) // H
    s = np.empty((g.num_dst,) + el.shape[1:], np.float32)
    ex = np.empty((g.num_edges,) + el.shape[1:], np.float32)
    ret = np.empty((g.num_dst,) + ft.shape[1:], np.float32)
    ip, ix, ei = g.in_csr
    lib().orc_fused_gat_forward(ctypes.c_int64(g.num_dst), _p(ip), _p(ix), _p(ei),
                                ctypes.c_int64(H), ctypes.c_int64(F), _p(ft), _p(el), _p(er),
                                ctypes.c_float(slope), _p(s), _p(ex), _p(ret))
    return s, ex, ret


def fused_gat_backward(g, ft, el, er, s, ex, ret, gout, slope):
    """BackwardFusedGatKernelImpl (binary_reduce_impl.cu:123-213, 1248-1308)."""
    ft, el, er, s, ex, ret, gout = [_f32(a) for a in (ft, el, er, s, ex, ret, gout)]
    H = int(np.prod(el.shape[1:]))
    F = int(np.prod(ft.shape[1:])) // H
    gft, gel, ger = np.empty_like(ft), np.empty_like(el), np.empty_like(er)
    ip, ix, ei = g.out_csr
    lib().orc_fused_gat_backward(ctypes.c_int64(g.num_src), ctypes.c_int64(g.num_dst), _p(ip),
                                 _p(ix), _p(ei), ctypes.c_int64(H), ctypes.c_int64(F), _p(ft),
                                 _p(el), _p(er), _p(s), _p(ex), _p(ret), _p(gout),
                                 ctypes.c_float(slope), _p(gft), _p(gel), _p(ger))
    return gft, gel, ger


# ------------------------------------------------------------------ RGCN hack ops (a15)
# PARITY UNPINNED: CUDA-only kernels in the reference, no reference test, and the fork's library
# cannot link them on CPU; restated from the CUDA source, cross-checked in tests against a dense
# numpy formula and the builtin composition (RelGraphConv-style u_mul_e + sum per relation).
def typed_csr(g, etypes, transpose):
    """Graph::GetCsrSortedByEdgeType(transpose) (src/graph/graph.cc:690-741) ->
    (ranges, ids, eids, types), rows ordered by (type, edge id)."""
    et = _i64(etypes)
    key, other = (g.src, g.dst) if transpose else (g.dst, g.src)
    n_rows = g.num_src if transpose else g.num_dst
    indptr = np.empty(n_rows + 1, np.int64)
    ids, eids, types = (np.empty(g.num_edges, np.int64) for _ in range(3))
    lib().orc_typed_csr(ctypes.c_int64(n_rows), ctypes.c_int64(g.num_edges), _p(key), _p(other),
                        _p(et), _p(indptr), _p(ids), _p(eids), _p(types))
    return indptr, ids, eids, types


def rgcn_layer0(g, etypes, weight, norm):
    """RgcnLayer0Impl (binary_reduce_impl.cu:913-975). weight [R,N,F], norm [E,1] -> [N,F]."""
    weight, norm = _f32(weight), _f32(norm).reshape(-1)
    ip, ids, eids, ty = typed_csr(g, etypes, False)
    ret = np.empty((g.num_dst, weight.shape[2]), np.float32)
    lib().orc_rgcn_layer0(ctypes.c_int64(g.num_dst), _p(ip), _p(ids), _p(eids), _p(ty),
                          ctypes.c_int64(weight.shape[1]), ctypes.c_int64(weight.shape[2]),
                          _p(weight), _p(norm), _p(ret))
    return ret


def rgcn_layer0_backward(g, etypes, grad_out, norm, weight_shape, accumulate=True):
    """RgcnLayer0BackwardImpl (binary_reduce_impl.cu:982-1046); accumulate=False = as written."""
    grad_out, norm = _f32(grad_out), _f32(norm).reshape(-1)
    ip, ids, eids, ty = typed_csr(g, etypes, True)
    gw = np.zeros(tuple(weight_shape), np.float32)
    lib().orc_rgcn_layer0_backward(ctypes.c_int64(g.num_src), _p(ip), _p(ids), _p(eids), _p(ty),
                                   ctypes.c_int64(weight_shape[1]),
                                   ctypes.c_int64(weight_shape[2]), _p(grad_out), _p(norm),
                                   ctypes.c_int(int(bool(accumulate))), _p(gw))
    return gw


def rgcn_layer1(g, etypes, hidden, weight, norm):
    """RgcnLayer1Impl (binary_reduce_impl.cu:1082-1155). hidden [N,in], weight [R,in,out]."""
    hidden, weight, norm = _f32(hidden), _f32(weight), _f32(norm).reshape(-1)
    ip, ids, eids, ty = typed_csr(g, etypes, False)
    ret = np.empty((g.num_dst, weight.shape[2]), np.float32)
    lib().orc_rgcn_layer1(ctypes.c_int64(g.num_dst), _p(ip), _p(ids), _p(eids), _p(ty),
                          ctypes.c_int64(weight.shape[1]), ctypes.c_int64(weight.shape[2]),
                          _p(hidden), _p(weight), _p(norm), _p(ret))
    return ret


def rgcn_layer1_backward(g, etypes, hidden, weight, norm, grad_out):
    """RgcnLayer1BackwardImpl (binary_reduce_impl.cu:1157-1246) -> (grad_hidden, grad_weight)."""
    hidden, weight, norm, grad_out = (_f32(hidden), _f32(weight), _f32(norm).reshape(-1),
                                      _f32(grad_out))
    ip, ids, eids, ty = typed_csr(g, etypes, True)
    gh, gw = np.empty_like(hidden), np.empty_like(weight)
    lib().orc_rgcn_layer1_backward(ctypes.c_int64(g.num_src), ctypes.c_int64(weight.shape[0]),
                                   _p(ip), _p(ids), _p(eids), _p(ty),
                                   ctypes.c_int64(weight.shape[1]),
                                   ctypes.c_int64(weight.shape[2]), _p(hidden), _p(weight),
                                   _p(norm), _p(grad_out), _p(gh), _p(gw))
    return gh, gw

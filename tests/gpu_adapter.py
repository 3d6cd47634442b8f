"""Adapter exposing the CUDA product (through the C ABI) via the golden_util replay interface:
numpy in, numpy out, same call meaning as the oracle adapter."""
import numpy as np
import torch

import dgl_hack_b200 as B
from dgl_hack_b200 import kernel as K

DEV = "cuda:0"


def _t(a):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(DEV)


def _m(a):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a).astype(np.int32)).to(DEV)


def _shape(op, lhs, rhs):
    feat = K.infer_binary_feature_shape(op, lhs, rhs)
    return feat, (feat[:-1] if op == "dot" else feat)


class GpuImpl(object):
    last = None

    @staticmethod
    def make_graph(ns, nd, src, dst):
        return B.GraphIndex.from_coo(ns, nd, np.asarray(src, np.int64), np.asarray(dst, np.int64),
                                     DEV)

    @staticmethod
    def binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out_rows, lm, rm, om):
        lhs, rhs = _t(lhs), _t(rhs)
        _, oshape = _shape(op, lhs, rhs)
        out = torch.full((out_rows,) + oshape, float("nan"), device=DEV)
        K.binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out, _m(lm), _m(rm), _m(om))
        GpuImpl.last = K.last_kernel()
        return out.cpu().numpy()

    @staticmethod
    def _bwd(fn, red, op, g, lt, rt, lhs, rhs, out, gout, lm, rm, om, rows_of):
        lhs, rhs = _t(lhs), _t(rhs)
        feat, _ = _shape(op, lhs, rhs)
        rows = (lhs if rows_of == 0 else rhs).shape[0]
        grad = torch.full((rows,) + feat, float("nan"), device=DEV)
        fn(red, op, g, lt, rt, lhs, rhs, _t(out), _t(gout), grad, _m(lm), _m(rm), _m(om))
        GpuImpl.last = K.last_kernel()
        return grad.cpu().numpy()

    @staticmethod
    def backward_lhs_binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out, gout, lm, rm, om):
        return GpuImpl._bwd(K.backward_lhs_binary_op_reduce, red, op, g, lt, rt, lhs, rhs, out,
                            gout, lm, rm, om, 0)

    @staticmethod
    def backward_rhs_binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out, gout, lm, rm, om):
        return GpuImpl._bwd(K.backward_rhs_binary_op_reduce, red, op, g, lt, rt, lhs, rhs, out,
                            gout, lm, rm, om, 1)

    @staticmethod
    def copy_reduce(red, g, tgt, x, out_rows, im, om):
        x = _t(x)
        out = torch.full((out_rows,) + tuple(x.shape[1:]), float("nan"), device=DEV)
        K.copy_reduce(red, g, tgt, x, out, _m(im), _m(om))
        GpuImpl.last = K.last_kernel()
        return out.cpu().numpy()

    @staticmethod
    def backward_copy_reduce(red, g, tgt, x, out, gout, im, om):
        x = _t(x)
        grad = torch.full(tuple(x.shape), float("nan"), device=DEV)
        K.backward_copy_reduce(red, g, tgt, x, _t(out), _t(gout), grad, _m(im), _m(om))
        GpuImpl.last = K.last_kernel()
        return grad.cpu().numpy()

"""Adapter exposing the CPU oracle through the golden_util replay interface."""
from oracle import oracle as O


class OracleImpl(object):
    make_graph = staticmethod(O.Graph)

    @staticmethod
    def binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out_rows, lm, rm, om):
        return O.binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out_rows, lm, rm, om)

    @staticmethod
    def backward_lhs_binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out, gout, lm, rm, om):
        return O.backward_lhs_binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out, gout,
                                               lm, rm, om)

    @staticmethod
    def backward_rhs_binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out, gout, lm, rm, om):
        return O.backward_rhs_binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out, gout,
                                               lm, rm, om)

    @staticmethod
    def copy_reduce(red, g, tgt, x, out_rows, im, om):
        return O.copy_reduce(red, g, tgt, x, out_rows, im, om)

    @staticmethod
    def backward_copy_reduce(red, g, tgt, x, out, gout, im, om):
        return O.backward_copy_reduce(red, g, tgt, x, out, gout, im, om)

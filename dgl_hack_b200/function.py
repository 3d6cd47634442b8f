"""Builtin message / reduce function descriptors (`dgl.function`).

Same names, argument order and meaning as the reference's python/dgl/function/message.py:101-302
(copy_u / copy_e / copy_src / copy_edge / src_mul_edge and the generated `{u,v,e}_{add,sub,mul,
div,dot}_{u,v,e}` family) and function/reducer.py:60-96 (sum / max / min / mean / prod). In the
reference a descriptor's `_invoke` appends nodes to the scheduler IR (runtime/ir); here a
descriptor is plain data and `Graph.update_all / apply_edges` (graph.py) turn a
(message, reduce) pair straight into ONE `binary_reduce` / `copy_reduce` kernel call -- the
§8(f) "scheduler bypass" row. Target codes: python/dgl/function/base.py:7-21.
"""
import sys

__all__ = ["TargetCode", "copy_u", "copy_e", "copy_src", "copy_edge", "src_mul_edge"]


class TargetCode(object):
    SRC = 0
    DST = 1
    EDGE = 2

    CODE2STR = {0: "u", 1: "v", 2: "e"}


_STR2CODE = {"u": TargetCode.SRC, "v": TargetCode.DST, "e": TargetCode.EDGE}


class BuiltinFunction(object):
    @property
    def name(self):
        raise NotImplementedError


class MessageFunction(BuiltinFunction):
    pass


class BinaryMessageFunction(MessageFunction):
    """out[e] = lhs_field[lhs end of e] (op) rhs_field[rhs end of e]."""

    def __init__(self, binary_op, lhs, rhs, lhs_field, rhs_field, out_field):
        self.binary_op = binary_op
        self.lhs, self.rhs = lhs, rhs
        self.lhs_field, self.rhs_field = lhs_field, rhs_field
        self.out_field = out_field

    @property
    def name(self):
        return "{}_{}_{}".format(TargetCode.CODE2STR[self.lhs], self.binary_op,
                                 TargetCode.CODE2STR[self.rhs])


class CopyMessageFunction(MessageFunction):
    """out[e] = in_field[target end of e]."""

    def __init__(self, target, in_field, out_field):
        self.target = target
        self.in_field = in_field
        self.out_field = out_field

    @property
    def name(self):
        return "copy_{}".format(TargetCode.CODE2STR[self.target])


class ReduceFunction(BuiltinFunction):
    pass


class SimpleReduceFunction(ReduceFunction):
    """out_field[v] = reduce over the in-edges of v of msg_field."""

    def __init__(self, name, msg_field, out_field):
        self._name = name
        self.msg_field = msg_field
        self.out_field = out_field

    @property
    def name(self):
        return self._name


def copy_u(u, out):
    return CopyMessageFunction(TargetCode.SRC, u, out)


def copy_e(e, out):
    return CopyMessageFunction(TargetCode.EDGE, e, out)


def copy_src(src, out):
    return copy_u(src, out)


def copy_edge(edge, out):
    return copy_e(edge, out)


def _gen_message_builtin(lhs, rhs, binary_op):
    name = "{}_{}_{}".format(lhs, binary_op, rhs)

    def func(lhs_field, rhs_field, out):
        return BinaryMessageFunction(binary_op, _STR2CODE[lhs], _STR2CODE[rhs], lhs_field,
                                     rhs_field, out)

    func.__name__ = name
    func.__doc__ = "Builtin message function: out = {}[{}] {} {}[{}] on every edge.".format(
        lhs, "lhs_field", binary_op, rhs, "rhs_field")
    return func


def _register():
    mod = sys.modules[__name__]
    for lhs, rhs in (("u", "v"), ("u", "e"), ("v", "u"), ("v", "e"), ("e", "u"), ("e", "v")):
        for op in ("add", "sub", "mul", "div", "dot"):
            f = _gen_message_builtin(lhs, rhs, op)
            setattr(mod, f.__name__, f)
            __all__.append(f.__name__)
    for red in ("max", "min", "sum", "mean", "prod"):
        def make(red):
            def func(msg, out):
                return SimpleReduceFunction(red, msg, out)
            func.__name__ = red
            func.__doc__ = "Builtin reduce function: out[v] = {} of msg over v's in-edges.".format(red)
            return func
        setattr(mod, red, make(red))
        __all__.append(red)


_register()


def src_mul_edge(src, edge, out):
    return getattr(sys.modules[__name__], "u_mul_e")(src, edge, out)

"""Reference-side binding: route DGL 0.4's kernel wrappers to the B200 library.

`install()` rebinds the functions of the reference's `python/dgl/kernel.py` (the module every
backend calls, `python/dgl/backend/pytorch/tensor.py:298-360, 524-562`) to dgl_hack_b200.kernel,
so `update_all / apply_edges / edge_softmax / GraphConv / SAGEConv / GATConv / RelGraphConv` run
on the sm_100a kernels unchanged. Needs the reference's `dgl` package importable (it is not
vendored here); nothing in this package depends on this module.

What changes hands at the boundary:
  * tensors: the reference passes `dgl.ndarray.NDArray` built zero-copy from torch via DLPack
    (`zerocopy_to_dgl_ndarray`, tensor.py:283). We go back to the torch tensor with the same
    zero-copy hop (`to_dlpack` -> `torch.utils.dlpack.from_dlpack`).
  * graph: the reference passes its C++ graph index handle. Its two CSRs are read once per
    (graph, device) through the public PackedFunc `_CAPI_DGLGraphGetAdj(g, transpose, "csr")`
    (src/graph/graph_apis.cc:285-292; `_CAPI_DGLHeteroGetAdj` for unit graphs) and wrapped with
    GraphIndex.from_csr; the result is cached on the Python graph-index object.
"""
import torch
from torch.utils import dlpack as _dlpack

from . import kernel as K
from .graph_index import GraphIndex

_CACHE_ATTR = "_b200_graph_index"


def _to_torch(nd_or_tensor):
    if nd_or_tensor is None or isinstance(nd_or_tensor, torch.Tensor):
        return nd_or_tensor
    if nd_or_tensor.shape[0] == 0 and len(nd_or_tensor.shape) == 1:
        return None  # nd.null()
    return _dlpack.from_dlpack(nd_or_tensor.to_dlpack())


def reference_csrs(gidx):
    """(num_src, num_dst, in_csr, out_csr) of a reference graph index, as numpy int arrays."""
    import dgl  # noqa: F401
    from dgl import graph_index, heterograph_index

    def adj(rst):
        return tuple(rst(i).asnumpy() for i in range(3))

    if isinstance(gidx, graph_index.GraphIndex):
        n = gidx.number_of_nodes()
        return (n, n, adj(graph_index._CAPI_DGLGraphGetAdj(gidx, False, "csr")),
                adj(graph_index._CAPI_DGLGraphGetAdj(gidx, True, "csr")))
    ns, nd = gidx.number_of_nodes(0), gidx.number_of_nodes(1)
    return (ns, nd, adj(heterograph_index._CAPI_DGLHeteroGetAdj(gidx, 0, False, "csr")),
            adj(heterograph_index._CAPI_DGLHeteroGetAdj(gidx, 0, True, "csr")))


def to_b200_graph(gidx, device):
    cache = gidx.__dict__.setdefault(_CACHE_ATTR, {})
    key = str(device)
    if key not in cache:
        ns, nd, in_csr, out_csr = reference_csrs(gidx)
        cache[key] = GraphIndex.from_csr(ns, nd, in_csr, out_csr, device)
    return cache[key]


def _wrap(fn, graph_pos, tensor_pos):
    def call(*args):
        args = list(args)
        tensors = [_to_torch(args[i]) for i in tensor_pos if i < len(args)]
        dev = next(t.device for t in tensors if t is not None)
        for i, t in zip(tensor_pos, tensors):
            args[i] = t
        args[graph_pos] = to_b200_graph(args[graph_pos], dev)
        return fn(*args)
    return call


def install():
    """Rebind dgl.kernel.* (and the alias the torch backend imported as `K`)."""
    import dgl.kernel as ref_kernel
    table = {
        "binary_op_reduce": _wrap(K.binary_op_reduce, 2, range(5, 11)),
        "backward_lhs_binary_op_reduce": _wrap(K.backward_lhs_binary_op_reduce, 2, range(5, 13)),
        "backward_rhs_binary_op_reduce": _wrap(K.backward_rhs_binary_op_reduce, 2, range(5, 13)),
        "copy_reduce": _wrap(K.copy_reduce, 1, range(3, 7)),
        "backward_copy_reduce": _wrap(K.backward_copy_reduce, 1, range(3, 9)),
        "infer_binary_feature_shape":
            lambda op, lhs, rhs: K.infer_binary_feature_shape(op, _to_torch(lhs), _to_torch(rhs)),
    }
    for name, fn in table.items():
        setattr(ref_kernel, name, fn)
    return sorted(table)

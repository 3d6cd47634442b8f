"""Reference-side binding: make DGL 0.4's own Python stack run on the B200 library.

`install()` registers the library's DGLPackedCFunc shims (csrc/packed.cu, DGLB200RegisterInto)
into the global function table of the reference's libdgl with override = 1
(include/dgl/runtime/c_runtime_api.h:319-346), under the names python/dgl/kernel.py binds with
`_init_api("dgl.kernel")` (kernel.py:438). From then on the UNMODIFIED reference code path --
`DGLGraph.update_all / apply_edges`, the scheduler IR, `backend/pytorch/tensor.py`'s autograd
`BinaryReduce / CopyReduce`, `dgl.nn.pytorch.{GraphConv, SAGEConv, GATConv, RelGraphConv}`,
`edge_softmax` -- calls the sm_100a kernels whenever its tensors are CUDA tensors, and the
reference's own CPU kernels otherwise (the shims forward those calls to the functions that were
registered before). No function of the reference's Python package is replaced.

Needs the reference's `dgl` package importable; nothing else in this package depends on this
module.

What crosses the boundary:
  * tensors: `dgl.ndarray.NDArray` containers built zero-copy from torch through DLPack
    (`zerocopy_to_dgl_ndarray`, backend/pytorch/tensor.py:283); the shims read the DLTensor header.
  * graph: the C++ graph object handle; the shims fetch its CSR arrays once per (graph, device)
    with the public PackedFunc `_CAPI_DGLGraphGetAdj(g, transpose, "csr")`
    (src/graph/graph_apis.cc:285-292; `_CAPI_DGLHeteroGetAdj` for unit graphs) and keep the
    device copy (DGLB200GraphCreateFromCSR), replacing ImmutableGraph::CopyTo
    (src/graph/immutable_graph.cc:548-559).

One accommodation, and only when the loaded libdgl was built WITHOUT CUDA (the build available
in this environment; a CUDA build needs none): `GraphIndex.get_immutable_gidx(gpu ctx)` /
`HeteroGraphIndex.get_unitgraph(gpu ctx)` would ask that library to copy the graph to a device it
has no API for, so they hand out the host index instead -- the kernels only read the CSR arrays;
and `dgl.ndarray.array(arr, ctx=gpu)` (the id maps of partial message passing) is built with torch
and wrapped through DLPack, because that library cannot allocate device memory either.
"""
import ctypes
import os
import sys

from ._lib import LIB, check_call

_state = {"installed": False, "cpu_only_libdgl": None}


def _find_libdgl(path=None):
    """dlopen handle of the reference library: the one `dgl` loaded, else `path`, else
    $DGL_LIBRARY_PATH/libdgl.so (the lookup of python/dgl/_ffi/libinfo.py)."""
    if "dgl._ffi.base" in sys.modules:
        return sys.modules["dgl._ffi.base"]._LIB._handle
    if path is None:
        path = os.path.join(os.environ.get("DGL_LIBRARY_PATH", ""), "libdgl.so")
    if not os.path.exists(path):
        raise FileNotFoundError("libdgl.so not found (pass its path or set DGL_LIBRARY_PATH)")
    # the same file dlopen'ed again by `import dgl` yields the same library instance
    return ctypes.CDLL(path, ctypes.RTLD_GLOBAL)._handle


def register(libdgl=None):
    """Registers the PackedFunc shims; `libdgl` = path, ctypes.CDLL, dlopen handle or None.
    Call BEFORE `import dgl` (python/dgl/kernel.py copies the global functions when it is
    imported), or use install(), which also rebinds an already imported dgl.kernel."""
    if isinstance(libdgl, ctypes.CDLL):
        handle = libdgl._handle
    elif isinstance(libdgl, int):
        handle = libdgl
    else:
        handle = _find_libdgl(libdgl)
    check_call(LIB.DGLB200RegisterInto(ctypes.c_void_p(handle)))
    return [LIB.DGLB200PackedFuncName(i).decode() for i in range(LIB.DGLB200PackedFuncCount())]


def _libdgl_has_cuda():
    from dgl._ffi.function import get_global_func
    return get_global_func("device_api.gpu", allow_missing=True) is not None


def _route_host_index():
    """CPU-only libdgl: GPU contexts get the host graph index (see the module docstring)."""
    from dgl import graph_index, heterograph_index
    from dgl import ndarray as nd

    def host_ctx(fn):
        def get(self, *args):              # get_immutable_gidx(ctx) / get_unitgraph(etype, ctx)
            ctx = args[-1]
            if ctx.device_type != 1:       # not kDLCPU
                args = args[:-1] + (nd.cpu(),)
            return fn(self, *args)
        get._b200_host_routed = True
        return get
    for cls, name in ((graph_index.GraphIndex, "get_immutable_gidx"),
                      (heterograph_index.HeteroGraphIndex, "get_unitgraph")):
        fn = getattr(cls, name, None)
        # (functools.wraps in utils.cached_member already sets __wrapped__: use an own marker)
        if fn is not None and not getattr(fn, "_b200_host_routed", False):
            setattr(cls, name, host_ctx(fn))


def _route_gpu_arrays():
    """CPU-only libdgl: `dgl.ndarray.array(arr, ctx=gpu)` (the id maps of send_and_recv / pull,
    python/dgl/runtime/spmv.py:180) would allocate through the library's absent CUDA device API.
    Build the array with torch and hand it over zero-copy through DLPack instead
    (DGLArrayFromDLPack needs no device API), the way every feature tensor already arrives."""
    import numpy as np
    import torch
    from torch.utils import dlpack
    from dgl import ndarray as nd
    if getattr(nd.array, "_b200_routed", False):
        return
    ref_array = nd.array

    def array(arr, ctx=nd.cpu(0)):
        if ctx.device_type == 1:
            return ref_array(arr, ctx)
        if isinstance(arr, nd.NDArray):
            arr = arr.asnumpy()
        t = torch.as_tensor(np.ascontiguousarray(arr)).to(torch.device("cuda", ctx.device_id))
        return nd.from_dlpack(dlpack.to_dlpack(t))
    array._b200_routed = True
    nd.array = array


def install(libdgl=None):
    """register() + rebinding of an already imported `dgl.kernel` (a re-run of the reference's
    own `_init_api("dgl.kernel")`) + the host-index routing for a CPU-only libdgl. Returns the
    registered global names."""
    names = register(libdgl)
    import dgl  # noqa: F401
    from dgl._ffi.function import _init_api
    _init_api("dgl.kernel")
    if _state["cpu_only_libdgl"] is None:
        _state["cpu_only_libdgl"] = not _libdgl_has_cuda()
    if _state["cpu_only_libdgl"]:
        _route_host_index()
        _route_gpu_arrays()
    _state["installed"] = True
    return names


def set_stream(stream=None):
    """The CUDA stream the shims launch on (thread-local; the reference's DGLSetStream role).
    None = torch's current stream."""
    import torch
    if stream is None:
        stream = torch.cuda.current_stream()
    check_call(LIB.DGLB200SetStream(ctypes.c_void_p(stream.cuda_stream)))

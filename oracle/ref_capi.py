"""TEST INFRASTRUCTURE ONLY -- drives the UNMODIFIED reference CPU library oracle/_ref/libdgl.so
(built by oracle/ref_build/Makefile from the sources under /root/reference) through the
reference's own C runtime API: DGLFuncGetGlobal / DGLFuncCall / DGLArrayAlloc
(include/dgl/runtime/c_runtime_api.h:255-260, 362, 388-445). No file of the reference's Python
package is needed, so this works on the GPU box where /root/reference does not exist.

Used by: bench.py's `cpu_baseline` leg and `--impl reference` arm (timing the reference's
OpenMP/minigun CPU kernels on the host cores) and tests that cross-check the oracle restatement
against the real reference. The product package never imports this module.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libdgl.so")

# DGLTypeCode (c_runtime_api.h:66-90); kDLInt=0, kDLFloat=2 from dlpack
K_INT, K_FLOAT, K_HANDLE, K_NULL, K_ARRAY, K_OBJECT, K_STR, K_NDARRAY = 0, 2, 3, 4, 7, 8, 11, 13


class _Value(ctypes.Union):
    _fields_ = [("v_int64", ctypes.c_int64), ("v_float64", ctypes.c_double),
                ("v_handle", ctypes.c_void_p), ("v_str", ctypes.c_char_p)]


def available():
    return os.path.exists(REF_SO)


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(REF_SO, ctypes.RTLD_GLOBAL)
        _lib.DGLGetLastError.restype = ctypes.c_char_p
    return _lib


def set_num_threads(n):
    """omp_set_num_threads on the libgomp the reference library is linked against, so the CPU arm
    uses the same number of host threads whatever OMP_NUM_THREADS the launcher exported
    (torch.distributed.run sets it to 1). Returns the number now in effect."""
    lib()
    gomp = ctypes.CDLL("libgomp.so.1", ctypes.RTLD_GLOBAL)
    gomp.omp_set_num_threads(int(n))
    return int(gomp.omp_get_max_threads())


def _check(rc):
    if rc != 0:
        raise RuntimeError(lib().DGLGetLastError().decode("utf-8", "replace"))


class RefArray(object):
    """A reference NDArray on the CPU (DGLArrayAlloc), filled from / read into numpy."""

    def __init__(self, arr=None, shape=None, dtype=None):
        if arr is not None:
            arr = np.ascontiguousarray(arr)
            shape, dtype = arr.shape, arr.dtype
        dtype = np.dtype(dtype)
        self.shape, self.dtype = tuple(shape), dtype
        code = K_FLOAT if dtype.kind == "f" else K_INT
        shp = (ctypes.c_int64 * max(len(shape), 1))(*shape)
        self.handle = ctypes.c_void_p()
        _check(lib().DGLArrayAlloc(shp, len(shape), code, dtype.itemsize * 8, 1, 1, 0,
                                   ctypes.byref(self.handle)))
        if arr is not None and arr.nbytes:
            _check(lib().DGLArrayCopyFromBytes(self.handle, arr.ctypes.data_as(ctypes.c_void_p),
                                               ctypes.c_size_t(arr.nbytes)))

    @classmethod
    def from_dlpack(cls, tensor):
        """A reference NDArray sharing a torch tensor's memory (CPU or CUDA), the way the
        reference's torch backend hands tensors to the kernels (`zerocopy_to_dgl_ndarray` =
        nd.from_dlpack(to_dlpack(x)), python/dgl/backend/pytorch/tensor.py:283;
        python/dgl/_ffi/_ctypes/ndarray.py `_from_dlpack` -> DGLArrayFromDLPack)."""
        from torch.utils import dlpack
        tensor = tensor.contiguous()
        cap = dlpack.to_dlpack(tensor)
        api = ctypes.pythonapi
        api.PyCapsule_GetPointer.restype = ctypes.c_void_p
        api.PyCapsule_GetPointer.argtypes = [ctypes.py_object, ctypes.c_char_p]
        api.PyCapsule_SetName.argtypes = [ctypes.py_object, ctypes.c_char_p]
        api.PyCapsule_SetDestructor.argtypes = [ctypes.py_object, ctypes.c_void_p]
        ptr = api.PyCapsule_GetPointer(cap, b"dltensor")
        self = cls.__new__(cls)
        self.shape = tuple(tensor.shape)
        self.dtype = None
        self.handle = ctypes.c_void_p()
        _check(lib().DGLArrayFromDLPack(ctypes.c_void_p(ptr), ctypes.byref(self.handle)))
        # the NDArray now owns the managed tensor (its deleter releases torch's reference)
        api.PyCapsule_SetName(cap, _USED)
        api.PyCapsule_SetDestructor(cap, None)
        self._keep = (tensor, cap)
        return self

    def numpy(self):
        out = np.empty(self.shape, self.dtype)
        if out.nbytes:
            _check(lib().DGLArrayCopyToBytes(self.handle, out.ctypes.data_as(ctypes.c_void_p),
                                             ctypes.c_size_t(out.nbytes)))
        return out

    def __del__(self):
        try:
            if self.handle:
                lib().DGLArrayFree(self.handle)
        except Exception:
            pass


_USED = ctypes.create_string_buffer(b"used_dltensor")   # capsule names are borrowed pointers
_funcs = {}


def forget_functions():
    """Drops the cached global-function handles (DGLFuncGetGlobal returns a COPY of the registered
    function, so a handle looked up before an override keeps calling the old body)."""
    _funcs.clear()


def call(name, *args):
    """Call a registered PackedFunc by global name. Returns (value, type_code)."""
    if name not in _funcs:
        h = ctypes.c_void_p()
        _check(lib().DGLFuncGetGlobal(name.encode(), ctypes.byref(h)))
        if not h:
            raise RuntimeError("reference global function %s not found" % name)
        _funcs[name] = h
    n = len(args)
    vals = (_Value * max(n, 1))()
    codes = (ctypes.c_int * max(n, 1))()
    keep = []
    for i, a in enumerate(args):
        if isinstance(a, RefArray):
            vals[i].v_handle, codes[i] = a.handle, K_NDARRAY
        elif isinstance(a, RefObject):
            vals[i].v_handle, codes[i] = a.handle, K_OBJECT
        elif isinstance(a, (bool, int, np.integer)):
            vals[i].v_int64, codes[i] = int(a), K_INT
        elif isinstance(a, float):
            vals[i].v_float64, codes[i] = a, K_FLOAT
        elif isinstance(a, str):
            b = a.encode()
            keep.append(b)
            vals[i].v_str, codes[i] = b, K_STR
        else:
            raise TypeError("unsupported PackedFunc argument %r" % (a,))
    ret, rcode = _Value(), ctypes.c_int()
    _check(lib().DGLFuncCall(_funcs[name], vals, codes, n, ctypes.byref(ret), ctypes.byref(rcode)))
    return ret, rcode.value


def returned_int64_array(ret, code):
    """A 1-D int64 CPU NDArray returned by a PackedFunc -> python list (and release it)."""
    assert code == K_NDARRAY, code

    class _DL(ctypes.Structure):
        _fields_ = [("data", ctypes.c_void_p), ("device_type", ctypes.c_int32),
                    ("device_id", ctypes.c_int32), ("ndim", ctypes.c_int32),
                    ("code", ctypes.c_uint8), ("bits", ctypes.c_uint8), ("lanes", ctypes.c_uint16),
                    ("shape", ctypes.POINTER(ctypes.c_int64)),
                    ("strides", ctypes.POINTER(ctypes.c_int64)), ("byte_offset", ctypes.c_uint64)]
    t = ctypes.cast(ret.v_handle, ctypes.POINTER(_DL)).contents
    n = t.shape[0] if t.ndim else 0
    vals = list((ctypes.c_int64 * n).from_address(t.data + t.byte_offset)) if n else []
    lib().DGLArrayFree(ctypes.c_void_p(ret.v_handle))
    return vals


class RefObject(object):
    def __init__(self, handle):
        self.handle = ctypes.c_void_p(handle)


_NULL = None


def null():
    global _NULL
    if _NULL is None:
        _NULL = RefArray(np.zeros((0,), np.int64))  # python/dgl/ndarray.py:97-106
    return _NULL


class RefGraph(object):
    """ImmutableGraph of the reference, created from COO exactly like
    graph_index.from_coo(...) -> _CAPI_DGLGraphCreate(src, dst, num_nodes, readonly=True)
    (src/graph/graph_apis.cc:29-40); int64 ids on the CPU."""

    def __init__(self, num_nodes, src, dst):
        self.num_nodes = int(num_nodes)
        self.num_edges = len(src)
        s, d = RefArray(np.asarray(src, np.int64)), RefArray(np.asarray(dst, np.int64))
        ret, code = call("graph_index._CAPI_DGLGraphCreate", s, d, self.num_nodes, True)
        assert code == K_OBJECT, code
        self.obj = RefObject(ret.v_handle)

    def materialize(self):
        """Force both CSRs (what get_immutable_gidx does before the first kernel call)."""
        for transpose in (False, True):
            call("graph_index._CAPI_DGLGraphGetAdj", self.obj, transpose, "csr")


def copy_reduce(reducer, g, target, x, out):
    """kernel._CAPI_DGLKernelCopyReduce(reducer, graph, target, in, out, in_map, out_map)."""
    call("kernel._CAPI_DGLKernelCopyReduce", reducer, g.obj, int(target), x, out, null(), null())


def backward_copy_reduce(reducer, g, target, x, out, grad_out, grad_in):
    call("kernel._CAPI_DGLKernelBackwardCopyReduce", reducer, g.obj, int(target), x, out,
         grad_out, grad_in, null(), null())


def binary_op_reduce(reducer, op, g, lt, rt, lhs, rhs, out):
    call("kernel._CAPI_DGLKernelBinaryOpReduce", reducer, op, g.obj, int(lt), int(rt), lhs, rhs,
         out, null(), null(), null())

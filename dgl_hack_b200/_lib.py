"""ctypes binding of libdglb200.so (include/dgl_b200.h).

This is the reference-side FFI stub a maintainer would add next to python/dgl/_ffi: it plays the
role of python/dgl/_ffi/base.py (_load_lib, check_call, DGLError) for the B200 kernel library.
The library is built in-tree by `make -C dgl_hack_b200/csrc` (see __graft_entry__.build()).
There is NO fallback: if the shared object is missing, importing this module raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdglb200.so")


class DGLError(RuntimeError):
    """Error raised by the kernel library (python/dgl/_ffi/base.py:50 DGLError)."""


class Context(ctypes.Structure):
    _fields_ = [("device_type", ctypes.c_int32), ("device_id", ctypes.c_int32)]


class DataType(ctypes.Structure):
    _fields_ = [("code", ctypes.c_uint8), ("bits", ctypes.c_uint8), ("lanes", ctypes.c_uint16)]


class Tensor(ctypes.Structure):
    """Layout-identical to DLTensor (third_party/dlpack/include/dlpack/dlpack.h)."""
    _fields_ = [("data", ctypes.c_void_p), ("ctx", Context), ("ndim", ctypes.c_int32),
                ("dtype", DataType), ("shape", ctypes.POINTER(ctypes.c_int64)),
                ("strides", ctypes.POINTER(ctypes.c_int64)), ("byte_offset", ctypes.c_uint64)]


# tensors cross the boundary as plain addresses of DLTensor-layout structs (see ndarray.NDArray)
TensorP = ctypes.c_void_p
GraphHandle = ctypes.c_void_p

# every symbol include/dgl_b200.h declares: name -> (restype, argtypes)
_c = ctypes
SYMBOLS = {
    "DGLB200GetLastError": (_c.c_char_p, []),
    "DGLB200Version": (_c.c_int, []),
    "DGLB200LastKernel": (_c.c_char_p, []),
    "DGLB200LaunchCount": (_c.c_int64, []),
    "DGLB200GraphCreateFromCOO": (_c.c_int, [_c.c_int64, _c.c_int64, _c.c_int64, _c.c_void_p,
                                             _c.c_void_p, _c.c_int, _c.c_void_p,
                                             _c.POINTER(GraphHandle)]),
    "DGLB200GraphCreateFromCSR": (_c.c_int, [_c.c_int64, _c.c_int64, _c.c_int64, _c.c_int,
                                             _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p,
                                             _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_void_p,
                                             _c.POINTER(GraphHandle)]),
    "DGLB200GraphFree": (_c.c_int, [GraphHandle]),
    "DGLB200GraphIndexBytes": (_c.c_int, [GraphHandle, _c.POINTER(_c.c_int64)]),
    "DGLB200GraphInfo": (_c.c_int, [GraphHandle, _c.POINTER(_c.c_int64)]),
    "DGLB200GraphGetCSR": (_c.c_int, [GraphHandle, _c.c_int, _c.POINTER(_c.c_void_p),
                                      _c.POINTER(_c.c_void_p), _c.POINTER(_c.c_void_p)]),
    "DGLB200GraphCopyCSR": (_c.c_int, [GraphHandle, _c.c_int, _c.c_void_p, _c.c_void_p,
                                       _c.c_void_p, _c.c_void_p]),
    "DGLB200GraphDegrees": (_c.c_int, [GraphHandle, _c.c_int, _c.c_void_p, _c.c_void_p]),
    "DGLB200COOToCSR": (_c.c_int, [_c.c_int64, _c.c_int64, _c.c_void_p, _c.c_void_p,
                                   _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p]),
    "DGLB200CSRTranspose": (_c.c_int, [_c.c_int64, _c.c_int64, _c.c_int64, _c.c_void_p,
                                       _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p,
                                       _c.c_void_p, _c.c_void_p]),
    "DGLB200CSRToCOORows": (_c.c_int, [_c.c_int64, _c.c_int64, _c.c_void_p, _c.c_void_p,
                                       _c.c_void_p]),
    "DGLB200AsInt32": (_c.c_int, [_c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_void_p]),
    "DGLB200KernelInferBinaryFeatureShape": (_c.c_int, [_c.c_char_p, TensorP, TensorP,
                                                        _c.POINTER(_c.c_int32),
                                                        _c.POINTER(_c.c_int64)]),
    "DGLB200KernelBinaryOpReduce": (_c.c_int, [_c.c_char_p, _c.c_char_p, GraphHandle, _c.c_int,
                                               _c.c_int, TensorP, TensorP, TensorP, TensorP,
                                               TensorP, TensorP, _c.c_void_p]),
    "DGLB200KernelBackwardLhsBinaryOpReduce": (_c.c_int, [_c.c_char_p, _c.c_char_p, GraphHandle,
                                                          _c.c_int, _c.c_int, TensorP, TensorP,
                                                          TensorP, TensorP, TensorP, TensorP,
                                                          TensorP, TensorP, _c.c_void_p]),
    "DGLB200KernelBackwardRhsBinaryOpReduce": (_c.c_int, [_c.c_char_p, _c.c_char_p, GraphHandle,
                                                          _c.c_int, _c.c_int, TensorP, TensorP,
                                                          TensorP, TensorP, TensorP, TensorP,
                                                          TensorP, TensorP, _c.c_void_p]),
    "DGLB200KernelCopyReduce": (_c.c_int, [_c.c_char_p, GraphHandle, _c.c_int, TensorP, TensorP,
                                           TensorP, TensorP, _c.c_void_p]),
    "DGLB200FusedGatKernel": (_c.c_int, [GraphHandle, TensorP, TensorP, TensorP, TensorP, TensorP,
                                         TensorP, _c.c_float, _c.c_void_p]),
    "DGLB200KernelBackwardFusedGat": (_c.c_int, [GraphHandle] + [TensorP] * 10 +
                                      [_c.c_float, _c.c_void_p]),
    "DGLB200GraphSetEdgeTypes": (_c.c_int, [GraphHandle, _c.c_void_p, _c.c_int, _c.c_int64,
                                            _c.c_void_p]),
    "DGLB200GraphCopyCsrSortedByEdgeType": (_c.c_int, [GraphHandle, _c.c_int, _c.c_void_p,
                                                       _c.c_void_p, _c.c_void_p, _c.c_void_p,
                                                       _c.c_void_p]),
    "DGLB200RgcnLayer0": (_c.c_int, [GraphHandle, TensorP, TensorP, TensorP, _c.c_void_p]),
    "DGLB200RgcnLayer0Backward": (_c.c_int, [GraphHandle, TensorP, TensorP, TensorP, _c.c_int,
                                             _c.c_void_p]),
    "DGLB200RgcnLayer1": (_c.c_int, [GraphHandle, TensorP, TensorP, TensorP, TensorP,
                                     _c.c_void_p]),
    "DGLB200RgcnLayer1Backward": (_c.c_int, [GraphHandle] + [TensorP] * 6 + [_c.c_void_p]),
    "DGLB200PeerAlloc": (_c.c_int, [_c.c_int64, _c.c_int, _c.POINTER(_c.c_void_p), _c.c_void_p]),
    "DGLB200PeerOpen": (_c.c_int, [_c.c_void_p, _c.c_int, _c.POINTER(_c.c_void_p)]),
    "DGLB200PeerClose": (_c.c_int, [_c.c_void_p]),
    "DGLB200PeerFree": (_c.c_int, [_c.c_void_p]),
    "DGLB200HaloPush": (_c.c_int, [_c.c_void_p, _c.c_int64, _c.c_int, _c.c_void_p, _c.c_void_p,
                                   _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_void_p]),
    "DGLB200HaloScatter": (_c.c_int, [_c.c_int, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p,
                                      _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_int, _c.c_int,
                                      _c.c_void_p]),
    "DGLB200HaloPull": (_c.c_int, [_c.c_int, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int64,
                                   _c.c_int64, _c.c_void_p, _c.c_int, _c.c_int, _c.c_void_p]),
    "DGLB200MetisPartition": (_c.c_int, [_c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_void_p,
                                         _c.c_int, _c.c_uint64, _c.c_void_p,
                                         _c.POINTER(_c.c_int64)]),
    "DGLB200SymmetrizeCOO": (_c.c_int, [_c.c_int64, _c.c_int64, _c.c_void_p, _c.c_void_p,
                                        _c.c_void_p, _c.c_void_p, _c.POINTER(_c.c_int64)]),
    "DGLB200GatAttnProj": (_c.c_int, [TensorP] * 5 + [_c.c_void_p]),
    "DGLB200GatAttnProjBackward": (_c.c_int, [TensorP] * 8 + [_c.c_void_p]),
    "DGLB200GatAttnProjBackwardNoAlloc": (_c.c_int, [TensorP] * 8 + [_c.c_void_p]),
    "DGLB200SampleCounts": (_c.c_int, [GraphHandle, _c.c_void_p, _c.c_int64, _c.c_int, _c.c_int,
                                       _c.c_void_p, _c.c_void_p]),
    "DGLB200SampleNeighbors": (_c.c_int, [GraphHandle, _c.c_void_p, _c.c_int64, _c.c_int, _c.c_int,
                                          _c.c_uint64, _c.c_void_p, _c.c_void_p, _c.c_void_p,
                                          _c.c_void_p, _c.c_void_p, _c.c_void_p]),
    "DGLB200CompactBlock": (_c.c_int, [_c.c_void_p, _c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_int64,
                                       _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p,
                                       _c.c_void_p]),
    "DGLB200KernelScaledCopySum": (_c.c_int, [GraphHandle, _c.c_int, TensorP, TensorP, TensorP,
                                              TensorP, _c.c_void_p]),
    "DGLB200KernelScaledCopySumAdd": (_c.c_int, [GraphHandle, _c.c_int, TensorP, TensorP, TensorP,
                                                 TensorP, TensorP, _c.c_void_p]),
    "DGLB200GraphCreateFromTypedCSR": (_c.c_int, [_c.c_int64, _c.c_int64, _c.c_int, _c.c_void_p,
                                                  _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int64,
                                                  _c.c_int, _c.c_void_p, _c.POINTER(GraphHandle)]),
    "DGLB200GraphCreateRelationSegments": (_c.c_int, [GraphHandle, _c.c_void_p,
                                                      _c.POINTER(GraphHandle)]),
    "DGLB200GraphCopyRelationSegments": (_c.c_int, [GraphHandle, _c.c_void_p, _c.c_void_p,
                                                    _c.c_void_p]),
    "DGLB200KernelCopySumBF16": (_c.c_int, [GraphHandle, _c.c_int, TensorP, TensorP, TensorP,
                                            _c.c_int, TensorP, _c.c_void_p]),
    "DGLB200DenseTF32x3": (_c.c_int, [TensorP, TensorP, _c.c_int, TensorP, _c.c_int, TensorP,
                                      _c.c_void_p]),
    "DGLB200DenseTF32x3WeightGrad": (_c.c_int, [TensorP, TensorP, TensorP, _c.c_void_p]),
    "DGLB200GroupedDenseTF32x3": (_c.c_int, [TensorP, TensorP, _c.c_int, _c.c_void_p, _c.c_int,
                                             TensorP, _c.c_void_p]),
    "DGLB200RegisterInto": (_c.c_int, [_c.c_void_p]),
    "DGLB200SetStream": (_c.c_int, [_c.c_void_p]),
    "DGLB200PackedFuncCount": (_c.c_int, []),
    "DGLB200PackedFuncName": (_c.c_char_p, [_c.c_int]),
    "DGLB200PackedFuncGet": (_c.c_void_p, [_c.c_int]),
    "DGLB200PackedCacheClear": (_c.c_int, []),
    "DGLB200PackedStats": (_c.c_int, [_c.POINTER(_c.c_int64)]),
    "DGLB200KernelPeerCopySum": (_c.c_int, [GraphHandle, _c.c_int, _c.c_int, _c.c_void_p, TensorP,
                                            TensorP, TensorP, _c.c_void_p]),
    "DGLB200EdgeSoftmaxForward": (_c.c_int, [GraphHandle, TensorP, TensorP, _c.c_void_p]),
    "DGLB200EdgeSoftmaxBackward": (_c.c_int, [GraphHandle, TensorP, TensorP, TensorP,
                                              _c.c_void_p]),
    "DGLB200KernelCopyReduceMean": (_c.c_int, [GraphHandle, _c.c_int, TensorP, TensorP,
                                               _c.c_void_p]),
    "DGLB200KernelBackwardCopyReduce": (_c.c_int, [_c.c_char_p, GraphHandle, _c.c_int, TensorP,
                                                   TensorP, TensorP, TensorP, TensorP, TensorP,
                                                   _c.c_void_p]),
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `make -C dgl_hack_b200/csrc` "
            "(or __graft_entry__.build()). There is no CPU / PyTorch fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    return lib


LIB = _load()


def check_call(ret):
    """python/dgl/_ffi/base.py:64 check_call."""
    if ret != 0:
        raise DGLError(LIB.DGLB200GetLastError().decode("utf-8", "replace"))


def last_kernel():
    return LIB.DGLB200LastKernel().decode()


def launch_count():
    return int(LIB.DGLB200LaunchCount())

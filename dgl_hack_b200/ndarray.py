"""Tensor views handed to the C ABI.

Mirrors python/dgl/ndarray.py + zerocopy_to_dgl_ndarray (python/dgl/backend/pytorch/tensor.py:
283-284): a torch tensor becomes a DLTensor-compatible view without a copy (a copy only if it is
not contiguous, exactly as the reference's `.contiguous()`), and `null()` is the empty int64
array the reference uses for "no mapping" (python/dgl/ndarray.py:97-106).
"""
import ctypes

import torch

from . import _lib

_DTYPES = {
    torch.float32: (2, 32), torch.float64: (2, 64), torch.float16: (2, 16),
    torch.int32: (0, 32), torch.int64: (0, 64), torch.uint8: (1, 8), torch.int8: (0, 8),
    torch.int16: (0, 16), torch.bool: (1, 8),
}


class NDArray(object):
    """A DLTensor-compatible view of a torch tensor (keeps the tensor alive)."""

    def __init__(self, tensor):
        tensor = tensor.contiguous()
        if tensor.dtype == torch.bfloat16:
            raise _lib.DGLError("bfloat16 tensors are not supported by the fp32 kernels")
        self.tensor = tensor
        self._shape = (ctypes.c_int64 * max(tensor.dim(), 1))(*tensor.shape)
        code, bits = _DTYPES[tensor.dtype]
        dev = tensor.device
        self.handle = _lib.Tensor(
            data=ctypes.c_void_p(tensor.data_ptr()),
            ctx=_lib.Context(2 if dev.type == "cuda" else 1,
                             (dev.index or 0) if dev.type == "cuda" else 0),
            ndim=tensor.dim(), dtype=_lib.DataType(code, bits, 1),
            shape=ctypes.cast(self._shape, ctypes.POINTER(ctypes.c_int64)),
            strides=None, byte_offset=0)

    @property
    def shape(self):
        return tuple(self.tensor.shape)

    @property
    def ptr(self):
        return ctypes.byref(self.handle)


def zerocopy_to_dgl_ndarray(tensor):
    """tensor.py:283 -- zero-copy view of a (contiguous) torch tensor."""
    return tensor if isinstance(tensor, NDArray) else NDArray(tensor)


_NULL = None


def null():
    """python/dgl/ndarray.py:97 null(): 1-D empty int64 array on the CPU."""
    global _NULL
    if _NULL is None:
        _NULL = NDArray(torch.empty((0,), dtype=torch.int64))
    return _NULL

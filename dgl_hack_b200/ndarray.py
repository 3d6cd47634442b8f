"""Tensor views handed to the C ABI.

Mirrors python/dgl/ndarray.py + zerocopy_to_dgl_ndarray (python/dgl/backend/pytorch/tensor.py:
283-284): a torch tensor becomes a DLTensor-compatible view without a copy (a copy only if it is
not contiguous, exactly as the reference's `.contiguous()`), and `null()` is the empty int64
array the reference uses for "no mapping" (python/dgl/ndarray.py:97-106).
"""
import ctypes
import struct

import torch

from . import _lib

_DTYPES = {
    torch.float32: (2, 32), torch.float64: (2, 64), torch.float16: (2, 16),
    torch.int32: (0, 32), torch.int64: (0, 64), torch.uint8: (1, 8), torch.int8: (0, 8),
    torch.int16: (0, 16), torch.bool: (1, 8),
    torch.bfloat16: (4, 16),   # kDLBfloat: accepted by the explicit bf16 entry points only
}


_ARR = {}  # ndim -> ctypes array type of the packed view
_SHP = {}  # ndim -> struct.Struct of the shape array
_HDR = struct.Struct("<QiiiBBHQQQ")  # DLTensor: data, ctx{type,id}, ndim, dtype{code,bits,lanes},
#                                       shape*, strides*, byte_offset  (48 bytes)


class NDArray(object):
    """A DLTensor-compatible view of a torch tensor (keeps the tensor alive). The 48-byte header
    and the shape array are packed into one buffer with struct (an order of magnitude cheaper
    than building ctypes Structures on every kernel call)."""
    __slots__ = ("tensor", "_buf", "_cbuf", "ptr")

    def __init__(self, tensor):
        tensor = tensor.contiguous()
        self.tensor = tensor
        shape = tensor.shape
        nd = len(shape)
        buf = bytearray(48 + 8 * max(nd, 1))
        self._buf = buf
        arr_t = _ARR.get(nd)
        if arr_t is None:
            arr_t = _ARR[nd] = ctypes.c_char * len(buf)
        self._cbuf = arr_t.from_buffer(buf)
        addr = ctypes.addressof(self._cbuf)
        code, bits = _DTYPES[tensor.dtype]
        dev = tensor.device
        cuda = dev.type == "cuda"
        _HDR.pack_into(buf, 0, tensor.data_ptr(), 2 if cuda else 1,
                       (dev.index or 0) if cuda else 0, nd, code, bits, 1, addr + 48, 0, 0)
        if nd:
            shp = _SHP.get(nd)
            if shp is None:
                shp = _SHP[nd] = struct.Struct("<%dq" % nd)
            shp.pack_into(buf, 48, *shape)
        self.ptr = addr  # passed as void* (DGLB200Tensor*)

    @property
    def shape(self):
        return tuple(self.tensor.shape)

    @property
    def handle(self):
        """The view as a ctypes DLTensor struct (for callers that want field access)."""
        return _lib.Tensor.from_buffer(self._buf)


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

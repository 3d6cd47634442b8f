"""B200-native (sm_100a) builtin message-passing kernels behind DGL 0.4's kernel interface.

Only what the hot path needs: the C-ABI library (csrc/, include/dgl_b200.h), wrappers with the
reference's python/dgl/kernel.py names (kernel), the torch autograd dispatch with the reference's
backend names (backend) and the device-resident graph index (graph_index). Importing this package
loads libdglb200.so and fails loudly if it has not been built -- there is no fallback path.
"""
from . import _lib, backend, kernel  # noqa: F401
from ._lib import DGLError  # noqa: F401
from .backend import binary_reduce, copy_reduce  # noqa: F401
from .graph_index import GraphIndex  # noqa: F401

__all__ = ["GraphIndex", "binary_reduce", "copy_reduce", "kernel", "backend", "DGLError"]

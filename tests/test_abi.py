"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/dgl_b200.h declares, its pure-host functions agree with the oracle, and compute entry
points refuse CPU tensors loudly (there is no CPU fallback). No kernel is launched here."""
import ctypes
import itertools
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from dgl_hack_b200 import _lib
    header = open(os.path.join(ROOT, "include", "dgl_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(DGLB200[A-Za-z0-9]+)\s*\(", header))
    assert len(declared) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), "library does not export %s" % name
    assert declared == set(_lib.SYMBOLS), "ctypes binding and header disagree"
    assert _lib.LIB.DGLB200Version() == 1


def test_infer_binary_feature_shape_matches_oracle():
    from oracle import oracle as O
    from dgl_hack_b200 import kernel as K
    shapes = [(3,), (1,), (2, 3), (2, 1), (1, 3), (4, 2, 3), (4, 1, 3), (1, 1, 1), (5, 1, 4, 1)]
    n = 0
    for op in ("add", "sub", "mul", "div", "dot"):
        for ls, rs in itertools.product(shapes, shapes):
            lhs, rhs = torch.zeros((6,) + ls), torch.zeros((9,) + rs)
            if op == "dot" and ls[-1] != rs[-1]:
                continue
            try:
                want = tuple(O.infer_binary_feature_shape(op, lhs.shape, rhs.shape))
            except ValueError:
                with pytest.raises(Exception, match="Invalid broadcasting"):
                    K.infer_binary_feature_shape(op, lhs, rhs)
                continue
            assert K.infer_binary_feature_shape(op, lhs, rhs) == want
            n += 1
    assert n > 100


def test_no_cpu_path_and_error_convention():
    """Errors come back as return code + last-error string (DGLError), like the reference's
    check_call; a null graph / CPU tensors never reach a kernel."""
    from dgl_hack_b200 import DGLError, _lib
    from dgl_hack_b200.ndarray import NDArray, null
    x = NDArray(torch.zeros(4, 3))
    out = NDArray(torch.zeros(4, 3))
    rc = _lib.LIB.DGLB200KernelCopyReduce(b"sum", None, 0, x.ptr, out.ptr, null().ptr,
                                          null().ptr, None)
    assert rc == -1
    assert b"NULL" in _lib.LIB.DGLB200GetLastError()
    with pytest.raises(DGLError):
        _lib.check_call(rc)
    if not torch.cuda.is_available():
        from dgl_hack_b200 import GraphIndex
        with pytest.raises(DGLError):
            GraphIndex.from_coo(3, 3, [0, 1], [1, 2], device="cpu")


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    import importlib
    from dgl_hack_b200 import _lib
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU / PyTorch fallback"):
        _lib._load()
    importlib.reload(_lib)


def test_product_does_not_import_oracle():
    """The product package must never route through oracle/ (test infrastructure only)."""
    pkg = os.path.join(ROOT, "dgl_hack_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "libdgl_oracle" not in text and "oracle/_ref" not in text, f

"""The PackedFunc face of the library (csrc/packed.cu) without a GPU: the shims register into
the reference's own libdgl (oracle/_ref/libdgl.so, built from /root/reference by
oracle/ref_build/Makefile) under the 12 `kernel._CAPI_*` names the reference's Python binds, the
host-only one (InferBinaryFeatureShape) answers through the reference's DGLFuncCall, and calls
with CPU tensors are forwarded to the reference's CPU kernels (same results as before the
override). The reference's Python stack driven through the shims: ref_python_driver.py."""
import ctypes
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_capi as R  # noqa: E402

REF_NAMES = [   # src/kernel/binary_reduce.cc:281, 357, 502, 600, 649, 697, 380, 529, 412-450
    "kernel._CAPI_DGLKernelInferBinaryFeatureShape", "kernel._CAPI_DGLKernelBinaryOpReduce",
    "kernel._CAPI_DGLKernelBackwardLhsBinaryOpReduce",
    "kernel._CAPI_DGLKernelBackwardRhsBinaryOpReduce", "kernel._CAPI_DGLKernelCopyReduce",
    "kernel._CAPI_DGLKernelBackwardCopyReduce", "kernel._CAPI_DGLFusedGatKernel",
    "kernel._CAPI_DGLKernelBackwardFusedGat", "kernel._CAPI_DGLRgcnLayer0",
    "kernel._CAPI_DGLRgcnLayer0Backward", "kernel._CAPI_DGLRgcnLayer1",
    "kernel._CAPI_DGLRgcnLayer1Backward"]

needs_ref = pytest.mark.skipif(not R.available(), reason="oracle/_ref/libdgl.so not built")


def test_shim_table_names_the_reference_globals():
    from dgl_hack_b200._lib import LIB
    n = LIB.DGLB200PackedFuncCount()
    names = [LIB.DGLB200PackedFuncName(i).decode() for i in range(n)]
    assert names == REF_NAMES
    assert all(LIB.DGLB200PackedFuncGet(i) for i in range(n))
    assert LIB.DGLB200PackedFuncGet(n) is None and LIB.DGLB200PackedFuncName(-1) is None


@needs_ref
def test_register_into_reference_and_call_through_dglfunccall():
    from dgl_hack_b200._lib import LIB, check_call
    # before the override: the reference's own CPU kernel (one thread: its `omp atomic` sums are
    # order-dependent in the last ulp otherwise)
    R.set_num_threads(1)
    rs = np.random.RandomState(0)
    n, e = 50, 400
    src, dst = rs.randint(0, n, e), rs.randint(0, n, e)
    x = rs.uniform(-1, 1, (n, 5)).astype(np.float32)
    g = R.RefGraph(n, src, dst)
    X = R.RefArray(x)
    before = R.RefArray(shape=x.shape, dtype=np.float32)
    R.copy_reduce("sum", g, 0, X, before)
    stats0 = (ctypes.c_int64 * 3)()
    LIB.DGLB200PackedStats(stats0)
    check_call(LIB.DGLB200RegisterInto(ctypes.c_void_p(R.lib()._handle)))
    check_call(LIB.DGLB200RegisterInto(ctypes.c_void_p(R.lib()._handle)))   # idempotent
    R.forget_functions()
    # every reference name now resolves (to the shim)
    for name in REF_NAMES:
        h = ctypes.c_void_p()
        assert R.lib().DGLFuncGetGlobal(name.encode(), ctypes.byref(h)) == 0 and h
    # host-only function answered by the shim through the reference's calling convention
    a = R.RefArray(np.zeros((4, 2, 3), np.float32))
    b = R.RefArray(np.zeros((4, 1, 3), np.float32))
    ret, code = R.call("kernel._CAPI_DGLKernelInferBinaryFeatureShape", "mul", a, b)
    assert R.returned_int64_array(ret, code) == [2, 3]
    # CPU tensors: forwarded to the function registered before = identical result
    after = R.RefArray(shape=x.shape, dtype=np.float32)
    R.copy_reduce("sum", g, 0, X, after)
    assert np.array_equal(after.numpy(), before.numpy())
    stats = (ctypes.c_int64 * 3)()
    LIB.DGLB200PackedStats(stats)
    assert stats[0] - stats0[0] >= 2 and stats[1] - stats0[1] >= 1
    # errors come back the reference's way: DGLFuncCall returns -1 + DGLGetLastError
    with pytest.raises(RuntimeError, match="expected 7 arguments"):
        R.call("kernel._CAPI_DGLKernelCopyReduce", "sum", g.obj, 0)
    with pytest.raises(RuntimeError, match="expected str|must be a string"):
        R.call("kernel._CAPI_DGLKernelCopyReduce", 3, g.obj, 0, X, after, R.null(), R.null())


@needs_ref
@pytest.mark.parametrize("register_first", [1, 0])
def test_reference_python_stack_through_the_shims_cpu(register_first):
    """The reference's own DGLGraph / nn modules with the shims registered (before `import dgl`,
    or afterwards through install()): every kernel call goes through a shim and, with CPU
    tensors, on to the reference's CPU kernels."""
    pkg = os.path.join(ROOT, "oracle", "_ref", "python")
    if not os.path.isdir(os.path.join(pkg, "dgl")):
        pytest.skip("reference python package not installed under oracle/_ref/python")
    env = dict(os.environ, OMP_NUM_THREADS="1", DGLBACKEND="pytorch",
               DGL_LIBRARY_PATH=os.path.join(ROOT, "oracle", "_ref"), PYTHONPATH=pkg)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_python_driver.py"),
                          "--device", "cpu", "--register-first", str(register_first)],
                         env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    rec = json.loads(out.stdout.strip().splitlines()[-1])
    assert rec["registered"] == REF_NAMES
    assert all(v == 0.0 for v in rec["cases"].values()), rec["cases"]
    assert rec["shim_calls"] > 50 and rec["forwarded_to_reference_cpu"] > 50
    assert rec["b200_launches"] == 0

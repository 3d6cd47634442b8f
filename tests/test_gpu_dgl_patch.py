"""The drop-in, proven on the GPU: with the PackedFunc shims registered into the reference's
libdgl (DGLB200RegisterInto, csrc/packed.cu),

  * the reference's calling convention itself -- DGLFuncGetGlobal + DGLFuncCall on
    `kernel._CAPI_DGLKernel*` with CUDA NDArrays made by DGLArrayFromDLPack and the reference's
    own graph object (src/kernel/binary_reduce.cc:357-378, 502-527, 600-626, 649-665, 697-716)
    -- runs the sm_100a kernels and matches the oracle;
  * the reference's own Python stack (oracle/_ref/python/dgl, installed from /root/reference by
    oracle/ref_build/Makefile): DGLGraph.update_all / apply_edges / send_and_recv / pull,
    edge_softmax, dgl.nn.pytorch.GraphConv / SAGEConv / GATConv / RelGraphConv with CUDA tensors
    matches the same calls with CPU tensors (reference CPU kernels), values and gradients.

Test infrastructure only: the product never imports the reference."""
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
import cases  # noqa: E402

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not R.available(), reason="oracle/_ref/libdgl.so not built")]


@pytest.fixture(scope="module")
def registered():
    from dgl_hack_b200._lib import LIB, check_call
    R.set_num_threads(1)
    check_call(LIB.DGLB200RegisterInto(ctypes.c_void_p(R.lib()._handle)))
    R.forget_functions()
    return LIB


def _stats(LIB):
    s = (ctypes.c_int64 * 3)()
    LIB.DGLB200PackedStats(s)
    return list(s)


def test_dglfunccall_with_cuda_ndarrays(registered):
    import torch
    from oracle import oracle as O
    from dgl_hack_b200 import kernel as K
    LIB = registered
    ns = nd = 400
    ne = 9000
    src, dst = cases.powerlaw_graph(ns, nd, ne, 0.6, 5)
    og = O.Graph(ns, nd, src, dst)
    g = R.RefGraph(ns, src, dst)               # the REFERENCE's graph object (host index)
    rs = np.random.RandomState(1)
    x = rs.uniform(-1, 1, (ns, 8, 4)).astype(np.float32)
    w = rs.uniform(0.1, 1, (ne, 8, 1)).astype(np.float32)
    go = rs.uniform(-1, 1, (nd, 8, 4)).astype(np.float32)
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    xt, wt, got = t(x), t(w), t(go)
    s0, l0 = _stats(LIB), K.launch_count()

    # copy_src + sum / max forward and backward
    for red in ("sum", "max"):
        out = torch.empty((nd, 8, 4), device=dev)
        R.call("kernel._CAPI_DGLKernelCopyReduce", red, g.obj, 0, R.RefArray.from_dlpack(xt),
               R.RefArray.from_dlpack(out), R.null(), R.null())
        want = O.copy_reduce(red, og, 0, x, nd)
        assert np.array_equal(out.cpu().numpy(), want) if red == "max" else \
            np.allclose(out.cpu().numpy(), want, rtol=1e-5, atol=1e-5)
        gin = torch.empty_like(xt)
        R.call("kernel._CAPI_DGLKernelBackwardCopyReduce", red, g.obj, 0,
               R.RefArray.from_dlpack(xt), R.RefArray.from_dlpack(out),
               R.RefArray.from_dlpack(got), R.RefArray.from_dlpack(gin), R.null(), R.null())
        gwant = O.backward_copy_reduce(red, og, 0, x, want, go)
        assert np.allclose(gin.cpu().numpy(), gwant, rtol=1e-5, atol=1e-5)

    # u_mul_e + sum with a broadcast edge operand, and both gradients
    ret, code = R.call("kernel._CAPI_DGLKernelInferBinaryFeatureShape", "mul",
                       R.RefArray.from_dlpack(xt), R.RefArray.from_dlpack(wt))
    assert R.returned_int64_array(ret, code) == [8, 4]
    out = torch.empty((nd, 8, 4), device=dev)
    R.call("kernel._CAPI_DGLKernelBinaryOpReduce", "sum", "mul", g.obj, 0, 2,
           R.RefArray.from_dlpack(xt), R.RefArray.from_dlpack(wt), R.RefArray.from_dlpack(out),
           R.null(), R.null(), R.null())
    want = O.binary_op_reduce("sum", "mul", og, 0, 2, x, w, nd)
    assert np.allclose(out.cpu().numpy(), want, rtol=1e-5, atol=1e-5)
    gl, gr = torch.empty((ns, 8, 4), device=dev), torch.empty((ne, 8, 4), device=dev)
    for name, grad in (("kernel._CAPI_DGLKernelBackwardLhsBinaryOpReduce", gl),
                       ("kernel._CAPI_DGLKernelBackwardRhsBinaryOpReduce", gr)):
        R.call(name, "sum", "mul", g.obj, 0, 2, R.null(), R.null(), R.null(),
               R.RefArray.from_dlpack(xt), R.RefArray.from_dlpack(wt),
               R.RefArray.from_dlpack(out), R.RefArray.from_dlpack(got),
               R.RefArray.from_dlpack(grad))
    wl = O.backward_lhs_binary_op_reduce("sum", "mul", og, 0, 2, x, w, want, go)
    wr = O.backward_rhs_binary_op_reduce("sum", "mul", og, 0, 2, x, w, want, go)
    assert np.allclose(gl.cpu().numpy(), wl, rtol=1e-5, atol=1e-5)
    assert np.allclose(gr.cpu().numpy(), wr, rtol=1e-5, atol=1e-5)

    s1 = _stats(LIB)
    assert s1[0] - s0[0] == 8 and s1[1] == s0[1], "every call went through a shim, none to the CPU"
    assert s1[2] - s0[2] == 1, "the reference graph was indexed once and cached"
    assert K.launch_count() > l0, "the B200 kernels ran"

    # errors travel the reference's way (DGLFuncCall -> -1, DGLGetLastError)
    bad = torch.empty((nd + 1, 8, 4), device=dev)
    with pytest.raises(RuntimeError):
        R.call("kernel._CAPI_DGLKernelCopyReduce", "sum", g.obj, 0, R.RefArray.from_dlpack(xt),
               R.RefArray.from_dlpack(bad.double()), R.null(), R.null())
    with pytest.raises(RuntimeError, match="mean"):
        R.call("kernel._CAPI_DGLKernelCopyReduce", "mean", g.obj, 0, R.RefArray.from_dlpack(xt),
               R.RefArray.from_dlpack(out), R.null(), R.null())


def test_typed_csr_constructor_matches_set_edge_types():
    """DGLB200GraphCreateFromTypedCSR (what the RGCN shims feed with the reference's
    _CAPI_DGLGraphGetCsrSortedByEdgeType arrays) rebuilds the same typed graph."""
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200._lib import LIB, check_call, GraphHandle
    n, e, r = 300, 5000, 7
    rs = np.random.RandomState(2)
    src, dst, ety = rs.randint(0, n, e), rs.randint(0, n, e), rs.randint(0, r, e)
    g = B.GraphIndex.from_coo_with_type(n, src, dst, ety, r, "cuda:0")
    arrs = g.get_csr_sorted_by_edge_type(False)
    for bits, conv in ((32, lambda a: a), (64, lambda a: a.long().cpu())):
        a4 = [conv(a).contiguous() for a in arrs]
        h = GraphHandle()
        check_call(LIB.DGLB200GraphCreateFromTypedCSR(
            n, e, bits, *[ctypes.c_void_p(a.data_ptr()) for a in a4], -1, 0, None,
            ctypes.byref(h)))
        g2 = B.GraphIndex(h, torch.device("cuda:0"))
        g2.num_types = r
        for t in (False, True):
            for x, y in zip(g.get_csr_sorted_by_edge_type(t), g2.get_csr_sorted_by_edge_type(t)):
                assert torch.equal(x, y)
        for x, y in zip(g.in_csr + g.out_csr, g2.in_csr + g2.out_csr):
            assert np.array_equal(x, y)


def test_reference_python_stack_on_b200_kernels():
    pkg = os.path.join(ROOT, "oracle", "_ref", "python")
    if not os.path.isdir(os.path.join(pkg, "dgl")):
        pytest.skip("reference python package not installed under oracle/_ref/python")
    env = dict(os.environ, OMP_NUM_THREADS="1", DGLBACKEND="pytorch",
               DGL_LIBRARY_PATH=os.path.join(ROOT, "oracle", "_ref"), PYTHONPATH=pkg)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_python_driver.py"),
                          "--device", "cuda:0"], env=env, capture_output=True, text=True,
                         timeout=900)
    assert out.returncode == 0, out.stderr[-3000:]
    rec = json.loads(out.stdout.strip().splitlines()[-1])
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "ref_python_on_b200.json"), "w") as f:
        json.dump(rec, f, indent=1)
    bad = {k: v for k, v in rec["cases"].items() if not (isinstance(v, float) and v <= 2e-5)}
    assert not bad, bad
    assert rec["b200_launches"] > 50 and rec["graphs_built"] >= 3
    # roughly half of the shim calls (the CUDA half) stayed in this library
    assert rec["shim_calls"] - rec["forwarded_to_reference_cpu"] > 50

"""Index preparation on the device (COO->CSR, transpose, degrees, narrowing): bit-exact against
the reference's known-answer fixtures (tests/cpp/test_spmat.cc) and against the oracle."""
import ctypes

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a, np.int32)).cuda()


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _coo_to_csr(n_rows, row, col):
    import torch
    from dgl_hack_b200._lib import LIB, check_call
    row, col = _dev(row), _dev(col)
    nnz = row.shape[0]
    ip = torch.empty(n_rows + 1, dtype=torch.int32, device="cuda")
    ix = torch.empty(nnz, dtype=torch.int32, device="cuda")
    dt = torch.empty(nnz, dtype=torch.int32, device="cuda")
    check_call(LIB.DGLB200COOToCSR(n_rows, nnz, _p(row), _p(col), _p(ip), _p(ix), _p(dt), None))
    torch.cuda.synchronize()
    return ip.cpu().tolist(), ix.cpu().tolist(), dt.cpu().tolist()


def _transpose(n, m, indptr, indices, data):
    import torch
    from dgl_hack_b200._lib import LIB, check_call
    ip, ix = _dev(indptr), _dev(indices)
    dt = _dev(data) if data is not None else None
    nnz = ix.shape[0]
    bp = torch.empty(m + 1, dtype=torch.int32, device="cuda")
    bi = torch.empty(nnz, dtype=torch.int32, device="cuda")
    bx = torch.empty(nnz, dtype=torch.int32, device="cuda")
    check_call(LIB.DGLB200CSRTranspose(n, m, nnz, _p(ip), _p(ix), _p(dt) if dt is not None else None,
                                       _p(bp), _p(bi), _p(bx), None))
    torch.cuda.synchronize()
    return bp.cpu().tolist(), bi.cpu().tolist(), bx.cpu().tolist()


CSR1 = dict(n=4, m=5, indptr=[0, 2, 3, 5, 5], indices=[1, 2, 0, 2, 3], data=[0, 2, 3, 1, 4])
CSR2 = dict(n=4, m=5, indptr=[0, 3, 4, 6, 6], indices=[1, 2, 2, 0, 2, 3], data=[0, 2, 5, 3, 1, 4])


def test_coo_to_csr_known_answer():
    # TestCOOToCSR, tests/cpp/test_spmat.cc:357-421
    assert _coo_to_csr(4, [0, 2, 0, 1, 2], [1, 2, 2, 0, 3]) == \
        (CSR1["indptr"], CSR1["indices"], CSR1["data"])
    assert _coo_to_csr(4, [0, 2, 0, 1, 2, 0], [1, 2, 2, 0, 3, 2]) == \
        (CSR2["indptr"], CSR2["indices"], CSR2["data"])


def test_csr_transpose_known_answer():
    # TestCSRTranspose, tests/cpp/test_spmat.cc:229-252
    assert _transpose(4, 5, CSR2["indptr"], CSR2["indices"], CSR2["data"]) == \
        ([0, 1, 2, 5, 6, 6], [1, 0, 0, 0, 2, 2], [3, 0, 2, 5, 1, 4])


def test_csr_to_coo_and_narrow():
    import torch
    from dgl_hack_b200._lib import LIB, check_call
    for csr, want in ((CSR1, [0, 0, 1, 2, 2]), (CSR2, [0, 0, 0, 1, 2, 2])):
        ip = _dev(csr["indptr"])
        row = torch.empty(len(want), dtype=torch.int32, device="cuda")
        check_call(LIB.DGLB200CSRToCOORows(csr["n"], len(want), _p(ip), _p(row), None))
        assert row.cpu().tolist() == want  # TestCSRToCOO, test_spmat.cc:254-281
    a = torch.arange(-5, 1000, dtype=torch.int64, device="cuda") * 7
    b = torch.empty(a.shape[0], dtype=torch.int32, device="cuda")
    check_call(LIB.DGLB200AsInt32(a.shape[0], _p(a), _p(b), None))
    assert torch.equal(b.to(torch.int64), a)  # TestAsNumBits, tests/cpp/test_aten.cc:64-121


@pytest.mark.parametrize("ns,nd,ne,alpha,seed", [
    (1, 1, 1, 0.0, 1), (5, 3, 0, 0.0, 2), (300, 300, 6000, 0.6, 11), (200, 350, 5000, 0.4, 12),
    (70000, 66000, 900000, 0.8, 13),   # > 65536 rows: 3 radix passes, multi-block scans
    (50, 40, 200000, 1.2, 14),         # long rows, many duplicate edges
])
def test_graph_index_vs_oracle(ns, nd, ne, alpha, seed):
    from oracle import oracle as O
    import dgl_hack_b200 as B
    src, dst = cases.powerlaw_graph(ns, nd, ne, alpha, seed)
    og = O.Graph(ns, nd, src, dst)
    gg = B.GraphIndex.from_coo(ns, nd, src, dst)
    for got, want in ((gg.in_csr, og.in_csr), (gg.out_csr, og.out_csr)):
        for a, b in zip(got, want):
            assert np.array_equal(a.astype(np.int64), b)
    assert np.array_equal(gg.in_degrees().cpu().numpy(), og.in_degrees())
    assert np.array_equal(gg.out_degrees().cpu().numpy(), og.out_degrees())
    # from_csr (wrapping reference-made CSRs) gives the same index
    g2 = B.GraphIndex.from_csr(ns, nd, og.in_csr, og.out_csr)
    for a, b in zip(g2.in_csr + g2.out_csr, og.in_csr + og.out_csr):
        assert np.array_equal(a.astype(np.int64), b)


def test_device_resident_coo_input():
    import torch
    import dgl_hack_b200 as B
    from oracle import oracle as O
    src, dst = cases.powerlaw_graph(1000, 1000, 30000, 0.5, 3)
    gg = B.GraphIndex.from_coo(1000, 1000, torch.from_numpy(src).cuda(), torch.from_numpy(dst).cuda())
    og = O.Graph(1000, 1000, src, dst)
    for a, b in zip(gg.in_csr + gg.out_csr, og.in_csr + og.out_csr):
        assert np.array_equal(a.astype(np.int64), b)


@pytest.mark.parametrize("bad", [("src", 1000), ("dst", -1), ("src", 2 ** 31 + 5), ("dst", 2 ** 40)])
@pytest.mark.parametrize("on_device", [False, True])
def test_out_of_range_ids_are_rejected(bad, on_device):
    """Ids outside [0, num_src) x [0, num_dst) (negative, too large, high bits set) must be an
    error, never an out-of-bounds device write: host arrays fail in the constructor, device
    arrays at the handle's first synchronisation point (no extra sync on the sampling path)."""
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200._lib import DGLError
    src, dst = cases.powerlaw_graph(1000, 1000, 30000, 0.5, 3)
    src, dst = src.copy(), dst.copy()
    (src if bad[0] == "src" else dst)[123] = bad[1]
    s, d = torch.from_numpy(src), torch.from_numpy(dst)
    if on_device:
        s, d = s.cuda(), d.cuda()
        g = B.GraphIndex.from_coo(1000, 1000, s, d)
        with pytest.raises(DGLError):
            g.in_degrees()
    else:
        with pytest.raises(DGLError):
            B.GraphIndex.from_coo(1000, 1000, s, d)
    # the library is still healthy afterwards
    ok = B.GraphIndex.from_coo(10, 10, torch.arange(10), torch.arange(10))
    assert int(ok.in_degrees().sum()) == 10


def test_bad_csr_arrays_are_rejected():
    import dgl_hack_b200 as B
    from dgl_hack_b200._lib import DGLError
    from oracle import oracle as O
    src, dst = cases.powerlaw_graph(50, 60, 500, 0.5, 4)
    og = O.Graph(50, 60, src, dst)
    for which, pos, val in ((1, 7, 50), (0, 3, 10 ** 6), (2, 0, -2), (3, 50, 499)):
        arrs = [a.copy() for a in og.in_csr + og.out_csr]
        arrs[which][pos] = val
        with pytest.raises(DGLError):
            B.GraphIndex.from_csr(50, 60, tuple(arrs[:3]), tuple(arrs[3:]))


def test_index_memory_is_two_csrs():
    """a graph built from COO keeps 4 x E ids + two indptr arrays resident (what
    the reference's in-CSR + out-CSR hold, immutable_graph.h:242-246) through forward + backward
    message passing; the plain in-CSR order, rebuilt on demand, is still the reference's."""
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F
    from oracle import oracle as O
    ns, nd, ne = 4000, 3000, 90000
    src, dst = cases.powerlaw_graph(ns, nd, ne, 0.6, 11)
    g = B.GraphIndex.from_coo(ns, nd, src, dst, "cuda:0")
    two_csrs = 4 * (4 * ne + (ns + 1) + (nd + 1))
    assert g.index_bytes() == two_csrs
    x = torch.rand(ns, 64, device="cuda").requires_grad_(True)
    for red in ("sum", "max", "mean"):
        out = F.copy_reduce(red, g, 0, x, nd)
        out.backward(torch.rand_like(out))
    w = torch.rand(ne, 1, device="cuda")
    F.binary_reduce("sum", "mul", g, 0, 2, x.detach(), w, nd)
    assert g.index_bytes() == two_csrs, "message passing must not grow the index"
    og = O.Graph(ns, nd, src, dst)
    for a, b in zip(g.in_csr + g.out_csr, og.in_csr + og.out_csr):
        assert np.array_equal(a.astype(np.int64), b)
    # + the plain in-CSR (and the out-CSR's row expansion its rebuild used), now resident
    assert two_csrs + 4 * 2 * ne <= g.index_bytes() <= two_csrs + 4 * 3 * ne

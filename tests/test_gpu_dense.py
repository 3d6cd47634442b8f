"""Opt-in dense projections (dgl_hack_b200.dense: fp32 GEMM emulated with BF16 tensor-core products
through the toolkit's cuBLASLt 12.9): values and gradients against an fp64 product within the
path's fp32 tolerance (1e-5 relative; measured ~5e-7), every transpose form, and GraphConv with
B200_DENSE=bf16x9 against the default route."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def dense():
    from dgl_hack_b200 import dense as D
    try:
        D._Lt.get()
    except Exception as ex:  # the toolkit's cuBLASLt >= 12.9 is an optional dependency
        pytest.skip("emulated-fp32 cuBLASLt not available: %s" % ex)
    return D


def _rel(got, want):
    return float((got.double() - want).abs().max() / want.abs().max())


@pytest.mark.parametrize("M,K,N", [(300, 100, 256), (1000, 256, 48), (77, 13, 5), (4096, 256, 256)])
def test_matmul_values_and_gradients(dense, M, K, N):
    torch.manual_seed(M)
    a = (torch.rand(M, K, device=DEV) * 2 - 1).requires_grad_(True)
    b = (torch.rand(K, N, device=DEV) * 2 - 1).requires_grad_(True)
    g = torch.rand(M, N, device=DEV) * 2 - 1
    c = dense.matmul(a, b)
    c.backward(g)
    ad, bd, gd = a.detach().double(), b.detach().double(), g.double()
    assert _rel(c.detach(), ad @ bd) < 1e-5
    assert _rel(a.grad, gd @ bd.t()) < 1e-5
    assert _rel(b.grad, ad.t() @ gd) < 1e-5


def test_gemm_edge_cases_and_errors(dense):
    z = dense.gemm(torch.zeros(0, 4, device=DEV), torch.zeros(4, 3, device=DEV))
    assert z.shape == (0, 3)
    z = dense.gemm(torch.zeros(5, 0, device=DEV), torch.zeros(0, 3, device=DEV))
    assert z.shape == (5, 3) and float(z.abs().sum()) == 0
    with pytest.raises(Exception, match="inner dimensions"):
        dense.gemm(torch.zeros(5, 4, device=DEV), torch.zeros(3, 3, device=DEV))
    with pytest.raises(Exception, match="no CPU path"):
        dense.gemm(torch.zeros(5, 4), torch.zeros(4, 3))


def test_graphconv_route(dense, monkeypatch):
    import dgl_hack_b200.nn as nnb
    from dgl_hack_b200.graph import Graph
    torch.manual_seed(1)
    n, e = 2000, 30000
    g = Graph(torch.randint(0, n, (e,)), torch.randint(0, n, (e,)), num_nodes=n, device=DEV)
    outs = []
    for din, dout in ((100, 256), (256, 47)):
        conv = nnb.GraphConv(din, dout).to(DEV)
        x = torch.rand(n, din, device=DEV)
        monkeypatch.delenv("B200_DENSE", raising=False)
        y0 = conv(g, x)
        monkeypatch.setenv("B200_DENSE", "bf16x9")
        x1 = x.clone().requires_grad_(True)
        y1 = conv(g, x1)
        y1.sum().backward()
        assert x1.grad is not None and conv.weight.grad is not None
        outs.append(float((y0 - y1).abs().max() / y0.abs().max()))
    assert max(outs) < 1e-5, outs

"""Fused GAT kernels (C ABI: DGLB200FusedGatKernel / DGLB200KernelBackwardFusedGat) against the
oracle's restatement of the fork's CUDA kernels, and the autograd wrapper backend.fused_gat.
fp32 tolerance 1e-5 relative to the magnitude of the summed terms (expf differs by <= 2 ulp
between CUDA and glibc, so bit-exactness is not expected here)."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _close(name, got, want, scale):
    err = np.abs(got.astype(np.float64) - want.astype(np.float64))
    tol = 1e-5 * np.maximum(scale, 1.0) + 1e-6
    assert (err <= tol).all(), "%s: max err %g (tol %g)" % (name, err.max(), tol.max())


@pytest.mark.parametrize("H,F", [(8, 16), (1, 64), (4, 32), (8, 1), (2, 256), (3, 8), (5, 2),
                                 (1, 1), (16, 4)])
@pytest.mark.parametrize("graph", ["square", "bipartite"])
@pytest.mark.parametrize("path", ["lane", "warp"])
def test_fused_gat_vs_oracle(H, F, graph, path, monkeypatch):
    # "lane": head-slice-per-lane kernels where the shape allows (F % 4 == 0 ...), "warp": the
    # warp-per-row kernels for every shape
    monkeypatch.setenv("B200_GAT_PATH", path)
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import kernel as K
    from oracle import oracle as O
    ns, nd, ne, alpha, seed = (900, 900, 15000, 0.6, 21) if graph == "square" else (500, 800, 9000, 0.4, 22)
    src, dst = cases.powerlaw_graph(ns, nd, ne, alpha, seed)
    og = O.Graph(ns, nd, src, dst)
    gg = B.GraphIndex.from_coo(ns, nd, src, dst)
    rs = np.random.RandomState(H * 100 + F)
    ft = rs.uniform(-1, 1, (ns, H, F)).astype(np.float32)
    el = rs.uniform(-1, 1, (ns, H, 1)).astype(np.float32)
    er = rs.uniform(-1, 1, (nd, H, 1)).astype(np.float32)
    gout = rs.uniform(-1, 1, (nd, H, F)).astype(np.float32)
    s, ex, ret = O.fused_gat_forward(og, ft, el, er, 0.2)
    T = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    tft, tel, ter = T(ft), T(el), T(er)
    ts = torch.full((nd, H, 1), float("nan"), device="cuda")
    tex = torch.full((ne, H, 1), float("nan"), device="cuda")
    tret = torch.full((nd, H, F), float("nan"), device="cuda")
    K.fused_gat_kernel(gg, tft, tel, ter, ts, tex, tret, 0.2)
    assert K.last_kernel() == "gat_fwd"
    _close("exp", tex.cpu().numpy(), ex, np.abs(ex))
    _close("sum", ts.cpu().numpy(), s, np.abs(s))
    _close("ret", tret.cpu().numpy(), ret, 1.0)
    # backward on the ORACLE's forward state, so only the backward kernels are compared
    gft, gel, ger = O.fused_gat_backward(og, ft, el, er, s, ex, ret, gout, 0.2)
    gabs = O.fused_gat_backward(og, np.abs(ft) + np.abs(ret).max(), el, er, s, ex,
                                np.zeros_like(ret), np.abs(gout), 0.2)
    tg = [torch.full(a.shape, float("nan"), device="cuda") for a in (ft, el, er)]
    K.backward_fused_gat(gg, tft, tel, ter, T(s), T(ex), T(ret), T(gout), tg[0], tg[1], tg[2], 0.2)
    _close("grad_ft", tg[0].cpu().numpy(), gft, gabs[0])
    _close("grad_el", tg[1].cpu().numpy(), gel, gabs[1])
    _close("grad_er", tg[2].cpu().numpy(), ger, gabs[2])


def test_fused_gat_autograd_matches_builtin_composition():
    """backend.fused_gat == u_add_v -> leaky_relu -> exp -> copy_e+sum -> e_div_v -> u_mul_e+sum
    composed from our own builtin kernels (forward and all three gradients)."""
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F_
    n, ne = 2000, 30000
    src, dst = cases.powerlaw_graph(n, n, ne, 0.4, 5)
    g = B.GraphIndex.from_coo(n, n, src, dst)
    torch.manual_seed(0)
    mk = lambda *s: (torch.rand(*s, device="cuda") * 2 - 1).requires_grad_(True)  # noqa: E731
    ft, el, er = mk(n, 8, 16), mk(n, 8, 1), mk(n, 8, 1)
    gout = torch.rand(n, 8, 16, device="cuda")
    out = F_.fused_gat(g, ft, el, er, 0.2)
    out.backward(gout)
    got = [out.detach().clone(), ft.grad.clone(), el.grad.clone(), er.grad.clone()]
    for t in (ft, el, er):
        t.grad = None
    e = F_.binary_reduce("none", "add", g, 0, 1, el, er, ne)
    e = torch.nn.functional.leaky_relu(e, 0.2).exp()
    s = F_.copy_reduce("sum", g, 2, e, n)
    a = F_.binary_reduce("none", "div", g, 2, 1, e, s, ne)
    ref = F_.binary_reduce("sum", "mul", g, 0, 2, ft, a, n)
    ref.backward(gout)
    want = [ref.detach(), ft.grad, el.grad, er.grad]
    for name, a_, b_ in zip(("out", "grad_ft", "grad_el", "grad_er"), got, want):
        torch.testing.assert_close(a_, b_, rtol=2e-4, atol=2e-4, msg=lambda m: name + ": " + m)


def test_fused_gat_rejects_unsupported_width():
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import kernel as K
    g = B.GraphIndex.from_coo(4, 4, [0, 1, 2], [1, 2, 3])
    z = lambda *s: torch.zeros(*s, device="cuda")  # noqa: E731
    with pytest.raises(B.DGLError, match="power of two"):
        K.fused_gat_kernel(g, z(4, 2, 6), z(4, 2, 1), z(4, 2, 1), z(4, 2, 1), z(3, 2, 1),
                           z(4, 2, 6), 0.2)


@pytest.mark.parametrize("H,F", [(8, 16), (1, 64), (3, 5)])
def test_edge_scalar_gradient_is_fused_dot(H, F):
    """backend.binary_reduce backward w.r.t. a [E,H,1] edge operand of u_mul_e+sum: the fused
    per-(edge, head) dot equals the reference route (unreduced [E,H,F] gradient, then a sum over F)."""
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F_
    from dgl_hack_b200 import kernel as K
    n, ne = 1500, 20000
    src, dst = cases.powerlaw_graph(n, n, ne, 0.5, 9)
    g = B.GraphIndex.from_coo(n, n, src, dst)
    torch.manual_seed(1)
    ft = (torch.rand(n, H, F, device="cuda") * 2 - 1)
    a = torch.rand(ne, H, 1, device="cuda").requires_grad_(True)
    gout = torch.rand(n, H, F, device="cuda") * 2 - 1
    out = F_.binary_reduce("sum", "mul", g, 0, 2, ft, a, n)
    out.backward(gout)
    unreduced = torch.empty(ne, H, F, device="cuda")
    K.backward_rhs_binary_op_reduce("sum", "mul", g, 0, 2, ft, a.detach(), out.detach(), gout,
                                    unreduced)
    want = unreduced.sum(dim=-1, keepdim=True)
    torch.testing.assert_close(a.grad, want, rtol=1e-5, atol=1e-5 * F)

"""CUDA RGCN hack ops (typed adjacency + RgcnLayer0/1 fwd/bwd) against the oracle restatement."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _case(n, e, r, seed, hub=True):
    rng = np.random.RandomState(seed)
    src = rng.randint(0, n, e)
    dst = rng.randint(0, n, e)
    if hub:
        dst[: e // 6] = 2
        src[e // 2: e // 2 + e // 8] = 5
    # Zipf-like relation frequencies, some relations unused
    p = 1.0 / np.arange(1, r + 1) ** 1.1
    et = rng.choice(r, size=e, p=p / p.sum())
    et[et == r - 2] = 0 if r > 3 else et[et == r - 2]
    norm = rng.rand(e, 1).astype(np.float32)
    return src, dst, et, norm, rng


def _graphs(n, src, dst, et, r):
    import dgl_hack_b200 as B
    from oracle import oracle as O
    g = B.GraphIndex.from_coo_with_type(n, src, dst, et, r, DEV)
    return g, O.Graph(n, n, src, dst)


@pytest.mark.parametrize("n,e,r", [(50, 700, 7), (3000, 40000, 133), (10, 0, 3)])
def test_typed_csr_bit_exact(n, e, r):
    from oracle import oracle as O
    src, dst, et, _, _ = _case(n, e, r, 3, hub=e > 0)
    g, og = _graphs(n, src, dst, et, r)
    for transpose in (False, True):
        got = [t.cpu().numpy() for t in g.get_csr_sorted_by_edge_type(transpose)]
        exp = O.typed_csr(og, et, transpose)
        for a, b in zip(got, exp):
            assert a.dtype == np.int32
            np.testing.assert_array_equal(a, b)


def test_edge_type_out_of_range_is_refused():
    import dgl_hack_b200 as B
    g = B.GraphIndex.from_coo(4, 4, [0, 1, 2], [1, 2, 3], DEV)
    with pytest.raises(B.DGLError, match="out of range"):
        g.set_edge_types([0, 5, 1], 3)
    from dgl_hack_b200 import backend as F
    with pytest.raises(B.DGLError, match="no edge types"):
        F.rgcn_layer1(g, torch.zeros(4, 4, device=DEV), torch.zeros(3, 4, 4, device=DEV),
                      torch.ones(3, 1, device=DEV))


@pytest.mark.parametrize("fy,fx", [(16, 16), (5, 9), (32, 32), (8, 100), (64, 16), (7, 3)])
@pytest.mark.parametrize("n,e,r", [(300, 6000, 11), (5000, 30000, 133)])
def test_layer1_forward_backward(fy, fx, n, e, r):
    from dgl_hack_b200 import backend as F
    from oracle import oracle as O
    src, dst, et, norm, rng = _case(n, e, r, 5)
    g, og = _graphs(n, src, dst, et, r)
    h = (rng.rand(n, fy) * 2 - 1).astype(np.float32)
    w = (rng.rand(r, fy, fx) * 2 - 1).astype(np.float32)
    go = (rng.rand(n, fx) * 2 - 1).astype(np.float32)
    ht = torch.tensor(h, device=DEV, requires_grad=True)
    wt = torch.tensor(w, device=DEV, requires_grad=True)
    nt = torch.tensor(norm, device=DEV)
    ret = F.rgcn_layer1(g, ht, wt, nt)
    exp = O.rgcn_layer1(og, et, h, w, norm)
    # hub rows sum up to e/6 terms of magnitude <= fy in a different order
    tol = dict(rtol=1e-4, atol=1e-5 * fy * max(1, e // 600))
    np.testing.assert_allclose(ret.detach().cpu().numpy(), exp, **tol)
    if fy * fx > 1024:
        return
    ret.backward(torch.tensor(go, device=DEV))
    gh, gw = O.rgcn_layer1_backward(og, et, h, w, norm, go)
    np.testing.assert_allclose(ht.grad.cpu().numpy(), gh, rtol=1e-4, atol=1e-5 * fx * max(1, e // 600))
    np.testing.assert_allclose(wt.grad.cpu().numpy(), gw, rtol=2e-4, atol=2e-4 * max(1, e // 6000))
    assert np.all(wt.grad.cpu().numpy()[np.bincount(et, minlength=r) == 0] == 0)


def test_layer1_backward_is_deterministic():
    from dgl_hack_b200 import kernel as K
    src, dst, et, norm, rng = _case(4000, 90000, 40, 8)
    g, _ = _graphs(4000, src, dst, et, 40)
    h = torch.rand(4000, 16, device=DEV)
    w = torch.rand(40, 16, 16, device=DEV)
    go = torch.rand(4000, 16, device=DEV)
    nt = torch.tensor(norm, device=DEV)
    outs = []
    for _ in range(3):
        gh, gw = torch.empty_like(h), torch.empty_like(w)
        K.rgcn_layer1_backward(g, h, w, nt, go, gh, gw)
        outs.append((gh.clone(), gw.clone()))
    for gh, gw in outs[1:]:
        assert torch.equal(gh, outs[0][0]) and torch.equal(gw, outs[0][1])


@pytest.mark.parametrize("f", [4, 12, 16, 40, 130])
def test_layer0_forward_backward(f):
    from dgl_hack_b200 import backend as F
    from dgl_hack_b200 import kernel as K
    from oracle import oracle as O
    n, e, r = 200, 3000, 9
    src, dst, et, norm, rng = _case(n, e, r, 6)
    g, og = _graphs(n, src, dst, et, r)
    w = (rng.rand(r, n, f) * 2 - 1).astype(np.float32)
    go = (rng.rand(n, f) * 2 - 1).astype(np.float32)
    wt = torch.tensor(w, device=DEV, requires_grad=True)
    nt = torch.tensor(norm, device=DEV)
    ret = F.rgcn_layer0(g, wt, nt)
    np.testing.assert_allclose(ret.detach().cpu().numpy(), O.rgcn_layer0(og, et, w, norm),
                               rtol=1e-4, atol=1e-4)
    ret.backward(torch.tensor(go, device=DEV))
    np.testing.assert_allclose(wt.grad.cpu().numpy(),
                               O.rgcn_layer0_backward(og, et, go, norm, w.shape, True),
                               rtol=1e-4, atol=1e-4)
    # the reference's store-last behaviour, bit for bit (one product per written element)
    gw = torch.zeros_like(wt)
    K.rgcn_layer0_backward(g, torch.tensor(go, device=DEV), nt, gw, accumulate=False)
    np.testing.assert_array_equal(gw.cpu().numpy(),
                                  O.rgcn_layer0_backward(og, et, go, norm, w.shape, False))

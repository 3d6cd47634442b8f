"""Oracle restatement of the RGCN hack ops (a15) against independent formulas (CPU).

Parity for these ops is UNPINNED against the reference binary (CUDA-only kernels, no reference
test, typed-CSR builder hard-codes a GPU copy); the restatement is checked here against a dense
float64 formula and against the builtin per-relation composition of the pinned oracle kernels
(the RelGraphConv low-memory path, python/dgl/nn/pytorch/conv/relgraphconv.py:122-137).
"""
import numpy as np
import pytest

from oracle import oracle as O


def _case(n, e, r, seed):
    rng = np.random.RandomState(seed)
    src = rng.randint(0, n, e)
    dst = rng.randint(0, n, e)
    dst[: e // 6] = 2  # hub with many types
    et = rng.randint(0, r, e)
    norm = rng.rand(e, 1).astype(np.float32)
    return O.Graph(n, n, src, dst), src, dst, et, norm, rng


def test_typed_csr_order():
    g, src, dst, et, _, _ = _case(50, 600, 7, 0)
    for transpose in (False, True):
        ip, ids, eids, ty = O.typed_csr(g, et, transpose)
        key, other = (src, dst) if transpose else (dst, src)
        assert ip[-1] == 600 and sorted(eids.tolist()) == list(range(600))
        assert np.array_equal(ids, other[eids]) and np.array_equal(ty, et[eids])
        for v in range(50):
            row = eids[ip[v]:ip[v + 1]]
            assert np.all(key[row] == v)
            k = list(zip(et[row].tolist(), row.tolist()))
            assert k == sorted(k)  # (type, edge id)
        assert np.array_equal(np.diff(ip), np.bincount(key, minlength=50))


@pytest.mark.parametrize("fy,fx", [(16, 16), (5, 9), (32, 8)])
def test_layer1_matches_dense_and_builtin(fy, fx):
    n, e, r = 40, 500, 6
    g, src, dst, et, norm, rng = _case(n, e, r, 1)
    h = (rng.rand(n, fy) * 2 - 1).astype(np.float32)
    w = (rng.rand(r, fy, fx) * 2 - 1).astype(np.float32)
    go = (rng.rand(n, fx) * 2 - 1).astype(np.float32)
    ret = O.rgcn_layer1(g, et, h, w, norm)
    msg = np.einsum("ey,eyx->ex", h[src].astype(np.float64), w[et].astype(np.float64)) * norm
    dense = np.zeros((n, fx))
    np.add.at(dense, dst, msg)
    np.testing.assert_allclose(ret, dense, rtol=1e-4, atol=1e-4)
    # builtin composition: per-edge message (bmm) then copy_e + sum with the pinned oracle kernel
    built = O.copy_reduce("sum", g, O.T_EDGE, msg.astype(np.float32), n)
    np.testing.assert_allclose(ret, built, rtol=1e-4, atol=1e-4)
    gh, gw = O.rgcn_layer1_backward(g, et, h, w, norm, go)
    gmsg = go[dst].astype(np.float64) * norm
    gh_d = np.zeros((n, fy))
    np.add.at(gh_d, src, np.einsum("ex,eyx->ey", gmsg, w[et].astype(np.float64)))
    gw_d = np.zeros((r, fy, fx))
    np.add.at(gw_d, et, np.einsum("ey,ex->eyx", h[src].astype(np.float64), gmsg))
    np.testing.assert_allclose(gh, gh_d, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(gw, gw_d, rtol=1e-4, atol=2e-4)


def test_layer0_forward_backward():
    n, e, r, f = 30, 300, 4, 12
    g, src, dst, et, norm, rng = _case(n, e, r, 2)
    w = (rng.rand(r, n, f) * 2 - 1).astype(np.float32)
    go = (rng.rand(n, f) * 2 - 1).astype(np.float32)
    ret = O.rgcn_layer0(g, et, w, norm)
    dense = np.zeros((n, f))
    np.add.at(dense, dst, w[et, src].astype(np.float64) * norm)
    np.testing.assert_allclose(ret, dense, rtol=1e-4, atol=1e-4)
    gw = O.rgcn_layer0_backward(g, et, go, norm, w.shape, accumulate=True)
    gw_d = np.zeros(w.shape)
    np.add.at(gw_d, (et, src), go[dst].astype(np.float64) * norm)
    np.testing.assert_allclose(gw, gw_d, rtol=1e-4, atol=1e-4)
    # as written in the reference: the last edge (type, edge id order) of a (type, src) pair wins
    gw_s = O.rgcn_layer0_backward(g, et, go, norm, w.shape, accumulate=False)
    last = {}
    for eid in range(e):
        last[(et[eid], src[eid])] = eid
    exp = np.zeros(w.shape, np.float32)
    for (t, u), eid in last.items():
        exp[t, u] = go[dst[eid]] * norm[eid]
    np.testing.assert_array_equal(gw_s, exp)

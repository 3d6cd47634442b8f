"""GPU parity proper: the CUDA kernels, called through the C ABI, against
  (1) the committed golden outputs of the unmodified reference (tests/golden/*.npz), and
  (2) the CPU oracle on seeded inputs of sizes it finishes in seconds.
Tolerances (golden_util.RTOL): bit-exact for index arrays and max/min/none values, 1e-5 relative
for sums (north_star), 1e-4 for prod."""
import numpy as np
import pytest

import cases
import golden_util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def impl():
    from gpu_adapter import GpuImpl
    return GpuImpl


def test_kernel_graph_all_builtins(impl):
    # 450 binary cases (fwd + both grads) + 10 copy cases, test_kernel.py:200-362
    assert golden_util.replay_kernel_graph(impl) == 460


def test_mid_graphs(impl):
    golden_util.replay_mid(impl)


def test_id_mappings(impl):
    golden_util.replay_mapped(impl)


def test_empty_and_isolated(impl):
    g = impl.make_graph(3, 3, np.zeros(0, np.int64), np.zeros(0, np.int64))
    x = np.ones((3, 4), np.float32)
    assert np.array_equal(impl.copy_reduce("sum", g, 0, x, 3, None, None), np.zeros((3, 4)))
    assert (impl.copy_reduce("max", g, 0, x, 3, None, None) == np.finfo(np.float32).min).all()
    assert (impl.copy_reduce("min", g, 0, x, 3, None, None) == np.finfo(np.float32).max).all()
    assert (impl.copy_reduce("prod", g, 0, x, 3, None, None) == 1).all()
    e = np.zeros((0, 4), np.float32)
    assert impl.copy_reduce("none", g, 0, x, 0, None, None).shape == (0, 4)
    assert np.array_equal(impl.copy_reduce("sum", g, 2, e, 3, None, None), np.zeros((3, 4)))


@pytest.mark.parametrize("path", ["reg", "tma"])
@pytest.mark.parametrize("d", [1, 2, 3, 7, 16, 47, 100, 128, 130, 256, 602, 1000, 2048])
def test_spmm_widths_vs_oracle(impl, d, path, monkeypatch):
    """Every vector-width / lanes-per-row / slab shape of both SpMM kernels (register-staged and
    TMA bulk-copy staged; B200_SPMM_PATH forces one) against the oracle."""
    from oracle import oracle as O
    monkeypatch.setenv("B200_SPMM_PATH", path)
    ns, nd, ne = 700, 650, 9000
    src, dst = cases.powerlaw_graph(ns, nd, ne, 0.7, 100 + d)
    og, gg = O.Graph(ns, nd, src, dst), impl.make_graph(ns, nd, src, dst)
    rs = np.random.RandomState(d)
    x = rs.uniform(-1, 1, (ns, d)).astype(np.float32)
    w = rs.uniform(0.5, 1.5, (ne, d)).astype(np.float32)
    w1 = rs.uniform(0.5, 1.5, (ne, 1)).astype(np.float32)
    for red in ("sum", "max", "min", "prod"):
        want = O.copy_reduce(red, og, 0, x, nd)
        got = impl.copy_reduce(red, gg, 0, x, nd, None, None)
        assert impl.last.startswith("spmm")
        golden_util.check("copy_u_%s_d%d" % (red, d), got, want, red,
                          scale=None if red != "sum" else _abs_sum(og, x, nd))
    for name, b in (("full", w), ("scalar", w1)):
        if d == 1 and name == "scalar":
            continue
        for op in ("mul", "add", "sub", "div"):
            want = O.binary_op_reduce("sum", op, og, 0, 2, x, b, nd)
            got = impl.binary_op_reduce("sum", op, gg, 0, 2, x, b, nd, None, None, None)
            assert impl.last.startswith("spmm")
            golden_util.check("u_%s_e_%s_d%d" % (op, name, d), got, want, "sum",
                              scale=_abs_sum(og, np.abs(x) + 1.5, nd))
    # backward of copy_u+sum and u_mul_e+sum w.r.t. u: transposed SpMM over the out-CSR
    out = O.copy_reduce("sum", og, 0, x, nd)
    gout = cases.grad_out_for(out.shape)
    want = O.backward_copy_reduce("sum", og, 0, x, out, gout)
    got = impl.backward_copy_reduce("sum", gg, 0, x, out, gout, None, None)
    assert impl.last.startswith("spmm")
    # rows above the (size-dependent) segment length are regrouped: error relative to sum |terms|
    gscale = O.backward_copy_reduce("sum", og, 0, x, out, np.abs(gout))
    golden_util.check("bwd_copy_u_d%d" % d, got, want, "sum", grad=True, scale=gscale)
    out = O.binary_op_reduce("sum", "mul", og, 0, 2, x, w1, nd)
    want = O.backward_lhs_binary_op_reduce("sum", "mul", og, 0, 2, x, w1, out, gout)
    got = impl.backward_lhs_binary_op_reduce("sum", "mul", gg, 0, 2, x, w1, out, gout,
                                             None, None, None)
    golden_util.check("bwd_u_mul_e_d%d" % d, got, want, "sum", grad=True, scale=gscale * 1.5)
    if path == "tma" and d >= 33:
        assert impl.last == "spmm_tma"


def _abs_sum(og, x, nd):
    from oracle import oracle as O
    return O.copy_reduce("sum", og, 0, np.abs(x).astype(np.float32), nd).astype(np.float64)


@pytest.mark.parametrize("seg", ["0", "64", "4096"])
@pytest.mark.parametrize("path", ["reg", "tma"])
def test_hub_rows_are_segmented(impl, path, seg, monkeypatch):
    """A hub row far above the segment length is cut into segments reduced by separate warps and
    folded in order (deterministic; within the 1e-5 tolerance). With segmentation off
    (B200_SPMM_SEG=0) the whole row is reduced sequentially and must be bit-identical."""
    from oracle import oracle as O
    monkeypatch.setenv("B200_SPMM_PATH", path)
    monkeypatch.setenv("B200_SPMM_SEG", seg)
    rs = np.random.RandomState(5)
    ns = nd = 3000
    src = rs.randint(0, ns, 60000)
    dst = rs.randint(0, nd, 60000)
    dst[:20000] = 17  # hub: in-degree >= 20000 vs mean 20
    src[40000:] = 5   # and an out-degree hub for the transposed (backward) SpMM
    og, gg = O.Graph(ns, nd, src, dst), impl.make_graph(ns, nd, src, dst)
    for d in (64, 602, 5):
        x = rs.uniform(-1, 1, (ns, d)).astype(np.float32)
        w = rs.uniform(0.5, 1.5, (60000, 1)).astype(np.float32)
        for red in ("sum", "max", "prod"):
            xx = (np.abs(x) * 0.02 + 0.99).astype(np.float32) if red == "prod" else x
            want = O.copy_reduce(red, og, 0, xx, nd)
            got = impl.copy_reduce(red, gg, 0, xx, nd, None, None)
            if red == "sum":
                golden_util.check("hub", got, want, "sum", scale=_abs_sum(og, xx, nd))
            elif red == "max":
                assert np.array_equal(got, want)
            else:
                np.testing.assert_allclose(got, want, rtol=5e-3, atol=1e-30)  # 20000 terms, regrouped
            if seg == "0":
                assert np.array_equal(got, want), "sequential in-order reduce must be bit-exact"
        want = O.binary_op_reduce("sum", "mul", og, 0, 2, x, w, nd)
        got = impl.binary_op_reduce("sum", "mul", gg, 0, 2, x, w, nd, None, None, None)
        golden_util.check("hub_w", got, want, "sum", scale=_abs_sum(og, np.abs(x) * 1.5, nd))
        gout = cases.grad_out_for(want.shape)
        want = O.backward_copy_reduce("sum", og, 0, x, want, gout)
        got = impl.backward_copy_reduce("sum", gg, 0, x, want, gout, None, None)
        golden_util.check("hub_bwd", got, want, "sum",
                          scale=O.backward_copy_reduce("sum", og, 0, x, want, np.abs(gout)))
        if seg == "0":
            assert np.array_equal(got, want)


@pytest.mark.parametrize("d", [1, 2, 3, 8, 16, 47, 64, 130, 602])
def test_sddmm_widths_vs_oracle(impl, d):
    """Edge-output kernels (reducer none, edge-operand gradients under sum) at every vector width /
    lanes-per-edge shape and both traversal orders, against the oracle. Element-wise results have
    one rounding each -> exact; dot is reduced in a butterfly order -> tolerance."""
    from oracle import oracle as O
    ns, nd, ne = 300, 280, 5000
    src, dst = cases.powerlaw_graph(ns, nd, ne, 0.7, 200 + d)
    og, gg = O.Graph(ns, nd, src, dst), impl.make_graph(ns, nd, src, dst)
    rs = np.random.RandomState(d)
    xs = rs.uniform(0.5, 1.5, (ns, d)).astype(np.float32)
    xd = rs.uniform(0.5, 1.5, (nd, d)).astype(np.float32)
    xe = rs.uniform(0.5, 1.5, (ne, d)).astype(np.float32)
    for op in ("add", "sub", "mul", "div"):
        for (lt, rt, l, r) in ((0, 1, xs, xd), (1, 0, xd, xs), (2, 1, xe, xd), (0, 2, xs, xe)):
            want = O.binary_op_reduce("none", op, og, lt, rt, l, r, ne)
            got = impl.binary_op_reduce("none", op, gg, lt, rt, l, r, ne, None, None, None)
            assert impl.last == "sddmm", impl.last
            np.testing.assert_array_equal(got, want)
    for tgt, x in ((0, xs), (1, xd)):
        want = O.copy_reduce("none", og, tgt, x, ne)
        got = impl.copy_reduce("none", gg, tgt, x, ne, None, None)
        assert impl.last == "sddmm"
        np.testing.assert_array_equal(got, want)
    # dot over the last axis: [., H, d] . [., H, d] -> [E, H]
    for H in (1, 4):
        a = rs.uniform(-1, 1, (ns, H, d)).astype(np.float32)
        b = rs.uniform(-1, 1, (nd, H, d)).astype(np.float32)
        want = O.binary_op_reduce("none", "dot", og, 0, 1, a, b, ne)
        got = impl.binary_op_reduce("none", "dot", gg, 0, 1, a, b, ne, None, None, None)
        pow2 = (d % 4 == 0 and (d // 4) & (d // 4 - 1) == 0 and d // 4 <= 32) or d in (1, 2)
        if pow2:
            assert impl.last == "sddmm", impl.last
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5 * d)
    # gradients of the edge operand under sum: copy_e, u_mul_e (both sides), u_sub_e
    out = O.copy_reduce("sum", og, 2, xe, nd)
    gout = cases.grad_out_for(out.shape)
    want = O.backward_copy_reduce("sum", og, 2, xe, out, gout)
    got = impl.backward_copy_reduce("sum", gg, 2, xe, out, gout, None, None)
    assert impl.last == "sddmm"
    np.testing.assert_array_equal(got, want)
    for op in ("mul", "sub", "add"):
        out = O.binary_op_reduce("sum", op, og, 0, 2, xs, xe, nd)
        want = O.backward_rhs_binary_op_reduce("sum", op, og, 0, 2, xs, xe, out, gout)
        got = impl.backward_rhs_binary_op_reduce("sum", op, gg, 0, 2, xs, xe, out, gout,
                                                 None, None, None)
        assert impl.last == "sddmm"
        np.testing.assert_array_equal(got, want)
        out = O.binary_op_reduce("sum", op, og, 2, 0, xe, xs, nd)
        want = O.backward_lhs_binary_op_reduce("sum", op, og, 2, 0, xe, xs, out, gout)
        got = impl.backward_lhs_binary_op_reduce("sum", op, gg, 2, 0, xe, xs, out, gout,
                                                 None, None, None)
        assert impl.last == "sddmm"
        np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("d", [9, 47, 602])
def test_odd_width_repitch_is_bit_identical(d, monkeypatch):
    """Rows whose pitch is not a multiple of 16 bytes are gathered from a zero-padded copy with
    16-byte rows (spmm.cu, B200_SPMM_PAD): same terms, same order -> bit-identical to the
    scalar-load kernels, forward, scaled forward, gradient and max reducer. Hub rows are kept
    sequential here (B200_SPMM_SEG=0): the segment length of the default mode depends on the row
    pitch, so with segmentation the two routes agree to rounding, which is checked as well."""
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F, kernel as K
    ns, nd, ne = 2000, 1500, 60000          # mean in-degree 40: the padded path is taken
    src, dst = cases.powerlaw_graph(ns, nd, ne, 0.7, 21)
    g = B.GraphIndex.from_coo(ns, nd, src, dst, "cuda:0")
    rs = np.random.RandomState(d)
    x = torch.from_numpy(rs.uniform(-1, 1, (ns, d)).astype(np.float32)).cuda()
    w = torch.from_numpy(rs.uniform(0, 1, (ne, 1)).astype(np.float32)).cuda()
    go = torch.from_numpy(rs.uniform(-1, 1, (nd, d)).astype(np.float32)).cuda()
    cs = torch.from_numpy(rs.uniform(0.5, 1, ns).astype(np.float32)).cuda()
    rsc = torch.from_numpy(rs.uniform(0.5, 1, nd).astype(np.float32)).cuda()

    def run():
        xr = x.clone().requires_grad_(True)
        o1 = F.copy_reduce("sum", g, 0, xr, nd)
        o1.backward(go)
        o2 = F.copy_reduce("max", g, 0, x, nd)
        o3 = F.binary_reduce("sum", "mul", g, 0, 2, x, w, nd)
        o4 = F.scaled_copy_sum(g, x, cs, rsc)
        o5 = F.copy_reduce("mean", g, 0, x, nd)
        return [o1.detach(), xr.grad, o2, o3, o4, o5]
    monkeypatch.setenv("B200_SPMM_SEG", "0")
    monkeypatch.setenv("B200_SPMM_PAD", "0")
    plain = run()
    monkeypatch.setenv("B200_SPMM_PAD", "1")
    padded = run()
    for a, b in zip(plain, padded):
        assert torch.equal(a, b)
    monkeypatch.delenv("B200_SPMM_SEG")
    segmented = run()
    # every term is at most 1 in magnitude: sum|terms| <= the largest degree
    mag = float(max(g.in_degrees().max().item(), g.out_degrees().max().item()))
    for a, b in zip(plain, segmented):
        assert (a - b).abs().max().item() <= 1e-5 * max(mag, 1.0)


@pytest.mark.parametrize("d", [16, 48, 100, 256])
def test_short_row_kernel_is_bit_identical(d, monkeypatch):
    """Graphs with ~1-3 edges per row (relation-segment graphs, the halo-source pass of a
    partitioned aggregation) take the lean short-row kernel (spmm.cu spmm_short_rows_kernel):
    same sequential order per row, hub rows still through their segments -> bit-identical to
    the general kernel, forward, transposed, with degree scales and with an addend."""
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F, kernel as K
    ns, nd, ne = 30000, 40000, 70000
    rs_ = np.random.RandomState(d)
    src = rs_.randint(0, ns, ne)
    dst = rs_.randint(0, nd, ne)
    dst[:9000] = 77                       # one hub row: cut into segments in both routes
    g = B.GraphIndex.from_coo(ns, nd, src, dst, "cuda:0")
    x = torch.from_numpy(rs_.uniform(-1, 1, (ns, d)).astype(np.float32)).cuda()
    go = torch.from_numpy(rs_.uniform(-1, 1, (nd, d)).astype(np.float32)).cuda()
    cs = torch.from_numpy(rs_.uniform(0.5, 1, ns).astype(np.float32)).cuda()
    rsc = torch.from_numpy(rs_.uniform(0.5, 1, nd).astype(np.float32)).cuda()
    base = torch.from_numpy(rs_.uniform(-1, 1, (nd, d)).astype(np.float32)).cuda()

    def run():
        o1 = F.copy_reduce("sum", g, 0, x, nd)
        assert K.last_kernel().startswith("spmm")
        xr = x.clone().requires_grad_(True)
        F.copy_reduce("sum", g, 0, xr, nd).backward(go)
        o3 = torch.empty(nd, d, device="cuda")
        K.scaled_copy_sum(g, False, x, None, rsc, o3)
        o4 = base.clone()
        K.scaled_copy_sum(g, False, x, None, rsc, o4, addend=o4)
        o5 = torch.empty(nd, d, device="cuda")
        K.scaled_copy_sum(g, False, x, cs, rsc, o5)          # a scale per gathered row as well
        o6 = torch.empty(ns, d, device="cuda")
        K.scaled_copy_sum(g, True, go, rsc, cs, o6)           # its gradient: transposed walk
        return [o1, xr.grad, o3, o4, o5, o6]
    monkeypatch.setenv("B200_SPMM_SHORT", "0")
    general = run()
    monkeypatch.setenv("B200_SPMM_SHORT", "1")
    lean = run()
    for a, b in zip(general, lean):
        assert torch.equal(a, b)
    # and against the definition
    want = torch.zeros(nd, d, device="cuda", dtype=torch.float64).index_add_(
        0, torch.from_numpy(dst).cuda(), x.double()[torch.from_numpy(src).cuda()])
    assert (lean[0].double() - want).abs().max().item() <= 1e-5 * 9000
    assert torch.allclose(lean[3], (base + lean[0]) * rsc.view(-1, 1), rtol=1e-5, atol=1e-4)

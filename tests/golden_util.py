"""Replays the committed golden fixtures (tests/golden/*.npz, outputs of the unmodified
reference, see tests/golden/make_golden.py) against an implementation adapter.

An adapter exposes numpy-in / numpy-out functions with the reference's kernel.py meaning:
    make_graph(n_src, n_dst, src, dst) -> g with .in_csr/.out_csr as 3 int arrays each
    binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out_rows, lhs_map, rhs_map, out_map)
    backward_lhs_binary_op_reduce(red, op, g, lt, rt, lhs, rhs, out, gout, lmap, rmap, omap)
    backward_rhs_binary_op_reduce(... same ...)
    copy_reduce(red, g, tgt, x, out_rows, in_map, out_map)
    backward_copy_reduce(red, g, tgt, x, out, gout, in_map, out_map)
Both the oracle (tests/test_oracle_golden.py, CPU) and the CUDA product
(tests/test_gpu_golden.py, -m gpu) are checked through this one replay.
"""
import os

import numpy as np

import cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# Tolerances: bit-exact for index arrays and for max/min/none *values*; sums within 1e-5
# relative (north_star) of sum|terms| per output element (term_scale; no absolute floor);
# prod 1e-4 of the value.
RTOL = {"sum": 1e-5, "prod": 1e-4, "max": 0.0, "min": 0.0, "none": 0.0}


def load(name):
    return np.load(os.path.join(GOLDEN, name))


def _oracle_twin(g):
    """The oracle's graph of the same edges (cached on g), for the tolerance scales below."""
    from oracle import oracle as O
    if isinstance(g, O.Graph):
        return g
    twin = getattr(g, "_oracle_twin", None)
    if twin is None:
        indptr, indices, eids = (np.asarray(a).astype(np.int64) for a in g.in_csr)
        n_dst = indptr.shape[0] - 1
        e = indices.shape[0]
        src, dst = np.empty(e, np.int64), np.empty(e, np.int64)
        src[eids] = indices
        dst[eids] = np.repeat(np.arange(n_dst), np.diff(indptr))
        n_src = np.asarray(g.out_csr[0]).shape[0] - 1
        twin = O.Graph(n_src, n_dst, src, dst)
        g._oracle_twin = twin
    return twin


def _abs_op(op):
    return "add" if op in ("add", "sub") else op


def term_scale(g, red, op, lt, rt, a, b, out_rows, maps, grad=None, want=None, gout=None):
    """sum over the reduced terms of |term| for every output element: the magnitude north_star's
    1e-5 relative tolerance for sums is measured against (instead of max(|want|, 1), an absolute
    1e-5 floor). Evaluated by the oracle on |inputs| with the sum reducer (max/min gradients:
    an upper bound, ties only drop terms). None for prod (kept relative to |want|)."""
    from oracle import oracle as O
    if red == "prod":
        return None
    og = _oracle_twin(g)
    r = "none" if red == "none" else "sum"
    a_ = np.abs(a)
    b_ = None if b is None else np.abs(b)
    if grad is None:
        if b is None:
            return np.abs(O.copy_reduce(r, og, lt, a_, out_rows, *maps))
        return np.abs(O.binary_op_reduce(r, _abs_op(op), og, lt, rt, a_, b_, out_rows, *maps))
    w_, g_ = np.abs(want), np.abs(gout)
    if b is None:
        return np.abs(O.backward_copy_reduce(r, og, lt, a_, w_, g_, *maps))
    fn = O.backward_lhs_binary_op_reduce if grad == "lhs" else O.backward_rhs_binary_op_reduce
    return np.abs(fn(r, _abs_op(op), og, lt, rt, a_, b_, w_, g_, *maps))


def check(name, got, want, red, scale=None, grad=False):
    got = np.asarray(got)
    assert got.shape == want.shape, "%s: shape %s vs %s" % (name, got.shape, want.shape)
    rtol = RTOL[red]
    if grad and red in ("max", "min", "none"):
        # gradients of max/min/none still sum contributions (into src/dst rows)
        rtol = 1e-5
    if grad and red == "prod":
        rtol = 1e-3
    if red == "none" and not grad and "dot" in name:
        # a dot product is itself a sum; the CUDA kernel adds its terms in a fixed butterfly order
        rtol = 1e-5
    if rtol == 0.0:
        ok = np.array_equal(got, want)
        assert ok, "%s: not bit-exact, max abs diff %g" % (
            name, np.abs(got.astype(np.float64) - want).max())
        return
    if scale is None:      # prod: relative to the value itself (no additive term structure)
        tol = rtol * np.maximum(np.abs(want), 1.0) + 1e-6
    else:                  # 1e-5 of sum|terms| (+ a denormal-sized floor, not an absolute 1e-5)
        tol = rtol * np.maximum(scale, np.abs(want)) + 1e-30
    err = np.abs(got.astype(np.float64) - want.astype(np.float64))
    assert (err <= tol).all(), "%s: max err %g (tol %g)" % (name, err.max(), tol.max())


def replay_index(impl, gold, prefix, g):
    for side, csr in (("in", g.in_csr), ("out", g.out_csr)):
        for k, a in zip(("indptr", "indices", "eids"), csr):
            want = gold["%s%s_%s" % (prefix, side, k)]
            assert np.array_equal(np.asarray(a).astype(np.int64), want.astype(np.int64)), \
                "%s%s_%s differs" % (prefix, side, k)


def replay_kernel_graph(impl, subset=None):
    """All 450 binary cases + 10 copy cases on the test_kernel.py graph."""
    gold = load("kernel_graph.npz")
    n, src, dst = cases.kernel_graph()
    ne = src.shape[0]
    g = impl.make_graph(n, n, src, dst)
    replay_index(impl, gold, "kg/", g)
    count = 0
    for lhs, rhs, op, red, bc in cases.binary_cases():
        name = cases.case_name(lhs, rhs, op, red, bc)
        if subset is not None and not subset(lhs, rhs, op, red, bc):
            continue
        a, b = cases.binary_case_inputs(n, ne, lhs, rhs, op, bc)
        lt, rt = cases.TARGETS[lhs], cases.TARGETS[rhs]
        out_rows = ne if red == "none" else n
        want = gold["kg/%s/out" % name]
        no = (None, None, None)
        out = impl.binary_op_reduce(red, op, g, lt, rt, a, b, out_rows, None, None, None)
        check(name + "/out", out, want, red,
              scale=term_scale(g, red, op, lt, rt, a, b, out_rows, no))
        gout = cases.grad_out_for(want.shape)
        # feed the REFERENCE forward output to backward so max/min masks are comparable
        gl = impl.backward_lhs_binary_op_reduce(red, op, g, lt, rt, a, b, want, gout,
                                                None, None, None)
        check(name + "/gl", gl, gold["kg/%s/gl" % name], red, grad=True,
              scale=term_scale(g, red, op, lt, rt, a, b, out_rows, no, "lhs", want, gout))
        gr = impl.backward_rhs_binary_op_reduce(red, op, g, lt, rt, a, b, want, gout,
                                                None, None, None)
        check(name + "/gr", gr, gold["kg/%s/gr" % name], red, grad=True,
              scale=term_scale(g, red, op, lt, rt, a, b, out_rows, no, "rhs", want, gout))
        count += 1
    hu, hv, he = cases.generate_feature(n, ne)
    for tname, x in (("u", hu), ("e", he)):
        for red in cases.REDUCERS:
            name = "copy_%s__%s" % (tname, red)
            out_rows = ne if red == "none" else n
            want = gold["kg/%s/out" % name]
            tg, no2 = cases.TARGETS[tname], (None, None)
            out = impl.copy_reduce(red, g, tg, x, out_rows, None, None)
            check(name + "/out", out, want, red,
                  scale=term_scale(g, red, None, tg, None, x, None, out_rows, no2))
            gout = cases.grad_out_for(want.shape)
            gi = impl.backward_copy_reduce(red, g, tg, x, want, gout, None, None)
            check(name + "/gin", gi, gold["kg/%s/gin" % name], red, grad=True,
                  scale=term_scale(g, red, None, tg, None, x, None, out_rows, no2, "lhs", want,
                                   gout))
            count += 1
    return count


def _bin(impl, gold, key, red, op, g, lt, rt, a, b, out_rows, maps=(None,) * 3,
         bmaps=(None,) * 3):
    want = gold[key + "/out"]
    out = impl.binary_op_reduce(red, op, g, lt, rt, a, b, out_rows, *maps)
    check(key + "/out", out, want, red,
          scale=term_scale(g, red, op, lt, rt, a, b, out_rows, maps))
    gout = cases.grad_out_for(want.shape)
    gl = impl.backward_lhs_binary_op_reduce(red, op, g, lt, rt, a, b, want, gout, *bmaps)
    check(key + "/gl", gl, gold[key + "/gl"], red, grad=True,
          scale=term_scale(g, red, op, lt, rt, a, b, out_rows, bmaps, "lhs", want, gout))
    gr = impl.backward_rhs_binary_op_reduce(red, op, g, lt, rt, a, b, want, gout, *bmaps)
    check(key + "/gr", gr, gold[key + "/gr"], red, grad=True,
          scale=term_scale(g, red, op, lt, rt, a, b, out_rows, bmaps, "rhs", want, gout))


def _copy(impl, gold, key, red, g, tgt, x, out_rows, maps=(None,) * 2, bmaps=(None,) * 2):
    want = gold[key + "/out"]
    out = impl.copy_reduce(red, g, tgt, x, out_rows, *maps)
    check(key + "/out", out, want, red,
          scale=term_scale(g, red, None, tgt, None, x, None, out_rows, maps))
    gout = cases.grad_out_for(want.shape)
    gi = impl.backward_copy_reduce(red, g, tgt, x, want, gout, *bmaps)
    check(key + "/gin", gi, gold[key + "/gin"], red, grad=True,
          scale=term_scale(g, red, None, tgt, None, x, None, out_rows, bmaps, "lhs", want, gout))


def replay_mid(impl):
    """Same draw order as make_golden.gen_mid (one RandomState per graph)."""
    gold = load("mid.npz")
    for name, ns, nd_, ne, alpha, seed in cases.MID_GRAPHS:
        src, dst = cases.powerlaw_graph(ns, nd_, ne, alpha, seed)
        g = impl.make_graph(ns, nd_, src, dst)
        p = "mid/%s/" % name
        replay_index(impl, gold, p, g)
        rs = np.random.RandomState(seed)
        x = rs.uniform(-1, 1, (ns, 33)).astype(np.float32)
        m = rs.uniform(-1, 1, (ne, 10)).astype(np.float32)
        for red in ("sum", "max", "min"):
            _copy(impl, gold, p + "copy_u_" + red, red, g, 0, x, nd_)
            _copy(impl, gold, p + "copy_e_" + red, red, g, 2, m, nd_)
        ft = rs.uniform(-1, 1, (ns, 4, 8)).astype(np.float32)
        a = rs.uniform(0, 1, (ne, 4, 1)).astype(np.float32)
        _bin(impl, gold, p + "u_mul_e_sum", "sum", "mul", g, 0, 2, ft, a, nd_)
        x32 = rs.uniform(-1, 1, (ns, 32)).astype(np.float32)
        w = rs.uniform(0, 1, (ne, 1)).astype(np.float32)
        _bin(impl, gold, p + "u_mul_w_sum", "sum", "mul", g, 0, 2, x32, w, nd_)
        el = rs.uniform(-1, 1, (ns, 4, 1)).astype(np.float32)
        er = rs.uniform(-1, 1, (nd_, 4, 1)).astype(np.float32)
        _bin(impl, gold, p + "u_add_v", "none", "add", g, 0, 1, el, er, ne)
        ee = rs.uniform(-1, 1, (ne, 4, 1)).astype(np.float32)
        for op in ("sub", "div", "mul"):
            vv = (er + 3) / 2 if op == "div" else er
            _bin(impl, gold, p + "e_%s_v" % op, "none", op, g, 2, 1, ee, vv, ne)
        if ns == nd_:
            ft2 = rs.uniform(-1, 1, (nd_, 4, 8)).astype(np.float32)
            _bin(impl, gold, p + "u_dot_v", "none", "dot", g, 0, 1, ft, ft2, ne)


def replay_mapped(impl):
    gold = load("mapped.npz")
    name, ns, nd_, ne, alpha, seed = cases.MID_GRAPHS[1]
    src, dst = cases.powerlaw_graph(ns, nd_, ne, alpha, seed)
    g = impl.make_graph(ns, nd_, src, dst)
    rs = np.random.RandomState(99)
    big_n, big_e = 500, 7000
    src_map = rs.permutation(big_n)[:ns]
    dst_map = rs.permutation(big_n)[:nd_]
    eid_user = rs.permutation(big_e)[:ne]
    assert np.array_equal(src_map, gold["map/src_map"])
    fwd_emap = eid_user[np.asarray(g.out_csr[2]).astype(np.int64)]
    bwd_emap = eid_user[np.asarray(g.in_csr[2]).astype(np.int64)]
    xu = rs.uniform(-1, 1, (big_n, 6)).astype(np.float32)
    xe = rs.uniform(0.5, 1.5, (big_e, 6)).astype(np.float32)
    for red in ("sum", "max", "prod"):
        _bin(impl, gold, "map/u_mul_e_" + red, red, "mul", g, 0, 2, xu, xe, big_n,
             maps=(src_map, fwd_emap, dst_map), bmaps=(src_map, bwd_emap, dst_map))
    _bin(impl, gold, "map/v_sub_u_none", "none", "sub", g, 1, 0, xu, xu, big_e,
         maps=(dst_map, src_map, fwd_emap), bmaps=(dst_map, src_map, bwd_emap))
    _copy(impl, gold, "map/copy_e_sum", "sum", g, 2, xe, big_n,
          maps=(fwd_emap, dst_map), bmaps=(bwd_emap, dst_map))

"""Generate the committed golden fixtures by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference and oracle/_ref/libdgl.so):

    make -C oracle/ref_build -j8
    OMP_NUM_THREADS=1 DGLBACKEND=pytorch DGL_LIBRARY_PATH=oracle/_ref \
        PYTHONPATH=/root/reference/python python tests/golden/make_golden.py

Every array written here is an output of the reference's own C++ (`kernel._CAPI_DGLKernel*`
through python/dgl/kernel.py, `_CAPI_DGLGraphGetAdj` / `_CAPI_DGLHeteroGetAdj` for the index
arrays). Inputs are NOT stored: tests regenerate them from tests/cases.py seeds.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402

import dgl  # noqa: E402
import dgl.ndarray as nd  # noqa: E402
from dgl import kernel as K  # noqa: E402
from dgl import graph_index, heterograph_index, utils  # noqa: E402

CPU = nd.cpu()
OPS_FEAT = {"dot": lambda s: s[:-1]}


def to_nd(a):
    return nd.array(np.ascontiguousarray(a), ctx=CPU)


def _adj(rst):
    """GetAdj returns a list-like PackedFunc: rst(i) = indptr, indices, edge ids."""
    return [rst(i).asnumpy() for i in range(3)]


def ref_graph(n_src, n_dst, src, dst):
    """Reference graph index object accepted by the kernels + its CSR arrays."""
    if n_src == n_dst:
        gi = graph_index.from_coo(n_src, src, dst, True)
        g = gi.get_immutable_gidx(CPU)
        in_csr = _adj(graph_index._CAPI_DGLGraphGetAdj(g, False, "csr"))
        out_csr = _adj(graph_index._CAPI_DGLGraphGetAdj(g, True, "csr"))
    else:
        hg = heterograph_index.create_unitgraph_from_coo(
            2, n_src, n_dst, utils.toindex(src), utils.toindex(dst), "any")
        g = hg.get_unitgraph(0, CPU)
        in_csr = _adj(heterograph_index._CAPI_DGLHeteroGetAdj(g, 0, False, "csr"))
        out_csr = _adj(heterograph_index._CAPI_DGLHeteroGetAdj(g, 0, True, "csr"))
    return g, in_csr, out_csr


def maybe(a, bits=32):
    if a is None:
        return None
    return to_nd(a.astype(np.int32 if bits == 32 else np.int64))


def run_binary(g, red, op, lt, rt, lhs, rhs, out_rows, gout_seed=7, maps=(None, None, None),
               bmaps=(None, None, None)):
    """fwd + both unreduced grads through the reference kernels."""
    feat = tuple(K.infer_binary_feature_shape(op, to_nd(lhs), to_nd(rhs)))
    oshape = (out_rows,) + (feat[:-1] if op == "dot" else feat)
    out = nd.empty(oshape, "float32", CPU)
    K.binary_op_reduce(red, op, g, lt, rt, to_nd(lhs), to_nd(rhs), out,
                       maybe(maps[0]), maybe(maps[1]), maybe(maps[2]))
    out_np = out.asnumpy()
    gout = cases.grad_out_for(oshape, gout_seed)
    gl = nd.empty((lhs.shape[0],) + feat, "float32", CPU)
    gr = nd.empty((rhs.shape[0],) + feat, "float32", CPU)
    K.backward_lhs_binary_op_reduce(red, op, g, lt, rt, to_nd(lhs), to_nd(rhs), to_nd(out_np),
                                    to_nd(gout), gl, maybe(bmaps[0]), maybe(bmaps[1]),
                                    maybe(bmaps[2]))
    K.backward_rhs_binary_op_reduce(red, op, g, lt, rt, to_nd(lhs), to_nd(rhs), to_nd(out_np),
                                    to_nd(gout), gr, maybe(bmaps[0]), maybe(bmaps[1]),
                                    maybe(bmaps[2]))
    return out_np, gl.asnumpy(), gr.asnumpy()


def run_copy(g, red, tgt, x, out_rows, gout_seed=7, maps=(None, None), bmaps=(None, None)):
    out = nd.empty((out_rows,) + x.shape[1:], "float32", CPU)
    K.copy_reduce(red, g, tgt, to_nd(x), out, maybe(maps[0]), maybe(maps[1]))
    out_np = out.asnumpy()
    gout = cases.grad_out_for(out_np.shape, gout_seed)
    gin = nd.empty(x.shape, "float32", CPU)
    K.backward_copy_reduce(red, g, tgt, to_nd(x), to_nd(out_np), to_nd(gout), gin,
                           maybe(bmaps[0]), maybe(bmaps[1]))
    return out_np, gin.asnumpy()


def gen_kernel_graph(store):
    n, src, dst = cases.kernel_graph()
    g, in_csr, out_csr = ref_graph(n, n, src, dst)
    for k, a in zip(("indptr", "indices", "eids"), in_csr):
        store["kg/in_" + k] = a
    for k, a in zip(("indptr", "indices", "eids"), out_csr):
        store["kg/out_" + k] = a
    ne = src.shape[0]
    for lhs, rhs, op, red, bc in cases.binary_cases():
        a, b = cases.binary_case_inputs(n, ne, lhs, rhs, op, bc)
        out_rows = ne if red == "none" else n
        o, gl, gr = run_binary(g, red, op, cases.TARGETS[lhs], cases.TARGETS[rhs], a, b,
                               out_rows)
        name = "kg/" + cases.case_name(lhs, rhs, op, red, bc)
        store[name + "/out"], store[name + "/gl"], store[name + "/gr"] = o, gl, gr
    hu, hv, he = cases.generate_feature(n, ne)
    for tname, x in (("u", hu), ("e", he)):
        for red in cases.REDUCERS:
            out_rows = ne if red == "none" else n
            o, gi = run_copy(g, red, cases.TARGETS[tname], x, out_rows)
            store["kg/copy_%s__%s/out" % (tname, red)] = o
            store["kg/copy_%s__%s/gin" % (tname, red)] = gi


def gen_mid(store):
    for name, ns, nd_, ne, alpha, seed in cases.MID_GRAPHS:
        src, dst = cases.powerlaw_graph(ns, nd_, ne, alpha, seed)
        g, in_csr, out_csr = ref_graph(ns, nd_, src, dst)
        for k, a in zip(("indptr", "indices", "eids"), in_csr):
            store["mid/%s/in_%s" % (name, k)] = a
        for k, a in zip(("indptr", "indices", "eids"), out_csr):
            store["mid/%s/out_%s" % (name, k)] = a
        rs = np.random.RandomState(seed)
        x = rs.uniform(-1, 1, (ns, 33)).astype(np.float32)
        m = rs.uniform(-1, 1, (ne, 10)).astype(np.float32)
        for red in ("sum", "max", "min"):
            o, gi = run_copy(g, red, 0, x, nd_)
            store["mid/%s/copy_u_%s/out" % (name, red)] = o
            store["mid/%s/copy_u_%s/gin" % (name, red)] = gi
            o, gi = run_copy(g, red, 2, m, nd_)
            store["mid/%s/copy_e_%s/out" % (name, red)] = o
            store["mid/%s/copy_e_%s/gin" % (name, red)] = gi
        # GAT-shaped u_mul_e + sum: [N,4,8] x [E,4,1]
        ft = rs.uniform(-1, 1, (ns, 4, 8)).astype(np.float32)
        a = rs.uniform(0, 1, (ne, 4, 1)).astype(np.float32)
        o, gl, gr = run_binary(g, "sum", "mul", 0, 2, ft, a, nd_)
        store["mid/%s/u_mul_e_sum/out" % name] = o
        store["mid/%s/u_mul_e_sum/gl" % name] = gl
        store["mid/%s/u_mul_e_sum/gr" % name] = gr
        # scalar edge weight [E,1] against [N,32]
        x32 = rs.uniform(-1, 1, (ns, 32)).astype(np.float32)
        w = rs.uniform(0, 1, (ne, 1)).astype(np.float32)
        o, gl, gr = run_binary(g, "sum", "mul", 0, 2, x32, w, nd_)
        store["mid/%s/u_mul_w_sum/out" % name] = o
        store["mid/%s/u_mul_w_sum/gl" % name] = gl
        store["mid/%s/u_mul_w_sum/gr" % name] = gr
        # GAT SDDMM u_add_v and edge-softmax pieces on [.,4,1]
        el = rs.uniform(-1, 1, (ns, 4, 1)).astype(np.float32)
        er = rs.uniform(-1, 1, (nd_, 4, 1)).astype(np.float32)
        o, gl, gr = run_binary(g, "none", "add", 0, 1, el, er, ne)
        store["mid/%s/u_add_v/out" % name] = o
        store["mid/%s/u_add_v/gl" % name] = gl
        store["mid/%s/u_add_v/gr" % name] = gr
        ee = rs.uniform(-1, 1, (ne, 4, 1)).astype(np.float32)
        for op in ("sub", "div", "mul"):
            vv = (er + 3) / 2 if op == "div" else er
            o, gl, gr = run_binary(g, "none", op, 2, 1, ee, vv, ne)
            store["mid/%s/e_%s_v/out" % (name, op)] = o
            store["mid/%s/e_%s_v/gl" % (name, op)] = gl
            store["mid/%s/e_%s_v/gr" % (name, op)] = gr
        # u_dot_v SDDMM [N,4,8].[N,4,8] -> [E,4]
        if ns == nd_:
            ft2 = rs.uniform(-1, 1, (nd_, 4, 8)).astype(np.float32)
            o, gl, gr = run_binary(g, "none", "dot", 0, 1, ft, ft2, ne)
            store["mid/%s/u_dot_v/out" % name] = o
            store["mid/%s/u_dot_v/gl" % name] = gl
            store["mid/%s/u_dot_v/gr" % name] = gr


def gen_mapped(store):
    """Id-mapping (pull / send_and_recv style) cases on the bipartite mid graph: operands live
    in larger frames and are addressed through lhs/rhs/out mappings; edge-target maps are
    indexed by CSR position exactly as spmv.build_gidx_and_mapping_uv builds them
    (python/dgl/runtime/spmv.py:146-186): forward = eid[out_csr.data], backward = eid[in_csr.data]."""
    name, ns, nd_, ne, alpha, seed = cases.MID_GRAPHS[1]
    src, dst = cases.powerlaw_graph(ns, nd_, ne, alpha, seed)
    g, in_csr, out_csr = ref_graph(ns, nd_, src, dst)
    rs = np.random.RandomState(99)
    big_n, big_e = 500, 7000
    src_map = rs.permutation(big_n)[:ns]
    dst_map = rs.permutation(big_n)[:nd_]
    eid_user = rs.permutation(big_e)[:ne]
    fwd_emap = eid_user[out_csr[2]]
    bwd_emap = eid_user[in_csr[2]]
    for k, a in (("src_map", src_map), ("dst_map", dst_map), ("eid_user", eid_user)):
        store["map/" + k] = a.astype(np.int64)
    xu = rs.uniform(-1, 1, (big_n, 6)).astype(np.float32)
    xe = rs.uniform(0.5, 1.5, (big_e, 6)).astype(np.float32)
    for red in ("sum", "max", "prod"):
        o, gl, gr = run_binary(g, red, "mul", 0, 2, xu, xe, big_n,
                               maps=(src_map, fwd_emap, dst_map),
                               bmaps=(src_map, bwd_emap, dst_map))
        store["map/u_mul_e_%s/out" % red] = o
        store["map/u_mul_e_%s/gl" % red] = gl
        store["map/u_mul_e_%s/gr" % red] = gr
    o, gl, gr = run_binary(g, "none", "sub", 1, 0, xu, xu, big_e,
                           maps=(dst_map, src_map, fwd_emap), bmaps=(dst_map, src_map, bwd_emap))
    store["map/v_sub_u_none/out"] = o
    store["map/v_sub_u_none/gl"] = gl
    store["map/v_sub_u_none/gr"] = gr
    o, gi = run_copy(g, "sum", 2, xe, big_n, maps=(fwd_emap, dst_map), bmaps=(bwd_emap, dst_map))
    store["map/copy_e_sum/out"] = o
    store["map/copy_e_sum/gin"] = gi


def main():
    assert os.environ.get("OMP_NUM_THREADS") == "1", "run with OMP_NUM_THREADS=1 (deterministic)"
    for fn_, gen in (("kernel_graph.npz", gen_kernel_graph), ("mid.npz", gen_mid),
                     ("mapped.npz", gen_mapped)):
        store = {}
        gen(store)
        np.savez_compressed(os.path.join(HERE, fn_), **store)
        print(fn_, len(store), "arrays,",
              os.path.getsize(os.path.join(HERE, fn_)) // 1024, "KiB")


if __name__ == "__main__":
    main()

"""Stage-by-stage timing of the builtin-composed GATConv step on the arxiv-shaped graph (C3):
which kernel calls the module's 3 ms go to. CUDA events, each stage alone."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import dgl_hack_b200 as B  # noqa: E402
from dgl_hack_b200 import backend as F, kernel as K  # noqa: E402
import dgl_hack_b200.nn as nnb  # noqa: E402

DEV = torch.device("cuda:0")


def timeit(fn, reps=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    n, e, H, Fw = 169343, 1166243, 8, 16
    src, dst = bench.make_graph_coo(n, e, 0.4, 3, DEV)
    g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
    r = lambda *s: torch.rand(*s, device=DEV) * 2 - 1  # noqa: E731
    ft, el, er, a, ge, go = r(n, H, Fw), r(n, H, 1), r(n, H, 1), torch.rand(e, H, 1, device=DEV), r(e, H, 1), r(n, H, Fw)
    out = {}

    def rec(name, fn):
        ms = timeit(fn)
        out[name] = {"ms": round(ms, 4), "kernel": K.last_kernel()}
        print(name, out[name], flush=True)
    rec("fwd u_add_v", lambda: F.binary_reduce("none", "add", g, 0, 1, el, er, e))
    rec("fwd edge_softmax fused", lambda: nnb.edge_softmax(g, ge))
    rec("fwd u_mul_e+sum", lambda: F.binary_reduce("sum", "mul", g, 0, 2, ft, a, n))
    o = torch.empty(n, H, Fw, device=DEV)
    gl = torch.empty(n, H, Fw, device=DEV)
    rec("bwd u_mul_e wrt ft (transposed spmm)",
        lambda: K.backward_lhs_binary_op_reduce("sum", "mul", g, 0, 2, ft, a, o, go, gl))
    ga = torch.empty(e, H, device=DEV)
    rec("bwd u_mul_e wrt a (dot per edge,head)",
        lambda: K.binary_op_reduce("none", "dot", g, 0, 1, ft, go, ga))
    sm = nnb.edge_softmax(g, ge)
    gz = torch.empty_like(sm)
    rec("bwd edge_softmax fused", lambda: K.edge_softmax_backward(g, sm, ge, gz))
    eo = torch.empty(e, H, 1, device=DEV)
    gel = torch.empty(n, H, 1, device=DEV)
    rec("bwd u_add_v wrt el (src)",
        lambda: K.backward_lhs_binary_op_reduce("none", "add", g, 0, 1, el, er, eo, ge, gel))
    rec("bwd u_add_v wrt er (dst)",
        lambda: K.backward_rhs_binary_op_reduce("none", "add", g, 0, 1, el, er, eo, ge, gel))
    x = r(n, 128)
    lin = torch.nn.Linear(128, H * Fw, bias=False).to(DEV)
    gout = r(n, H * Fw)

    def fc():
        lin(x).backward(gout)
        lin.weight.grad = None
    rec("torch fc 128->128 fwd+bwd", fc)
    for cls, name in ((nnb.GATConv, "GATConv step (builtin composition)"),
                      (nnb.FusedGATConv, "GATConv step (fused kernel)")):
        conv = cls(128, Fw, H).to(DEV)

        def step():
            conv(Gr, x).backward(go)
            for p in conv.parameters():
                p.grad = None
        from dgl_hack_b200.graph import Graph
        Gr = Graph(src, dst, num_nodes=n, device=DEV, gidx=g)
        rec(name, step)
    path = os.path.join(ROOT, "gpurun_out", "r1_gat_conv_probe.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    json.dump(out, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()

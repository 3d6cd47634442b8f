"""Runs the REFERENCE's own Python stack (oracle/_ref/python/dgl: DGLGraph.update_all,
apply_edges, send_and_recv, pull, edge_softmax, dgl.nn.pytorch.GraphConv / SAGEConv / GATConv /
RelGraphConv) with the B200 PackedFunc shims registered into the reference's libdgl, once with CPU
tensors (the shims forward to the reference's CPU kernels) and once with tensors on --device
(cuda: the sm_100a kernels), and prints one JSON object {case: max relative error} (+ shim stats).
Test infrastructure: started as a subprocess by tests/test_packed_abi.py (device cpu) and
tests/test_gpu_dgl_patch.py (device cuda) with DGLBACKEND / DGL_LIBRARY_PATH / PYTHONPATH set.
"""
import argparse
import contextlib
import ctypes
import io
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel_err(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    if a.shape != b.shape:
        return float("inf")
    if a.numel() == 0:
        return 0.0
    fin = torch.isfinite(b) & (b.abs() < 1e30)   # reducer identities (+-FLT_MAX) must match exactly
    if not torch.equal(a[~fin], b[~fin]):
        return float("inf")
    return float(((a[fin] - b[fin]).abs().max() / b[fin].abs().max().clamp(min=1e-12)).item()) \
        if fin.any() else 0.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--register-first", type=int, default=1,
                    help="1: register the shims before `import dgl`; 0: install() afterwards")
    args = ap.parse_args()
    dev = torch.device(args.device)
    if not torch.cuda.is_available():
        # the fork's GATConv.forward calls th.cuda.synchronize() around its timing prints
        # (nn/pytorch/conv/gatconv.py:146): not part of the arithmetic, meaningless without a GPU
        torch.cuda.synchronize = lambda *a, **k: None
    from dgl_hack_b200 import dgl_patch
    from dgl_hack_b200._lib import LIB
    if args.register_first:
        dgl_patch.register()
    with contextlib.redirect_stdout(io.StringIO()):
        import dgl
        import dgl.function as fn
        import dgl.nn.pytorch as dglnn
        from dgl.nn.pytorch.softmax import edge_softmax
    names = dgl_patch.install()

    rs = np.random.RandomState(7)
    n, e = 300, 4000
    src = rs.randint(0, n, e)
    dst = rs.randint(0, n, e)
    dst[:400] = 5                      # a hub; node n-1 gets no in-edge
    dst[dst == n - 1] = 0
    etypes = rs.randint(0, 4, e)

    def graph(self_loops=False):
        g = dgl.DGLGraph()
        g.add_nodes(n)
        g.add_edges(src, dst)
        if self_loops:      # modules: no zero-in-degree node (max-pool of nothing is -FLT_MAX)
            g.add_edges(np.arange(n), np.arange(n))
        return g

    res = {}

    def both(name, fn_run):
        """fn_run(device) -> list of tensors (outputs and gradients)."""
        try:
            with contextlib.redirect_stdout(io.StringIO()):   # the fork prints timings
                want = fn_run(torch.device("cpu"))
                got = fn_run(dev)
            res[name] = max(rel_err(g_, w_) for g_, w_ in zip(got, want))
        except Exception as ex:  # reported per case
            res[name] = "error: %s: %s" % (type(ex).__name__, str(ex)[:300])

    def T(arr, d, grad=True):
        return torch.tensor(arr, dtype=torch.float32, device=d, requires_grad=grad)

    x_np = rs.uniform(-1, 1, (n, 6)).astype(np.float32)
    w_np = rs.uniform(0.1, 1, (e, 1)).astype(np.float32)
    h_np = rs.uniform(-1, 1, (n, 4, 3)).astype(np.float32)
    a_np = rs.uniform(-1, 1, (e, 4, 1)).astype(np.float32)

    for red in ("sum", "max", "mean"):
        def run(d, red=red):
            g = graph()
            x = T(x_np, d)
            g.ndata["h"] = x
            g.update_all(fn.copy_src("h", "m"), getattr(fn, red)("m", "o"))
            o = g.ndata["o"]
            (o[:n - 1] * o[:n - 1]).sum().backward()
            return [o, x.grad]
        both("update_all copy_src+%s" % red, run)

    def run_umule(d):
        g = graph()
        x, w = T(h_np, d), T(a_np, d)
        g.ndata["h"], g.edata["a"] = x, w
        g.update_all(fn.u_mul_e("h", "a", "m"), fn.sum("m", "o"))
        o = g.ndata["o"]
        (o * o).sum().backward()
        return [o, x.grad, w.grad]
    both("update_all u_mul_e+sum (broadcast [E,H,1])", run_umule)

    def run_apply(d):
        g = graph()
        x = T(x_np, d)
        g.ndata["h"] = x
        g.apply_edges(fn.u_add_v("h", "h", "s"))
        g.apply_edges(fn.u_dot_v("h", "h", "t"))
        s, t = g.edata["s"], g.edata["t"]
        ((s * s).sum() + (t * t).sum()).backward()
        return [s, t, x.grad]
    both("apply_edges u_add_v / u_dot_v", run_apply)

    def run_softmax(d):
        g = graph()
        z = T(a_np * 3, d)
        a = edge_softmax(g, z)
        (a * T(a_np, d, False)).sum().backward()
        return [a, z.grad]
    both("edge_softmax", run_softmax)

    def run_partial(d):
        g = graph()
        x = T(x_np, d)
        g.ndata["h"] = x
        g.ndata["o"] = torch.zeros(n, 6, device=d)
        g.send_and_recv(np.arange(0, e, 3), fn.copy_src("h", "m"), fn.sum("m", "o"))
        o1 = g.ndata["o"]
        g.pull(np.arange(0, n, 7), fn.copy_src("h", "m"), fn.max("m", "p"))
        o2 = g.ndata["p"]
        ((o1 * o1).sum() + (o2[np.arange(0, n, 7)] ** 2).sum()).backward()
        return [o1, o2[np.arange(0, n, 7)], x.grad]
    both("send_and_recv / pull (id-mapped kernels)", run_partial)

    def run_bipartite(d):
        g = dgl.bipartite((src % 50, dst), card=(50, n))
        x = T(x_np[:50], d)
        g.srcdata["h"] = x
        g.update_all(fn.copy_src("h", "m"), fn.sum("m", "o"))
        o = g.dstdata["o"]
        (o * o).sum().backward()
        return [o, x.grad]
    both("bipartite unit graph update_all", run_bipartite)

    def module_case(name, make, feat_np, call=None):
        def run(d):
            torch.manual_seed(3)
            m = make().to(d)
            g = graph(self_loops=True)
            x = T(feat_np, d)
            out = call(m, g, x, d) if call else m(g, x)
            (out * out).sum().backward()
            return [out, x.grad] + [p.grad for p in m.parameters() if p.grad is not None]
        both(name, run)

    feat16 = rs.uniform(-1, 1, (n, 16)).astype(np.float32)
    module_case("GraphConv(16, 8)", lambda: dglnn.GraphConv(16, 8), feat16)
    module_case("GraphConv(16, 32)", lambda: dglnn.GraphConv(16, 32), feat16)
    module_case("SAGEConv mean", lambda: dglnn.SAGEConv(16, 8, "mean"), feat16)
    module_case("SAGEConv pool", lambda: dglnn.SAGEConv(16, 8, "pool"), feat16)
    module_case("SAGEConv gcn", lambda: dglnn.SAGEConv(16, 8, "gcn"), feat16)
    module_case("GATConv 4 heads", lambda: dglnn.GATConv(16, 8, 4), feat16)
    module_case("RelGraphConv basis", lambda: dglnn.RelGraphConv(16, 8, 4, "basis", 2), feat16,
                call=lambda m, g, x, d: m(g, x, torch.tensor(np.concatenate([etypes, np.zeros(n, np.int64)]), device=d)))

    stats = (ctypes.c_int64 * 3)()
    LIB.DGLB200PackedStats(stats)
    from dgl_hack_b200 import kernel as K
    print(json.dumps({"device": str(dev), "registered": names, "cases": res,
                      "shim_calls": int(stats[0]), "forwarded_to_reference_cpu": int(stats[1]),
                      "graphs_built": int(stats[2]), "b200_launches": K.launch_count()}))


if __name__ == "__main__":
    main()

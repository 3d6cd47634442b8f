"""Per-config kernel timings for BASELINE.json configs 1, 3, 4, 5 (config 2 is bench.py).

    python profiles/bench_configs.py [--out gpurun_out/r1_configs.json] [--only c3,c5]

Every op goes through the public API (dgl_hack_b200.backend / kernel); time = CUDA events around
`reps` back-to-back calls after warm-up, on the current stream. Algorithmic bytes follow
SURVEY.md 8(d). Synthetic graphs are generated on the device with fixed seeds.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import dgl_hack_b200 as B  # noqa: E402
from dgl_hack_b200 import backend as F  # noqa: E402
from dgl_hack_b200 import kernel as K  # noqa: E402

DEV = torch.device("cuda:0")
PEAK, _ = bench.measured_peak()
L2 = None


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


def gather_bytes(n_src, e, d):
    s = n_src * d * 4
    return e * d * 4 if s > L2 else s


def rec(out, cfg, op, ms, e, nbytes, note=""):
    row = {"config": cfg, "op": op, "ms": round(ms, 4), "gedges_per_s": round(e / ms / 1e6, 3),
           "algorithmic_gb": round(nbytes / 1e9, 4), "gb_per_s": round(nbytes / ms / 1e6, 1),
           "frac_of_measured_hbm": round(nbytes / ms / 1e6 / PEAK, 3), "kernel": K.last_kernel(),
           "note": note}
    out.append(row)
    print(json.dumps(row), flush=True)


def rand(*shape):
    return torch.rand(*shape, device=DEV) * 2 - 1


def c1(out):
    n, e = 2708, 10556
    gen = torch.Generator(device=DEV)
    gen.manual_seed(1)
    src = torch.randint(0, n, (e,), generator=gen, device=DEV)
    dst = torch.randint(0, n, (e,), generator=gen, device=DEV)
    g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
    for d in (16, 7):
        x = rand(n, d)
        ms = timeit(lambda: F.copy_reduce("sum", g, 0, x, n), reps=200)
        rec(out, "C1 cora", "copy_src+sum D=%d" % d, ms, e,
            gather_bytes(n, e, d) + e * 4 + n * d * 4 + (n + 1) * 4, "launch/latency bound")


def c2(out):
    """Reddit-shaped operator sweep (bench.py times copy_src+mean fwd+bwd on the same graph)."""
    n, e, d = 232965, 114615892, 602
    src, dst = bench.make_graph_coo(n, e, 0.6, 2, DEV)
    g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
    del src, dst
    x = rand(n, d)
    gout = rand(n, d)
    w = torch.rand(e, 1, device=DEV)
    gb = gather_bytes(n, e, d)
    base = gb + e * 4 + n * d * 4 + (n + 1) * 4
    ms = timeit(lambda: F.copy_reduce("sum", g, 0, x, n), reps=5, warm=2)
    rec(out, "C2 reddit", "copy_src+sum fwd D=602", ms, e, base)
    ms = timeit(lambda: F.binary_reduce("sum", "mul", g, 0, 2, x, w, n), reps=5, warm=2)
    rec(out, "C2 reddit", "u_mul_e+sum fwd D=602, e [E,1]", ms, e, base + e * 8)
    out_t = F.binary_reduce("sum", "mul", g, 0, 2, x, w, n)
    gx = torch.empty_like(x)
    ms = timeit(lambda: K.backward_lhs_binary_op_reduce("sum", "mul", g, 0, 2, x, w, out_t, gout,
                                                        gx), reps=5, warm=2)
    rec(out, "C2 reddit", "u_mul_e+sum backward lhs (grad x)", ms, e, base + e * 8)
    # grad wrt the edge scalar: unreduced gradient is [E, 602]; the torch side sums it to [E,1].
    # Timed at 1/16 of the edges' worth of rows is not possible (kernel walks all edges), so the
    # full call is timed only when its [E,602] buffer fits (276 GB: it does not) -> skipped.
    ms = timeit(lambda: F.copy_reduce("max", g, 0, x, n), reps=3, warm=1)
    rec(out, "C2 reddit", "copy_src+max fwd D=602", ms, e, base)
    xs = rand(n, 1)
    es = rand(e, 1)
    ms = timeit(lambda: F.copy_reduce("sum", g, 2, es, n), reps=5, warm=2)
    rec(out, "C2 reddit", "copy_e+sum [E,1]", ms, e, e * 12 + n * 8)
    ms = timeit(lambda: F.binary_reduce("none", "add", g, 0, 1, xs, xs, e), reps=5, warm=2)
    rec(out, "C2 reddit", "u_add_v -> [E,1] (SDDMM)", ms, e, e * 12 + n * 8)
    x64 = rand(n, 64)
    ms = timeit(lambda: F.binary_reduce("none", "dot", g, 0, 1, x64, x64, e), reps=3, warm=1)
    rec(out, "C2 reddit", "u_dot_v D=64 -> [E,1] (SDDMM)", ms, e, e * 12 + 2 * n * 64 * 4)


def c3(out):
    n, e, H, Fw = 169343, 1166243, 8, 16
    src, dst = bench.make_graph_coo(n, e, 0.4, 3, DEV)
    g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
    ft, el, er = rand(n, H, Fw), rand(n, H, 1), rand(n, H, 1)
    ee = rand(e, H, 1)
    a = torch.rand(e, H, 1, device=DEV)
    gout = rand(n, H, Fw)
    ms = timeit(lambda: F.binary_reduce("none", "add", g, 0, 1, el, er, e))
    rec(out, "C3 arxiv", "u_add_v SDDMM [N,8,1]", ms, e, e * 8 + 2 * n * H * 4 + e * H * 4)

    def edge_softmax():
        # nn/pytorch/softmax.py:69-78 : 4 kernel calls + exp
        smax = F.copy_reduce("max", g, 2, ee, n)
        o = F.binary_reduce("none", "sub", g, 2, 1, ee, smax, e)
        o = torch.exp(o)
        ssum = F.copy_reduce("sum", g, 2, o, n)
        return F.binary_reduce("none", "div", g, 2, 1, o, ssum, e)
    ms = timeit(edge_softmax)
    rec(out, "C3 arxiv", "edge_softmax (reference 4-call form)", ms, e,
        4 * (e * 8 + e * H * 4 * 2 + n * H * 4), "4 kernels + torch.exp")
    ms = timeit(lambda: F.binary_reduce("sum", "mul", g, 0, 2, ft, a, n))
    rec(out, "C3 arxiv", "u_mul_e+sum [N,8,16]x[E,8,1]", ms, e,
        gather_bytes(n, e, H * Fw) + e * 8 + e * H * 4 + n * H * Fw * 4)
    ms = timeit(lambda: F.fused_gat(g, ft, el, er, 0.2))
    rec(out, "C3 arxiv", "fused GAT forward (u_add_v+leaky+exp+sum+div+u_mul_e)", ms, e,
        gather_bytes(n, e, H * Fw) + e * 8 + e * H * 4 + n * H * Fw * 4 + 3 * n * H * 4,
        "one kernel")
    ftg, elg, erg = [t.clone().requires_grad_(True) for t in (ft, el, er)]

    def gat_fb():
        o = F.fused_gat(g, ftg, elg, erg, 0.2)
        o.backward(gout)
        ftg.grad = elg.grad = erg.grad = None
    ms = timeit(gat_fb)
    rec(out, "C3 arxiv", "fused GAT fwd+bwd", ms, 2 * e,
        2 * (gather_bytes(n, e, H * Fw) * 2 + e * 8 + e * H * 8 + n * H * Fw * 4), "3 kernels")

    def gat_builtin_fb():
        # GATConv as the reference composes it from builtins (gatconv.py:122-170)
        ev = F.binary_reduce("none", "add", g, 0, 1, elg, erg, e)
        ev = torch.nn.functional.leaky_relu(ev, 0.2)
        smax = F.copy_reduce("max", g, 2, ev, n)
        ev = torch.exp(F.binary_reduce("none", "sub", g, 2, 1, ev, smax, e))
        ssum = F.copy_reduce("sum", g, 2, ev, n)
        att = F.binary_reduce("none", "div", g, 2, 1, ev, ssum, e)
        o = F.binary_reduce("sum", "mul", g, 0, 2, ftg, att, n)
        o.backward(gout)
        ftg.grad = elg.grad = erg.grad = None
    ms = timeit(gat_builtin_fb)
    rec(out, "C3 arxiv", "GAT attention from builtins fwd+bwd (edge_softmax with max)", ms, 2 * e,
        2 * (gather_bytes(n, e, H * Fw) * 2 + e * 8 + e * H * 8 + n * H * Fw * 4),
        "6 forward + 7 backward kernel calls + torch elementwise")


def c4(out):
    n, e, R, d = 1885136, 5668682, 133, 16
    gen = torch.Generator(device=DEV)
    gen.manual_seed(4)
    src = torch.randint(0, n, (e,), generator=gen, device=DEV)
    dst = torch.randint(0, n, (e,), generator=gen, device=DEV)
    w = (torch.arange(1, R + 1, device=DEV, dtype=torch.float64)) ** -1.1
    et = torch.multinomial(w, e, replacement=True, generator=gen)
    x = rand(n, d)
    graphs = []
    for r in range(R):
        m = et == r
        graphs.append(B.GraphIndex.from_coo(n, n, src[m], dst[m], DEV))

    def loop():
        acc = None
        for gr in graphs:
            o = F.copy_reduce("sum", gr, 0, x, n)
            acc = o if acc is None else acc + o
        return acc
    ms = timeit(loop, reps=3, warm=1)
    rec(out, "C4 AM", "133 x copy_src+sum D=16 (per-relation loop, HeteroGraphConv style)", ms, e,
        R * (n * d * 4 * 2) + e * 4, "one launch per relation + torch add")
    g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
    msg = rand(e, d)
    ms = timeit(lambda: F.copy_reduce("sum", g, 2, msg, n))
    rec(out, "C4 AM", "copy_e+sum [E,16] (RelGraphConv message reduce)", ms, e,
        e * d * 4 + e * 4 + n * d * 4 + (n + 1) * 4)
    # fused RGCN layer 1 (hack op): typed adjacency, W [133,16,16], norm [E,1]
    import time
    torch.cuda.synchronize()
    t0 = time.time()
    g.set_edge_types(et, R)
    torch.cuda.synchronize()
    prep_ms = (time.time() - t0) * 1e3
    w = rand(R, d, d)
    norm = torch.rand(e, 1, device=DEV)
    ms = timeit(lambda: F.rgcn_layer1(g, x, w, norm))
    # per edge: h row (gather) + 4 index/type/eid/norm words; per node: out row + indptr
    nb = gather_bytes(n, e, d) + e * 16 + n * d * 4 + (n + 1) * 4 + R * d * d * 4
    rec(out, "C4 AM", "rgcn_layer1 fwd (typed CSR, one kernel)", ms, e, nb,
        "typed adjacency build %.1f ms once" % prep_ms)
    xg = x.clone().requires_grad_(True)
    wg = w.clone().requires_grad_(True)
    gout = rand(n, d)

    def l1_fb():
        o = F.rgcn_layer1(g, xg, wg, norm)
        o.backward(gout)
        xg.grad = wg.grad = None
    ms = timeit(l1_fb)
    rec(out, "C4 AM", "rgcn_layer1 fwd+bwd (grad_h + grad_W, no atomics)", ms, 2 * e,
        2 * nb + e * (2 * d * 4 + 12), "5 kernels")

    def relgraphconv_lowmem():
        # relgraphconv.py:122-137 low-mem basis path: per-edge bmm message then copy_e+sum
        m = torch.bmm(x[src].unsqueeze(1), w[et]).squeeze(1) * norm
        return F.copy_reduce("sum", g, 2, m, n)
    ms = timeit(relgraphconv_lowmem, reps=5, warm=2)
    rec(out, "C4 AM", "RelGraphConv-style torch bmm message + copy_e+sum (comparison)", ms, e, nb,
        "torch gather+bmm then our copy_e kernel")


def graphed(out, cfg, op, step, e):
    """Same step captured once into a CUDA graph and replayed: the kernels' own time without the
    per-call Python / FFI / autograd host cost (the library allocates scratch stream-ordered and
    builds its plans on first use, so a warmed-up step is capturable)."""
    try:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                step()
        torch.cuda.current_stream().wait_stream(s)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            step()
        ms = timeit(gr.replay)
        rec(out, cfg, op, ms, e, 0, "torch.cuda.CUDAGraph replay")
    except Exception as ex:  # capture is an optimisation; report instead of failing the sweep
        torch.cuda.synchronize()
        print(json.dumps({"config": cfg, "op": op, "error": str(ex)[:300]}), flush=True)


def models(out):
    """The configs' models through dgl_hack_b200.nn (module forward + backward, CUDA events):
    C1 2-layer GCN (Cora-shaped), C3 GATConv builtin-composed (fused edge_softmax) vs FusedGATConv,
    edge_softmax fused vs 4-call, C4 RelGraphConv (basis, typed-adjacency kernel)."""
    import dgl_hack_b200.nn as nnb
    from dgl_hack_b200.graph import Graph
    # C1
    n, e = 2708, 10556
    gen = torch.Generator(device=DEV)
    gen.manual_seed(1)
    src = torch.cat([torch.randint(0, n, (e - n,), generator=gen, device=DEV),
                     torch.arange(n, device=DEV)])
    dst = torch.cat([torch.randint(0, n, (e - n,), generator=gen, device=DEV),
                     torch.arange(n, device=DEV)])
    g = Graph(src, dst, num_nodes=n, device=DEV)
    l1 = nnb.GraphConv(1433, 16, activation=torch.relu).to(DEV)
    l2 = nnb.GraphConv(16, 7).to(DEV)
    x = rand(n, 1433)
    y = torch.randint(0, 7, (n,), device=DEV)
    params = list(l1.parameters()) + list(l2.parameters())

    def gcn_step():
        loss = torch.nn.functional.cross_entropy(l2(g, l1(g, x)), y)
        loss.backward()
        for p_ in params:
            p_.grad = None
    ms = timeit(gcn_step)
    rec(out, "C1 cora", "2-layer GCN fwd+loss+bwd (nn.GraphConv via update_all)", ms, 4 * e, 0,
        "host/launch bound")
    graphed(out, "C1 cora", "2-layer GCN fwd+loss+bwd, CUDA-graph replay", gcn_step, 4 * e)
    # C3
    n, e, H, Fw = 169343, 1166243, 8, 16
    src, dst = bench.make_graph_coo(n, e, 0.4, 3, DEV)
    g = Graph(src, dst, num_nodes=n, device=DEV)
    z = rand(e, H, 1).requires_grad_(True)
    go = rand(e, H, 1)
    for form in ("fused", "builtin"):
        os.environ["B200_EDGE_SOFTMAX"] = form
        with torch.no_grad():
            ms = timeit(lambda: nnb.edge_softmax(g, z))
        rec(out, "C3 arxiv", "edge_softmax forward (%s)" % form, ms, e,
            2 * e * H * 4 + e * 8 + n * 4, "1 kernel" if form == "fused" else "4 kernels + exp")

        def sm_fb():
            nnb.edge_softmax(g, z).backward(go)
            z.grad = None
        ms = timeit(sm_fb)
        rec(out, "C3 arxiv", "edge_softmax fwd+bwd (%s)" % form, ms, 2 * e,
            5 * e * H * 4 + 2 * (e * 8 + n * 4))
    x = rand(n, 128)
    gout = rand(n, H, Fw)
    for form, cls in (("builtin composition, fused edge_softmax", nnb.GATConv),
                      ("fused GAT kernel", nnb.FusedGATConv)):
        os.environ["B200_EDGE_SOFTMAX"] = "fused"
        conv = cls(128, Fw, H).to(DEV)

        def gat_step():
            conv(g, x).backward(gout)
            for p_ in conv.parameters():
                p_.grad = None
        ms = timeit(gat_step)
        rec(out, "C3 arxiv", "GATConv 128->8x16 fwd+bwd (%s)" % form, ms, 2 * e, 0)
        graphed(out, "C3 arxiv", "GATConv 128->8x16 fwd+bwd (%s), CUDA-graph replay" % form,
                gat_step, 2 * e)
    os.environ["B200_EDGE_SOFTMAX"] = "builtin"
    conv = nnb.GATConv(128, Fw, H).to(DEV)
    ms = timeit(gat_step)
    rec(out, "C3 arxiv", "GATConv 128->8x16 fwd+bwd (builtin composition, 4-call edge_softmax)",
        ms, 2 * e, 0)
    os.environ.pop("B200_EDGE_SOFTMAX")
    # C4
    n, e, R, d = 1885136, 5668682, 133, 16
    gen.manual_seed(4)
    src = torch.randint(0, n, (e,), generator=gen, device=DEV)
    dst = torch.randint(0, n, (e,), generator=gen, device=DEV)
    w = (torch.arange(1, R + 1, device=DEV, dtype=torch.float64)) ** -1.1
    et = torch.multinomial(w, e, replacement=True, generator=gen)
    g = Graph(src, dst, num_nodes=n, device=DEV)
    conv = nnb.RelGraphConv(d, d, R, "basis", num_bases=R, self_loop=True).to(DEV)
    x = rand(n, d)
    norm = torch.rand(e, 1, device=DEV)
    gout = rand(n, d)

    def rgcn_step():
        conv(g, x, et, norm).backward(gout)
        for p_ in conv.parameters():
            p_.grad = None
    ms = timeit(rgcn_step)
    rec(out, "C4 AM", "RelGraphConv 16->16 R=133 fwd+bwd (typed-adjacency kernel)", ms, 2 * e, 0)
    graphed(out, "C4 AM", "RelGraphConv 16->16 R=133 fwd+bwd, CUDA-graph replay", rgcn_step, 2 * e)


def sbm_graph(n, e, blocks, p_in, alpha, seed):
    gen = torch.Generator(device=DEV)
    gen.manual_seed(seed)
    bsz = n // blocks
    sb = torch.randint(0, blocks, (e,), generator=gen, device=DEV)
    same = torch.rand(e, generator=gen, device=DEV) < p_in
    db = torch.where(same, sb, torch.randint(0, blocks, (e,), generator=gen, device=DEV))

    def within(_):
        u = torch.rand(e, generator=gen, device=DEV, dtype=torch.float64)
        r = ((bsz ** (1.0 - alpha) - 1.0) * u + 1.0) ** (1.0 / (1.0 - alpha)) - 1.0
        return r.floor().clamp_(0, bsz - 1).to(torch.int64)
    perm = torch.randperm(bsz, generator=gen, device=DEV)
    src = sb * bsz + perm[within(0)]
    dst = db * bsz + perm[within(1)]
    return src, dst


def c5(out):
    n, e = 2449029, 61859140
    src, dst = sbm_graph(n, e, 64, 0.85, 0.5, 5)
    g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
    del src, dst
    for d in (100, 256, 48, 47):
        x = rand(n, d).requires_grad_(True)
        gout = rand(n, d)
        ms = timeit(lambda: F.copy_reduce("sum", g, 0, x.detach(), n), reps=10, warm=3)
        nb = gather_bytes(n, e, d) + e * 4 + n * d * 4 + (n + 1) * 4
        rec(out, "C5 products", "copy_src+sum fwd D=%d" % d, ms, e, nb)

        def fb():
            o = F.copy_reduce("sum", g, 0, x, n)
            o.backward(gout)
            x.grad = None
        ms = timeit(fb, reps=10, warm=3)
        rec(out, "C5 products", "copy_src+sum fwd+bwd D=%d" % d, ms, 2 * e, 2 * nb)


def ref_cpu(out):
    """The unmodified reference CPU kernels (oracle/_ref/libdgl.so, all host cores) on the same
    shapes, one warm-up + one timed call each: the 'Ref CPU ms' column of BASELINE.md section 5."""
    import time
    import numpy as np
    sys.path.insert(0, ROOT)
    from oracle import ref_capi as R
    if not R.available():
        print(json.dumps({"ref_cpu": "oracle/_ref/libdgl.so not present"}))
        return
    cores = os.cpu_count()

    def t(fn):
        fn()
        t0 = time.perf_counter()
        fn()
        return (time.perf_counter() - t0) * 1e3

    def row(cfg, op, ms, e):
        r = {"config": cfg, "op": op, "ref_cpu_ms": round(ms, 2), "cores": cores,
             "gedges_per_s": round(e / ms / 1e6, 5)}
        out.append(r)
        print(json.dumps(r), flush=True)

    rs = np.random.RandomState(0)
    f32 = lambda *sh: rs.uniform(-1, 1, sh).astype(np.float32)  # noqa: E731
    # C1
    n, e = 2708, 10556
    g = R.RefGraph(n, rs.randint(0, n, e), rs.randint(0, n, e))
    g.materialize()
    X, O_ = R.RefArray(f32(n, 16)), R.RefArray(shape=(n, 16), dtype=np.float32)
    row("C1 cora", "copy_src+sum D=16", t(lambda: R.copy_reduce("sum", g, 0, X, O_)), e)
    # C3
    n, e = 169343, 1166243
    src, dst = bench.make_graph_coo(n, e, 0.4, 3, torch.device("cpu"))
    g = R.RefGraph(n, src.numpy(), dst.numpy())
    g.materialize()
    X, O_ = R.RefArray(f32(n, 128)), R.RefArray(shape=(n, 128), dtype=np.float32)
    row("C3 arxiv", "copy_src+sum D=128", t(lambda: R.copy_reduce("sum", g, 0, X, O_)), e)
    FT, A = R.RefArray(f32(n, 8, 16)), R.RefArray(np.abs(f32(e, 8, 1)))
    O3 = R.RefArray(shape=(n, 8, 16), dtype=np.float32)
    row("C3 arxiv", "u_mul_e+sum [N,8,16]x[E,8,1]",
        t(lambda: R.binary_op_reduce("sum", "mul", g, 0, 2, FT, A, O3)), e)
    EL, ER = R.RefArray(f32(n, 8, 1)), R.RefArray(f32(n, 8, 1))
    OE = R.RefArray(shape=(e, 8, 1), dtype=np.float32)
    row("C3 arxiv", "u_add_v SDDMM [N,8,1]",
        t(lambda: R.binary_op_reduce("none", "add", g, 0, 1, EL, ER, OE)), e)
    # C4
    n, e = 1885136, 5668682
    g = R.RefGraph(n, rs.randint(0, n, e), rs.randint(0, n, e))
    g.materialize()
    M, O4 = R.RefArray(f32(e, 16)), R.RefArray(shape=(n, 16), dtype=np.float32)
    row("C4 AM", "copy_e+sum [E,16]", t(lambda: R.copy_reduce("sum", g, 2, M, O4)), e)
    # C5 at 1/8 scale (same generator family; the full graph takes minutes on the CPU path)
    n, e = 2449029 // 8 // 64 * 64, 61859140 // 8
    dev0, globals()["DEV"] = DEV, torch.device("cpu")
    try:
        src, dst = sbm_graph(n, e, 64, 0.85, 0.5, 5)
    finally:
        globals()["DEV"] = dev0
    g = R.RefGraph(n, src.numpy(), dst.numpy())
    g.materialize()
    X, O5 = R.RefArray(f32(n, 100)), R.RefArray(shape=(n, 100), dtype=np.float32)
    row("C5 products (1/8 scale)", "copy_src+sum D=100",
        t(lambda: R.copy_reduce("sum", g, 0, X, O5)), e)


def main():
    global L2
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r1_configs.json"))
    ap.add_argument("--only", default="c1,c3,c4,c5")
    args = ap.parse_args()
    L2 = torch.cuda.get_device_properties(DEV).L2_cache_size
    out = []
    for name in args.only.split(","):
        {"ref": ref_cpu, "c1": c1, "c2": c2, "c3": c3, "c4": c4, "c5": c5, "models": models}[name](out)
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()

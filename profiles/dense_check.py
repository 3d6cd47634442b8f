"""One-shot check of dgl_hack_b200.dense (opt-in BF16x9 GEMM): values and gradients against torch
on small shapes with every transpose form, then GraphConv with B200_DENSE=bf16x9 against the
default, then timing of [1M,256]x[256,256] forward+backward."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dgl_hack_b200 import dense  # noqa: E402

DEV = "cuda:0"
out = {}
try:
    torch.manual_seed(0)
    worst = 0.0
    for M, K, N in ((300, 100, 256), (1000, 256, 48), (77, 13, 5), (4096, 256, 256)):
        a = (torch.rand(M, K, device=DEV) * 2 - 1).requires_grad_(True)
        b = (torch.rand(K, N, device=DEV) * 2 - 1).requires_grad_(True)
        g = torch.rand(M, N, device=DEV) * 2 - 1
        ref = a.double() @ b.double()
        c = dense.matmul(a, b)
        c.backward(g)
        errs = [float((c.double() - ref).abs().max() / ref.abs().max()),
                float((a.grad.double() - g.double() @ b.double().t()).abs().max() / (g.double() @ b.double().t()).abs().max()),
                float((b.grad.double() - a.double().t() @ g.double()).abs().max() / (a.double().t() @ g.double()).abs().max())]
        worst = max(worst, *errs)
        out["err_%dx%dx%d" % (M, K, N)] = errs
    out["worst_rel_err"] = worst
    out["computes"] = sorted({v[5] for v in dense._Lt.get().plans.values()})
    # GraphConv with and without
    import dgl_hack_b200.nn as nnb
    from dgl_hack_b200.graph import Graph
    n, e = 5000, 60000
    gr = Graph(torch.randint(0, n, (e,)), torch.randint(0, n, (e,)), num_nodes=n, device=DEV)
    conv = nnb.GraphConv(100, 256).to(DEV)
    x = torch.rand(n, 100, device=DEV)
    y0 = conv(gr, x)
    os.environ["B200_DENSE"] = "bf16x9"
    y1 = conv(gr, x)
    os.environ.pop("B200_DENSE")
    out["graphconv_rel_diff"] = float((y0 - y1).abs().max() / y0.abs().max())
    # timing
    M, K, N = 1 << 20, 256, 256
    a = (torch.rand(M, K, device=DEV)).requires_grad_(True)
    b = (torch.rand(K, N, device=DEV)).requires_grad_(True)
    g = torch.rand(M, N, device=DEV)

    def t(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 5

    def step(mm):
        def f():
            mm(a, b).backward(g)
            a.grad = b.grad = None
        return f
    out["ms_fwd_bwd_torch"] = round(t(step(torch.matmul)), 3)
    out["ms_fwd_bwd_bf16x9"] = round(t(step(dense.matmul)), 3)
except Exception as ex:
    out["error"] = repr(ex)[:400]
print(json.dumps(out), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "r1_dense_check.json"), "w"), indent=1)

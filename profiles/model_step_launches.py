"""One model step between cudaProfilerStart/Stop for an ncu launch list:
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/x.csv python profiles/model_step_launches.py gat_fused|gat|rgcn
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import dgl_hack_b200.nn as nnb  # noqa: E402
from dgl_hack_b200.graph import Graph  # noqa: E402

DEV = torch.device("cuda:0")
which = sys.argv[1] if len(sys.argv) > 1 else "gat_fused"
r = lambda *s: torch.rand(*s, device=DEV) * 2 - 1  # noqa: E731
if which.startswith("gat"):
    n, e, H, Fw = 169343, 1166243, 8, 16
    src, dst = bench.make_graph_coo(n, e, 0.4, 3, DEV)
    g = Graph(src, dst, num_nodes=n, device=DEV)
    conv = (nnb.FusedGATConv if which == "gat_fused" else nnb.GATConv)(128, Fw, H).to(DEV)
    x, go = r(n, 128), r(n, H, Fw)
    step = lambda: conv(g, x).backward(go)  # noqa: E731
else:
    n, e, R, d = 1885136, 5668682, 133, 16
    gen = torch.Generator(device=DEV)
    gen.manual_seed(4)
    src = torch.randint(0, n, (e,), generator=gen, device=DEV)
    dst = torch.randint(0, n, (e,), generator=gen, device=DEV)
    w = (torch.arange(1, R + 1, device=DEV, dtype=torch.float64)) ** -1.1
    et = torch.multinomial(w, e, replacement=True, generator=gen)
    g = Graph(src, dst, num_nodes=n, device=DEV)
    conv = nnb.RelGraphConv(d, d, R, "basis", num_bases=R, self_loop=True).to(DEV)
    x, norm, go = r(n, d), torch.rand(e, 1, device=DEV), r(n, d)
    step = lambda: conv(g, x, et, norm).backward(go)  # noqa: E731
for _ in range(3):
    step()
    for p in conv.parameters():
        p.grad = None
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
step()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()

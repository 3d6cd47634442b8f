"""Runs the round-1 caller-layer kernels once each between cudaProfilerStart/Stop (arxiv-shaped
graph) for `ncu --set full --profile-from-start off`:
edge_softmax fwd/bwd, attn_proj fwd/bwd, sample_neighbors, SpMM-routed u_add_v gradients."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from dgl_hack_b200 import backend as B, kernel as K, sampling as S  # noqa: E402
from dgl_hack_b200.graph import Graph  # noqa: E402
from dgl_hack_b200.nn import edge_softmax  # noqa: E402

DEV = torch.device("cuda:0")
n, e, H, Fw = 169343, 1166243, 8, 16
src, dst = bench.make_graph_coo(n, e, 0.4, 3, DEV)
g = Graph(src, dst, num_nodes=n, device=DEV)
r = lambda *s: torch.rand(*s, device=DEV) * 2 - 1  # noqa: E731
z = r(e, H, 1).requires_grad_(True)
go = r(e, H, 1)
ft = r(n, H, Fw).requires_grad_(True)
al, ar = r(1, H, Fw).requires_grad_(True), r(1, H, Fw).requires_grad_(True)
gn = r(n, H, 1)
seeds = torch.randperm(n, device=DEV)[:20000]


def once():
    edge_softmax(g, z).backward(go)
    el, er = B.gat_attn_proj(ft, al, ar)
    (el * gn + er * gn).sum().backward()
    S.sample_neighbors(g, seeds, 25, replace=True, rng_seed=3)
    S.sample_neighbors(g, seeds, 25, replace=False, rng_seed=3)
    z.grad = ft.grad = al.grad = ar.grad = None


for _ in range(3):
    once()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
once()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()

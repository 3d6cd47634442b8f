"""Host-side profile (cProfile) of the sampled GraphSAGE step of bench_sampling.py: which Python /
ctypes / torch calls the ~4 ms step spends its host time in."""
import cProfile
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import dgl_hack_b200.nn as nnb  # noqa: E402
from dgl_hack_b200 import sampling as S  # noqa: E402
from dgl_hack_b200.graph import Graph  # noqa: E402

DEV = torch.device("cuda:0")
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.25
n, e, d, hid, ncls = int(232965 * scale), int(114615892 * scale), 602, 16, 41
src, dst = bench.make_graph_coo(n, e, 0.6, 2, DEV)
g = Graph(src, dst, num_nodes=n, device=DEV)
x = torch.rand(n, d, device=DEV) * 2 - 1
y = torch.randint(0, ncls, (n,), device=DEV)
l1 = nnb.SAGEConv(d, hid, "mean", activation=torch.relu).to(DEV)
l2 = nnb.SAGEConv(hid, ncls, "mean").to(DEV)
opt = torch.optim.Adam(list(l1.parameters()) + list(l2.parameters()), lr=3e-3)


def step():
    seeds = torch.randperm(n, device=DEV)[:1000]
    blocks, cur = [], seeds
    for f in (10, 25):
        b = S.to_block(S.sample_neighbors(g, cur, f, replace=True), cur)
        cur = b.srcdata[S.NID]
        blocks.insert(0, b)
    h = x[blocks[0].srcdata[S.NID]]
    for layer, b in zip((l1, l2), blocks):
        h = layer(b, (h, h[:b.number_of_dst_nodes()]))
    loss = torch.nn.functional.cross_entropy(h, y[seeds])
    opt.zero_grad()
    loss.backward()
    opt.step()


for _ in range(5):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr, stream=sys.stdout)
st.sort_stats("tottime").print_stats(32)
st.sort_stats("cumulative").print_stats(40)

"""Target of the ncu capture of the GCN epoch's aggregation kernels on the products-shaped SBM
graph (1 GPU): scaled_copy_sum at D = 48 (short-row kernel), 100 and 256 (general kernel).

    ncu --set full --clock-control none --import-source on -k regex:spmm -c 6 \
        -o gpurun_out/r2_gcn_spmm_full python profiles/gcn_spmm_ncu_target.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_partitioned as BP  # noqa: E402
import dgl_hack_b200 as B  # noqa: E402
from dgl_hack_b200 import backend as F  # noqa: E402

dev = torch.device("cuda:0")
n = (BP.PRODUCTS_N // 64) * 64
src, dst = BP.sbm_graph(n, BP.PRODUCTS_E, 64, 0.85, 0.5, 5, dev)
g = B.GraphIndex.from_coo(n, n, src, dst, dev)
cs = torch.rand(n, device=dev) + 0.5
rs = torch.rand(n, device=dev) + 0.5
for d in (48, 100, 256):
    x = torch.rand(n, d, device=dev)
    for _ in range(2):
        out = F.scaled_copy_sum(g, x, cs, rs)
    torch.cuda.synchronize()
    del x, out
print("ok")

"""Target of the ncu capture of the own GEMMs: [1M,256] x [256,256] forward (bias + ReLU fused) and
the [1M,256]^T x [1M,256] weight gradient, two launches each.

    ncu --set full --clock-control none --import-source on -k regex:gemm_t -c 4 \
        -o gpurun_out/r2_gemm_full python profiles/gemm_ncu_target.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgl_hack_b200 import dense  # noqa: E402

dev = "cuda:0"
gen = torch.Generator(device=dev).manual_seed(0)
a = torch.rand(1 << 20, 256, generator=gen, device=dev) * 2 - 1
w = torch.rand(256, 256, generator=gen, device=dev) * 2 - 1
b = torch.rand(256, generator=gen, device=dev)
g = torch.rand(1 << 20, 256, generator=gen, device=dev) * 2 - 1
for _ in range(2):
    y = dense.tc_gemm(a, w, True, b, True)
    gw = dense.tc_weight_grad(a, g)
torch.cuda.synchronize()
print("ok", float(y[0, 0]), float(gw[0, 0]))

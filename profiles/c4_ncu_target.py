"""Target of the ncu capture of BASELINE config 4's batched per-relation copy_src+sum (ONE launch
on the relation-segment graph of the AM-shaped heterograph, D=16).

    ncu --set full --clock-control none --import-source on -k regex:spmm_kernel -c 3 \
        -o gpurun_out/r2_c4_full python profiles/c4_ncu_target.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgl_hack_b200 as B  # noqa: E402
from dgl_hack_b200 import backend as F  # noqa: E402

DEV = torch.device("cuda:0")
n, e, R, d = 1885136, 5668682, 133, 16
gen = torch.Generator(device=DEV)
gen.manual_seed(4)
src = torch.randint(0, n, (e,), generator=gen, device=DEV)
dst = torch.randint(0, n, (e,), generator=gen, device=DEV)
w = (torch.arange(1, R + 1, device=DEV, dtype=torch.float64)) ** -1.1
et = torch.multinomial(w, e, replacement=True, generator=gen)
x = torch.rand(n, d, device=DEV) * 2 - 1
g = B.GraphIndex.from_coo_with_type(n, src, dst, et, R, DEV)
sg, seg_dst, seg_type = g.relation_segments()
for _ in range(3):
    out = F.copy_reduce("sum", sg, 0, x, sg.num_dst)
torch.cuda.synchronize()
print("rows", sg.num_dst, float(out[0, 0]))

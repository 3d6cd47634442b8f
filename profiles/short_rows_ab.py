"""A/B of the short-row kernel's mean-degree threshold (B200_SPMM_SHORT_DEG) on graphs with 7 and
25 edges per row: copy_src+sum forward, arxiv-shaped D=128 and products-shaped D=100 / 48."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "profiles"))
import bench  # noqa: E402
import bench_configs as BC  # noqa: E402
import dgl_hack_b200 as B  # noqa: E402
from dgl_hack_b200 import backend as F  # noqa: E402

DEV = torch.device("cuda:0")
rows = []
for name, n, e, alpha, seed, widths in (("arxiv", 169343, 1166243, 0.4, 3, (128, 16)),
                                        ("products", 2449029, 61859140, 0.5, 5, (100, 48))):
    src, dst = bench.make_graph_coo(n, e, alpha, seed, DEV)
    g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
    for d in widths:
        x = torch.rand(n, d, device=DEV) * 2 - 1
        res = {}
        outs = {}
        for deg in ("1", "64"):
            os.environ["B200_SPMM_SHORT_DEG"] = deg
            outs[deg] = F.copy_reduce("sum", g, 0, x, n)
            res[deg] = round(BC.timeit(lambda: F.copy_reduce("sum", g, 0, x, n)), 4)
        row = {"graph": name, "mean_degree": round(e / n, 1), "D": d, "ms_general": res["1"],
               "ms_short_rows": res["64"], "identical": bool(torch.equal(outs["1"], outs["64"]))}
        rows.append(row)
        print(json.dumps(row), flush=True)
    del g
json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "r2_short_rows_ab.json"), "w"), indent=1)

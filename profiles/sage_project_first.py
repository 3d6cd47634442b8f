"""SAGEConv-mean layer of the Reddit-shaped config (C2: 602 -> 16) forward + backward, the
reference's order (aggregate 602-wide rows, then fc_neigh) against project_first (fc_neigh's weight
first, aggregate 16-wide rows): SURVEY 8(f) rank 2, "fusing the step either side"."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import dgl_hack_b200.nn as nnb  # noqa: E402
from dgl_hack_b200.graph import Graph  # noqa: E402

DEV = torch.device("cuda:0")


def timeit(fn, reps=10, warm=3):
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


def main():
    n, e, d, hid = 232965, 114615892, 602, 16
    src, dst = bench.make_graph_coo(n, e, 0.6, 2, DEV)
    g = Graph(src, dst, num_nodes=n, device=DEV)
    del src, dst
    x = (torch.rand(n, d, device=DEV) * 2 - 1).requires_grad_(True)
    go = torch.rand(n, hid, device=DEV)
    rows, outs = [], []
    torch.manual_seed(0)
    ref = nnb.SAGEConv(d, hid, "mean").to(DEV)
    fast = nnb.SAGEConv(d, hid, "mean", project_first=True).to(DEV)
    fast.load_state_dict(ref.state_dict())
    for name, m in (("reference order (aggregate 602-wide)", ref), ("project_first (aggregate 16-wide)", fast)):
        def step():
            m(g, x).backward(go)
            x.grad = None
            for p in m.parameters():
                p.grad = None
        with torch.no_grad():
            fwd = timeit(lambda: m(g, x))
            outs.append(m(g, x))
        rows.append({"config": "C2 reddit-shaped", "op": "SAGEConv-mean 602->16, " + name,
                     "ms_fwd": round(fwd, 3), "ms_fwd_bwd": round(timeit(step), 3)})
        print(json.dumps(rows[-1]), flush=True)
    err = float((outs[0] - outs[1]).abs().max() / outs[0].abs().max())
    rows.append({"max_abs_diff_over_max_abs": err})
    print(json.dumps(rows[-1]), flush=True)
    out = os.path.join(ROOT, "gpurun_out", "r1_sage_project_first.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump(rows, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()

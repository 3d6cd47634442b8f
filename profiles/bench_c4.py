"""BASELINE config 4 as worded: per-relation copy_src+sum over the AM-shaped heterograph
(N=1,885,136, E=5,668,682, R=133 Zipf(1.1) relations, D=16), looped (the reference's
HeteroGraphConv structure: one launch per relation, every launch writes all N rows) vs batched
(ONE launch on the relation-segment graph, every receiving (relation, node) row written once),
and HeteroGraphConv({rel: GraphConv(16, 16)}) generic loop vs the fused typed kernel.

    python profiles/bench_c4.py [--out gpurun_out/r2_configs_c4.json] [--scale 1]

Algorithmic bytes (SpMM fwd): G + 4E + 4*rows*D + 4*(rows+1), rows = receiving
(relation, node) pairs, G = 4*N*D if X fits L2 else 4*E*D. Parity batched == loop is asserted.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "profiles"))
import bench_configs as BC  # noqa: E402
import dgl_hack_b200 as B  # noqa: E402
import dgl_hack_b200.nn as bnn  # noqa: E402
from dgl_hack_b200 import backend as F  # noqa: E402
from dgl_hack_b200.hetero import HeteroGraph  # noqa: E402

DEV = BC.DEV


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r2_configs_c4.json"))
    ap.add_argument("--scale", type=int, default=1)
    args = ap.parse_args()
    BC.L2 = torch.cuda.get_device_properties(DEV).L2_cache_size
    n, e, R, d = 1885136 // args.scale, 5668682 // args.scale, 133, 16
    gen = torch.Generator(device=DEV)
    gen.manual_seed(4)
    src = torch.randint(0, n, (e,), generator=gen, device=DEV)
    dst = torch.randint(0, n, (e,), generator=gen, device=DEV)
    w = (torch.arange(1, R + 1, device=DEV, dtype=torch.float64)) ** -1.1
    et = torch.multinomial(w, e, replacement=True, generator=gen)
    x = BC.rand(n, d)
    out = []

    # ---- the kernel-level pair: looped vs batched copy_src+sum
    g = B.GraphIndex.from_coo_with_type(n, src, dst, et, R, DEV)
    sg, seg_dst, seg_type = g.relation_segments()
    rows = sg.num_dst
    graphs = [B.GraphIndex.from_coo(n, n, src[et == r], dst[et == r], DEV) for r in range(R)]

    def loop():
        return [F.copy_reduce("sum", gr, 0, x, n) for gr in graphs]
    ms_loop = BC.timeit(loop, reps=3, warm=1)
    BC.rec(out, "C4 AM", "133 x copy_src+sum D=16, one launch per relation (reference structure)",
           ms_loop, e, R * (n * d * 4) + BC.gather_bytes(n, e, d) + e * 4,
           "every launch writes all %d rows" % n)
    nbytes = BC.gather_bytes(n, e, d) + e * 4 + rows * d * 4 + (rows + 1) * 4
    ms = BC.timeit(lambda: F.copy_reduce("sum", sg, 0, x, rows))
    BC.rec(out, "C4 AM", "batched per-relation copy_src+sum D=16, ONE launch (relation segments)",
           ms, e, nbytes, "%d receiving (relation, node) rows, each written once" % rows)
    xg = x.clone().requires_grad_(True)
    go = BC.rand(rows, d)

    def fb():
        F.copy_reduce("sum", sg, 0, xg, rows).backward(go)
        xg.grad = None
    ms_fb = BC.timeit(fb)
    BC.rec(out, "C4 AM", "batched per-relation copy_src+sum fwd+bwd", ms_fb, 2 * e,
           nbytes + BC.gather_bytes(rows, e, d) + e * 4 + n * d * 4 + (n + 1) * 4)
    # parity: every relation's receiving rows of the loop == the batched rows (same order of sums)
    batched = F.copy_reduce("sum", sg, 0, x, rows)
    worst = 0.0
    for r, o in enumerate(loop()):
        sel = seg_type == r
        worst = max(worst, float((o[seg_dst[sel]] - batched[sel]).abs().max()) if sel.any() else 0.0)
    assert worst <= 1e-5, worst
    out.append({"config": "C4 AM", "op": "parity batched vs loop", "max_abs_diff": worst,
                "speedup_vs_loop": round(ms_loop / ms, 1)})
    del graphs, g, sg

    # ---- the module-level pair: HeteroGraphConv of GraphConv(16, 16), sum aggregate
    edges = {("n", "r%d" % r, "n"): (src[et == r], dst[et == r]) for r in range(R)}
    hg = HeteroGraph(edges, {"n": n}, DEV)
    torch.manual_seed(0)
    conv = bnn.HeteroGraphConv({"r%d" % r: bnn.GraphConv(d, d) for r in range(R)}, "sum").to(DEV)
    xin = x.clone().requires_grad_(True)
    gout = BC.rand(n, d)

    def step():
        conv(hg, {"n": xin})["n"].backward(gout)
        xin.grad = None
        conv.zero_grad(set_to_none=True)
    res = {}
    for mode in ("1", "0"):
        os.environ["B200_HETERO_FUSED"] = mode
        res[mode] = conv(hg, {"n": xin})["n"].detach()
        ms_m = BC.timeit(step, reps=3 if mode == "0" else 20, warm=1 if mode == "0" else 5)
        BC.rec(out, "C4 AM", "HeteroGraphConv{133 x GraphConv(16,16)} fwd+bwd, %s path"
               % conv.last_path, ms_m, 2 * e, 0)
    rel = float((res["1"] - res["0"]).abs().max() / res["0"].abs().max())
    assert rel <= 1e-4, rel
    out.append({"config": "C4 AM", "op": "parity fused vs generic HeteroGraphConv",
                "max_rel_diff": rel})
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump({"peak_gbs": BC.PEAK, "rows": out}, f, indent=1)
    print("wrote", args.out)


if __name__ == "__main__":
    main()

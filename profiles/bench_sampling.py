"""Sampled mini-batch GraphSAGE step on the Reddit-shaped graph, everything on the GPU
(examples/pytorch/graphsage/train_sampling_multi_gpu.py: batch 1000, fan-outs 10 and 25, hidden 16,
replace=True): neighbour sampling + to_block per layer, feature gather, 2-layer SAGEConv-mean
forward + loss + backward + Adam. CUDA events around whole steps; phases timed separately."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import dgl_hack_b200.nn as nnb  # noqa: E402
from dgl_hack_b200 import sampling as S  # noqa: E402
from dgl_hack_b200.graph import Graph  # noqa: E402

DEV = torch.device("cuda:0")


def ev_ms(fn, reps):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--batch", type=int, default=1000)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--project-first", action="store_true",
                    help="SAGEConv(project_first='auto'): weight before the mean aggregation")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r1_sampling.json"))
    args = ap.parse_args()
    n, e, d, hid, ncls = int(232965 * args.scale), int(114615892 * args.scale), 602, 16, 41
    src, dst = bench.make_graph_coo(n, e, 0.6, 2, DEV)
    g = Graph(src, dst, num_nodes=n, device=DEV)
    del src, dst
    x = torch.rand(n, d, device=DEV) * 2 - 1
    y = torch.randint(0, ncls, (n,), device=DEV)
    pf = "auto" if args.project_first else False
    l1 = nnb.SAGEConv(d, hid, "mean", activation=torch.relu, project_first=pf).to(DEV)
    l2 = nnb.SAGEConv(hid, ncls, "mean", project_first=pf).to(DEV)
    opt = torch.optim.Adam(list(l1.parameters()) + list(l2.parameters()), lr=3e-3)
    fanouts = [10, 25]
    gen = torch.Generator(device=DEV)
    gen.manual_seed(0)
    state = {}

    def sample():
        seeds = torch.randperm(n, generator=gen, device=DEV)[:args.batch]
        blocks, cur = [], seeds
        for f in fanouts:  # the example samples the output layer first (fanouts[0] = 10)
            b = S.to_block(S.sample_neighbors(g, cur, f, replace=True), cur)
            cur = b.srcdata[S.NID]
            blocks.insert(0, b)
        state["blocks"], state["seeds"] = blocks, seeds

    def train():
        blocks, seeds = state["blocks"], state["seeds"]
        h = x[blocks[0].srcdata[S.NID]]
        for layer, b in zip((l1, l2), blocks):
            h = layer(b, (h, h[:b.number_of_dst_nodes()]))
        loss = torch.nn.functional.cross_entropy(h, y[seeds])
        opt.zero_grad()
        loss.backward()
        opt.step()

    def step():
        sample()
        train()
    for _ in range(5):
        step()
    ms_step = ev_ms(step, args.steps)
    ms_sample = ev_ms(sample, args.steps)
    ms_train = ev_ms(train, args.steps)
    b0 = state["blocks"][0]
    line = {"metric": "sampled GraphSAGE step (sample + gather + fwd + bwd + Adam)",
            "config": {"workload": "Reddit-shaped N=%d E=%d D=%d, batch %d, fan-outs %s, replace"
                       % (n, e, d, args.batch, fanouts),
                       "project_first": bool(args.project_first)},
            "ms_per_step": round(ms_step, 3), "ms_sampling": round(ms_sample, 3),
            "ms_train": round(ms_train, 3),
            "seeds_per_s": round(args.batch / ms_step * 1e3, 1),
            "input_nodes_last_batch": b0.number_of_src_nodes(),
            "edges_last_batch": [b.number_of_edges() for b in state["blocks"]]}
    print(json.dumps(line), flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(line, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()

"""Full-graph 3-layer GCN epoch on the products-shaped graph at 1/2/4/8 GPUs (BASELINE config 5).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29521 profiles/bench_gcn_epoch.py [--scale 1.0] [--epochs 5]

Model = examples/pytorch/gcn (GraphConv, norm='both', python/dgl/nn/pytorch/conv/graphconv.py:
100-160): 100 -> 256 -> 256 -> 47, weight applied before the aggregation when it shrinks the
features. The aggregation is this repo's halo-exchanging copy_src+sum
(dgl_hack_b200.partition.PartitionedGraph); the dense layers are torch matmuls (cuBLAS, plain
library GEMMs); weight gradients are all-reduced (NCCL). One epoch = forward + cross-entropy on
synthetic labels + backward + Adam step over the whole graph. Time = CUDA events, max over ranks.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "profiles"))
from bench_multigpu import sbm_graph  # noqa: E402
from dgl_hack_b200.partition import (Partition, PartitionedGraph, range_partition,  # noqa: E402
                                      kway_partition_on_rank0)


FUSED = os.environ.get("B200_GRAPHCONV_FUSED", "1") != "0"


class GCN(torch.nn.Module):
    def __init__(self, dims):
        super().__init__()
        self.w = torch.nn.ParameterList()
        self.b = torch.nn.ParameterList()
        for i, o in zip(dims[:-1], dims[1:]):
            w = torch.empty(i, o)
            torch.nn.init.xavier_uniform_(w)
            self.w.append(torch.nn.Parameter(w))
            self.b.append(torch.nn.Parameter(torch.zeros(o)))

    def forward(self, pg, h, norm_src, norm_dst):
        n = len(self.w)
        for k, (w, b) in enumerate(zip(self.w, self.b)):
            if FUSED:
                # degree scaling inside the aggregation kernels; a width that is not a multiple of
                # 4 floats is aggregated zero-padded (what nn.GraphConv does)
                if w.shape[0] > w.shape[1]:
                    pad = (-w.shape[1]) % 4
                    h = h @ (torch.nn.functional.pad(w, (0, pad)) if pad else w)
                    h = pg.copy_src_sum(h, norm_src, norm_dst)[:, :w.shape[1]]
                else:
                    h = pg.copy_src_sum(h, norm_src, norm_dst) @ w
                h = h + b      # (torch.addmm materialises the broadcast bias first: slower here)
            else:
                if w.shape[0] > w.shape[1]:      # graphconv.py:138-147: shrink first
                    h = (h @ w) * norm_src
                    h = pg.copy_src_sum(h)
                else:
                    h = pg.copy_src_sum(h * norm_src) @ w
                h = h * norm_dst + b
            if k + 1 < n:
                h = torch.relu(h)
        return h


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--out", default=None)
    ap.add_argument("--partition", default="range", choices=["range", "kway"],
                    help="range: contiguous ranges balanced by in-edges; kway: the library's "
                         "multilevel k-way partitioner (the reference's metis_partition role)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if "MASTER_ADDR" not in os.environ:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", "29521"
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    n = int(2449029 * args.scale) // 64 * 64
    e = int(61859140 * args.scale)
    dims = [100, 256, 256, 47]
    src, dst = sbm_graph(n, e, 64, 0.85, 0.5, 5, dev)
    part_s = 0.0
    if args.partition == "kway" and world > 1:
        part, part_s = kway_partition_on_rank0(n, src, dst, world, rank, dev)
    else:
        part = range_partition(n, dst, world)
    p = Partition.build(n, src, dst, part, rank, world)
    out_deg = torch.bincount(src, minlength=n)[p.inner_global].clamp(min=1).float()
    in_deg = torch.bincount(dst, minlength=n)[p.inner_global].clamp(min=1).float()
    norm_src, norm_dst = out_deg.pow(-0.5).unsqueeze(1), in_deg.pow(-0.5).unsqueeze(1)
    pg = PartitionedGraph(p, dev)
    del src, dst
    torch.cuda.empty_cache()
    gen = torch.Generator(device=dev)
    gen.manual_seed(100 + rank)
    x = torch.rand(p.n_inner, dims[0], generator=gen, device=dev) * 2 - 1
    y = torch.randint(0, dims[-1], (p.n_inner,), generator=gen, device=dev)
    torch.manual_seed(0)                      # identical initial weights on every rank
    model = GCN(dims).to(dev)
    opt = torch.optim.Adam(model.parameters(), lr=1e-2)
    params = list(model.parameters())

    def epoch():
        opt.zero_grad(set_to_none=True)
        logits = model(pg, x, norm_src, norm_dst)
        loss = torch.nn.functional.cross_entropy(logits, y, reduction="sum") / n
        loss.backward()
        if world > 1:
            flat = torch.cat([q.grad.reshape(-1) for q in params])
            dist.all_reduce(flat)
            off = 0
            for q in params:
                q.grad.copy_(flat[off:off + q.numel()].view_as(q))
                off += q.numel()
        opt.step()
        return loss

    for _ in range(args.warmup):
        epoch()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    prof = os.environ.get("B200_PROFILE") == "1"   # ncu --profile-from-start off: timed epochs only
    if prof:
        torch.cuda.cudart().cudaProfilerStart()
    a.record()
    for _ in range(args.epochs):
        loss = epoch()
    b.record()
    torch.cuda.synchronize()
    if prof:
        torch.cuda.cudart().cudaProfilerStop()
    ms = torch.tensor([a.elapsed_time(b) / args.epochs], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    lsum = loss.detach().clone()
    dist.all_reduce(lsum)
    halo = torch.tensor([float(p.n_halo), float(p.cut_edges), float(p.local_edges)], device=dev)
    dist.all_reduce(halo)
    if rank == 0:
        line = {"metric": "GCN epoch ms (3 layers 100-256-256-47, fwd+bwd+Adam)", "n_gpus": world,
                "epoch_ms": round(ms.item(), 3), "scaling": "strong",
                "aggregated_edges_per_s_G": round(6 * e / ms.item() / 1e6, 3),
                "config": {"workload": "products-shaped SBM N=%d E=%d, %s partition" %
                           (n, e, args.partition)},
                "partition_seconds_rank0": round(part_s, 2),
                "loss": round(lsum.item(), 5), "halo_nodes_total": int(halo[0].item()),
                "cut_fraction": round(halo[1].item() / max(halo[2].item(), 1), 4),
                "epochs": args.epochs, "warmup": args.warmup, "fused_norm": FUSED}
        print(json.dumps(line), flush=True)
        if args.out:
            os.makedirs(os.path.dirname(args.out), exist_ok=True)
            with open(args.out, "a") as f:
                f.write(json.dumps(line) + "\n")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

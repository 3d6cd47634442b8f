"""Partitioned copy_src+sum with halo exchange on N GPUs (BASELINE config 5 shape).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 profiles/bench_multigpu.py [--scale 1.0] [--feat 100] [--check]

Strong scaling of ONE graph: the products-shaped SBM graph (64 blocks, p_in 0.85) is range
partitioned over the ranks (blocks are contiguous id ranges, so this is the METIS-like cut the
reference's `partition_graph_with_halo` would be given). Each step = forward + backward of the
halo-exchanging aggregation; time = CUDA events, max over ranks. `--check` compares rank results
with the single-GPU kernel on rank 0's device (small scales only).
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
import dgl_hack_b200 as B  # noqa: E402
from dgl_hack_b200 import backend as F  # noqa: E402
from dgl_hack_b200.partition import (Partition, PartitionedGraph, range_partition,  # noqa: E402
                                      kway_partition_on_rank0)


def sbm_graph(n, e, blocks, p_in, alpha, seed, dev):
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    bsz = n // blocks
    sb = torch.randint(0, blocks, (e,), generator=gen, device=dev)
    same = torch.rand(e, generator=gen, device=dev) < p_in
    db = torch.where(same, sb, torch.randint(0, blocks, (e,), generator=gen, device=dev))

    def within():
        u = torch.rand(e, generator=gen, device=dev, dtype=torch.float64)
        r = ((bsz ** (1.0 - alpha) - 1.0) * u + 1.0) ** (1.0 / (1.0 - alpha)) - 1.0
        return r.floor().clamp_(0, bsz - 1).to(torch.int64)
    perm = torch.randperm(bsz, generator=gen, device=dev)
    src = sb * bsz + perm[within()]
    dst = db * bsz + perm[within()]
    return src, dst


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--feat", type=int, default=100)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--breakdown", action="store_true",
                    help="also time every phase alone (synchronised; no overlap)")
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
    dist.init_process_group("nccl", device_id=dev)
    n = int(2449029 * args.scale) // 64 * 64
    e = int(61859140 * args.scale)
    d = args.feat
    src, dst = sbm_graph(n, e, 64, 0.85, 0.5, 5, dev)
    part_s = 0.0
    if args.partition == "kway" and world > 1:
        part, part_s = kway_partition_on_rank0(n, src, dst, world, rank, dev)
    else:
        part = range_partition(n, dst, world)
    p = Partition.build(n, src, dst, part, rank, world)
    pg = PartitionedGraph(p, dev)
    gen = torch.Generator(device=dev)
    gen.manual_seed(17)
    x_full = torch.rand(n, d, generator=gen, device=dev) * 2 - 1
    w_full = torch.rand(n, d, generator=gen, device=dev) * 2 - 1
    x = x_full[p.inner_global].clone().requires_grad_(True)
    gout = w_full[p.inner_global].clone()
    check = None
    if args.check:
        g = B.GraphIndex.from_coo(n, n, src, dst, dev)
        xg = x_full.clone().requires_grad_(True)
        ref = F.copy_reduce("sum", g, 0, xg, n)
        ref.backward(w_full)
        out = pg.copy_src_sum(x)
        out.backward(gout)
        ef = (out.detach() - ref.detach()[p.inner_global]).abs().max().item()
        eb = (x.grad - xg.grad[p.inner_global]).abs().max().item()
        scale_f = ref.detach().abs().max().item()
        check = {"fwd_max_abs_err": ef, "bwd_max_abs_err": eb, "ref_max_abs": scale_f}
        x.grad = None
        del g, xg, ref, out
    del src, dst, x_full, w_full
    torch.cuda.empty_cache()

    def step():
        o = pg.copy_src_sum(x)
        o.backward(gout)
        x.grad = None

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        step()
    b.record()
    torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b) / args.steps], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    phases = None
    if args.breakdown:
        def t_ms(fn, reps=5):
            fn()
            torch.cuda.synchronize()
            dist.barrier()
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record()
            for _ in range(reps):
                fn()
            b_.record()
            torch.cuda.synchronize()
            return a_.elapsed_time(b_) / reps
        xd = x.detach()
        g_halo = torch.rand(p.n_halo, d, device=dev)

        def ex_fwd():
            reqs, xh, keep = p.start_forward_exchange(xd)
            for r in reqs:
                r.wait()

        def ex_bwd():
            reqs, recv = p.start_backward_exchange(g_halo)
            for r in reqs:
                r.wait()
            return recv
        recv = ex_bwd()
        xh = torch.rand(max(p.n_halo, 1), d, device=dev)
        names = ["exchange_fwd", "agg_inner", "agg_halo", "agg_t_halo", "exchange_bwd",
                 "agg_t_inner", "reduce_received"]
        fns = [ex_fwd, lambda: pg._agg("inner", xd), lambda: pg._agg("halo", xh[:max(p.n_halo, 1)]),
               lambda: pg._agg_t("halo", gout), ex_bwd, lambda: pg._agg_t("inner", gout),
               lambda: p.reduce_received(torch.zeros_like(xd), recv) if torch.is_tensor(recv) else None]
        vals = torch.tensor([t_ms(f) for f in fns], device=dev)
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        phases = {k: round(v, 4) for k, v in zip(names, vals.tolist())}
    stats = torch.tensor([p.local_edges, p.cut_edges, p.n_inner, p.n_halo], device=dev,
                         dtype=torch.float64)
    allstats = [torch.zeros_like(stats) for _ in range(world)]
    dist.all_gather(allstats, stats)
    if check is not None:
        c = torch.tensor([check["fwd_max_abs_err"], check["bwd_max_abs_err"]], device=dev)
        dist.all_reduce(c, op=dist.ReduceOp.MAX)
        check["fwd_max_abs_err"], check["bwd_max_abs_err"] = c.tolist()
    if rank == 0:
        st = torch.stack(allstats).cpu()
        line = {"metric": "partitioned copy_src+sum fwd+bwd", "n_gpus": world, "scaling": "strong",
                "ms_per_step": round(ms.item(), 4),
                "gedges_per_s": round(2 * e / ms.item() / 1e6, 3),
                "config": {"workload": "products-shaped SBM N=%d E=%d D=%d, %s partition" %
                           (n, e, d, args.partition)},
                "partition_seconds_rank0": round(part_s, 2),
                "cut_fraction": round(float(st[:, 1].sum() / st[:, 0].sum()), 4),
                "max_local_edges": int(st[:, 0].max()), "max_halo_nodes": int(st[:, 3].max()),
                "halo_bytes_per_step_per_rank_max": int(st[:, 3].max()) * d * 4 * 2,
                "halo_path": pg.halo_path, "steps": args.steps, "warmup": args.warmup, "check": check,
                "phase_ms_max_over_ranks": phases}
        print(json.dumps(line), flush=True)
        if args.out:
            os.makedirs(os.path.dirname(args.out), exist_ok=True)
            with open(args.out, "a") as f:
                f.write(json.dumps(line) + "\n")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

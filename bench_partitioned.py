"""The partitioned multi-GPU leg of bench.py (BASELINE.json configs[4]).

ONE graph -- the ogbn-products-shaped SBM graph (2,449,029 nodes, 61,859,140 edges, 64 planted
blocks, 85 % intra-block edges, power-law endpoints) -- is split over the N ranks with the
library's multilevel k-way partitioner (the reference's `metis_partition` role,
python/dgl/transform.py:588-640, src/graph/metis_partition.cc:19-66); every rank owns its
nodes' in-edges (`partition_graph_with_halo` semantics, transform.py:551-586,
src/graph/graph_op.cc:403-441, tools/partition.py:8-53). Measured in the same run, strong scaling:

  (i)  copy_src+sum forward + backward (D = 100) through PartitionedGraph.copy_src_sum;
  (ii) the 3-layer GCN epoch (100 -> 256 -> 256 -> 47, GraphConv norm='both',
       python/dgl/nn/pytorch/conv/graphconv.py:128-176; forward + loss + backward + Adam).

Every run is also a PARITY run: the N-GPU output and gradient of (i) are compared on every rank
with the 1-GPU kernel on the unpartitioned graph (rows permuted back) and ASSERTED to agree
within 1e-5 of sum|terms|; the partitioned GCN forward loss is asserted against the whole-graph
forward with the same weights. `exposed_comm_ms` = step time minus the time of the same kernels
with every slot aliased to local memory (no NVLink traffic, no barrier).
"""
import os
import time

import torch
import torch.distributed as dist

from dgl_hack_b200 import dense

PRODUCTS_N, PRODUCTS_E, PRODUCTS_D = 2449029, 61859140, 100
GCN_DIMS = [100, 256, 256, 47]


def sbm_graph(n, e, blocks, p_in, alpha, seed, dev):
    """SURVEY 8(d) C5 generator on the device: stochastic block model with `blocks` equal blocks
    (p_in of the edges stay inside their block) and power-law endpoints inside a block."""
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


class GCN(torch.nn.Module):
    """examples/pytorch/gcn: GraphConv(norm='both') layers, the weight applied before the
    aggregation when it shrinks the features (graphconv.py:144-152); the degree scaling runs
    inside the aggregation kernels. A width that is not a multiple of 4 floats is aggregated
    zero-padded (what dgl_hack_b200.nn.GraphConv does)."""

    def __init__(self, dims):
        super().__init__()
        self.w = torch.nn.ParameterList()
        self.b = torch.nn.ParameterList()
        for i, o in zip(dims[:-1], dims[1:]):
            w = torch.empty(i, o)
            torch.nn.init.xavier_uniform_(w)
            self.w.append(torch.nn.Parameter(w))
            self.b.append(torch.nn.Parameter(torch.zeros(o)))

    def forward(self, agg, h):
        n = len(self.w)
        for k, (w, b) in enumerate(zip(self.w, self.b)):
            last = k + 1 == n
            if w.shape[0] > w.shape[1]:
                pad = (-w.shape[1]) % 4
                h = dense.linear(h, torch.nn.functional.pad(w, (0, pad)) if pad else w)
                h = agg(h)[:, :w.shape[1]] + b
                if not last:
                    h = torch.relu(h)
            else:
                # bias and ReLU ride on the GEMM's epilogue (B200_DENSE=tf32x3: fused in-kernel)
                h = dense.linear(agg(h), w, b, relu=not last)
        return h


def _max_over_ranks(vals, dev, world):
    t = torch.tensor(vals, device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def _timed(fn, steps, warmup, dev, world):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return _max_over_ranks([a.elapsed_time(b) / steps], dev, world)[0]


def run_partitioned(rank, world, dev, scale=1, steps=10, warmup=3, epochs=5, partition="kway"):
    """Returns the `partitioned` record (rank 0; None elsewhere). Collective over all ranks."""
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F
    from dgl_hack_b200.partition import (Partition, PartitionedGraph, LocalBlocks,
                                         range_partition, kway_partition_on_rank0)
    n = (PRODUCTS_N // scale) // 64 * 64
    e = PRODUCTS_E // scale
    d = PRODUCTS_D
    src, dst = sbm_graph(n, e, 64, 0.85, 0.5, 5, dev)
    part_s = 0.0
    if world > 1 and partition == "kway":
        part, part_s = kway_partition_on_rank0(n, src, dst, world, rank, dev)
    elif world > 1:
        part = range_partition(n, dst, world)
    else:
        part = torch.zeros(n, dtype=torch.int64, device=dev)
    t0 = time.perf_counter()
    p = Partition.build(n, src, dst, part, rank, world)
    pg = PartitionedGraph(p, dev)
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0

    # ---- inputs: identical on every rank (same generator), each rank keeps its inner rows
    gen = torch.Generator(device=dev)
    gen.manual_seed(17)
    x_full = torch.rand(n, d, generator=gen, device=dev) * 2 - 1
    w_full = torch.rand(n, d, generator=gen, device=dev) * 2 - 1
    y_full = torch.randint(0, GCN_DIMS[-1], (n,), generator=gen, device=dev)
    out_deg = torch.bincount(src, minlength=n).clamp(min=1).float()
    in_deg = torch.bincount(dst, minlength=n).clamp(min=1).float()
    ns_full, nd_full = out_deg.pow(-0.5), in_deg.pow(-0.5)
    ig = p.inner_global
    x = x_full[ig].clone().requires_grad_(True)
    gout = w_full[ig].clone()
    norm_src, norm_dst = ns_full[ig].contiguous(), nd_full[ig].contiguous()
    y = y_full[ig].clone()

    # ---- parity against the 1-GPU kernel on the unpartitioned graph (every rank checks its rows)
    g = B.GraphIndex.from_coo(n, n, src, dst, dev)
    del src, dst
    xg = x_full.clone().requires_grad_(True)
    ref = F.scaled_copy_sum(g, xg, ns_full, nd_full)
    ref.backward(w_full)
    with torch.no_grad():
        mag_f = F.scaled_copy_sum(g, x_full.abs(), ns_full, nd_full)
    out = pg.copy_src_sum(x, norm_src, norm_dst)
    out.backward(gout)
    tiny = 1e-30
    err_f = ((out.detach() - ref.detach()[ig]).abs() / (mag_f[ig] + tiny)).max().item() \
        if p.n_inner else 0.0
    # gradient magnitude: the transposed sum of |grad_out| with the scales swapped
    with torch.no_grad():
        mag_b = torch.empty_like(x_full)
        from dgl_hack_b200 import kernel as K
        K.scaled_copy_sum(g, True, w_full.abs().contiguous(), nd_full, ns_full, mag_b)
    err_b = ((x.grad - xg.grad[ig]).abs() / (mag_b[ig] + tiny)).max().item() if p.n_inner else 0.0
    x.grad = None
    del xg, ref, out, mag_f, mag_b
    parity = max(_max_over_ranks([err_f, err_b], dev, world))
    assert parity <= 1e-5, "partitioned aggregation differs from the 1-GPU kernel: %g" % parity

    # ---- GCN forward parity (same weights on every rank, whole graph vs partitioned)
    torch.manual_seed(0)
    model = GCN(GCN_DIMS).to(dev)
    with torch.no_grad():
        whole = model(lambda h: F.scaled_copy_sum(g, h, ns_full, nd_full), x_full)
        loss_whole = torch.nn.functional.cross_entropy(whole, y_full, reduction="sum") / n
        logits = model(lambda h: pg.copy_src_sum(h, norm_src, norm_dst), x.detach())
        loss_part = torch.nn.functional.cross_entropy(logits, y, reduction="sum") / n
        if world > 1:
            dist.all_reduce(loss_part)
        logit_err = ((logits - whole[ig]).abs().max() / whole.abs().max()).item() \
            if p.n_inner else 0.0
    loss_rel = abs(loss_part.item() - loss_whole.item()) / abs(loss_whole.item())
    logit_err = _max_over_ranks([logit_err], dev, world)[0]
    assert loss_rel <= 1e-5 and logit_err <= 1e-4, \
        "partitioned GCN forward differs from the whole-graph forward: loss %g logits %g" % (
            loss_rel, logit_err)
    del g, whole, logits, x_full, w_full, y_full, ns_full, nd_full, in_deg, out_deg
    torch.cuda.empty_cache()

    # ---- (i) aggregation step
    def agg_step():
        o = pg.copy_src_sum(x)
        o.backward(gout)
        x.grad = None
    agg_ms = _timed(agg_step, steps, warmup, dev, world)

    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        agg_step()
    agg_host_ms = (time.perf_counter() - t0) * 1e3 / steps
    torch.cuda.synchronize()
    if os.environ.get("B200_BENCH_PROFILE"):
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            for _ in range(4):
                agg_step()
            torch.cuda.synchronize()
        if rank == 0 or rank == world - 1:
            with open(os.environ["B200_BENCH_PROFILE"] + ".agg.rank%d" % rank, "w") as f:
                f.write("world %d, 4 agg steps, agg_ms %.4f host_enqueue_ms %.4f\n" % (
                    world, agg_ms, agg_host_ms))
                f.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30,
                                                  max_name_column_width=70))
            if rank == 0:
                prof.export_chrome_trace(os.environ["B200_BENCH_PROFILE"] + ".agg.trace.json")

    phases, local_ms = None, None
    if pg.gather is not None:
        gth = pg.gather
        xd = x.detach()

        def t_alone(fn):
            return _timed(fn, 5, 2, dev, world)
        bsel = [0]

        def stage_f():
            bsel[0] = gth.stage(0, xd, None)
        phases = {"stage_fwd": t_alone(stage_f), "barrier": t_alone(gth.blocks.barrier),
                  "gather_fwd": t_alone(lambda: gth.reduce(0, bsel[0], d, None, xd))}
        bsel[0] = gth.stage(1, gout, None)
        gth.blocks.barrier()
        phases["gather_bwd"] = t_alone(lambda: gth.reduce(1, bsel[0], d, None, xd))
        phases = {k: round(v, 4) for k, v in phases.items()}
        # the same kernels, indices and sizes with every slot aliased to local memory
        real = gth.blocks
        gth.blocks = LocalBlocks.group(world, p.rows_max, d, dev, alias_rank=rank,
                                       ext_max=p.ext_max)[rank]
        local_ms = _timed(agg_step, steps, warmup, dev, world)
        gth.blocks = real
        if world > 1:
            dist.barrier()

    # ---- (ii) GCN epoch
    opt = torch.optim.Adam(model.parameters(), lr=1e-2)
    params = list(model.parameters())

    def epoch():
        opt.zero_grad(set_to_none=True)
        lg = model(lambda h: pg.copy_src_sum(h, norm_src, norm_dst), x.detach())
        # per-row losses summed by a parallel reduction: torch's reduction="sum" path is a
        # one-block kernel (1.3 + 0.7 ms per epoch at 1.2 M rows in the round-2 profile)
        loss = torch.nn.functional.cross_entropy(lg, y, reduction="none").sum() / n
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
    last = [None]

    def run_epoch():
        last[0] = epoch()
    epoch_ms = _timed(run_epoch, epochs, 2, dev, world)
    # host time to ENQUEUE one epoch (no sync inside): close to epoch_ms = launch bound
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(epochs):
        run_epoch()
    host_ms = (time.perf_counter() - t0) * 1e3 / epochs
    torch.cuda.synchronize()
    prof_path = os.environ.get("B200_BENCH_PROFILE")
    if prof_path:
        # per-kernel device time of two epochs (torch.profiler / CUPTI), rank 0's table to a file
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            for _ in range(2):
                run_epoch()
            torch.cuda.synchronize()
        if rank == 0:
            with open(prof_path, "w") as f:
                f.write("world %d, 2 epochs, epoch_ms %.3f host_enqueue_ms %.3f\n" % (
                    world, epoch_ms, host_ms))
                f.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45,
                                                  max_name_column_width=70))
    # A/B of the other reduction modes on the same partition / blocks (B200_BENCH_MODES=a,b)
    mode_cmp = {}
    extra = [m for m in os.environ.get("B200_BENCH_MODES", "").split(",") if m]
    if pg.gather is not None and extra:
        from dgl_hack_b200.partition import PeerGather
        first = pg.gather
        mode_cmp[first.mode] = {"agg_ms_per_step": round(agg_ms, 4), "gcn_epoch_ms": round(epoch_ms, 3)}
        for m in extra:
            if m == first.mode:
                continue
            if "@" in m:     # "split@4": split mode with 4 pull CTAs per SM
                os.environ["B200_PULL_CTAS_PER_SM"] = m.split("@")[1]
            pg.gather = PeerGather(p, dev, first.blocks, mode=m.split("@")[0])
            a_ms = _timed(agg_step, steps, warmup, dev, world)
            e_ms = _timed(run_epoch, epochs, 2, dev, world)
            mode_cmp[m] = {"agg_ms_per_step": round(a_ms, 4), "gcn_epoch_ms": round(e_ms, 3)}
        pg.gather = first
    lsum = last[0].detach().clone()
    if world > 1:
        dist.all_reduce(lsum)

    st = torch.tensor([p.local_edges, p.cut_edges, p.n_inner, p.n_halo, p.n_ohalo,
                       int(p.bwd_coo[0].numel()) if p.bwd_coo is not None else 0],
                      device=dev, dtype=torch.float64)
    allst = [torch.zeros_like(st) for _ in range(world)]
    if world > 1:
        dist.all_gather(allst, st)
    else:
        allst = [st]
    if rank != 0:
        return None
    S = torch.stack(allst).cpu()
    cut = float(S[:, 1].sum())
    # rows crossing NVLink per rank and step: push / split / pull move every remote row once per
    # direction (forward halo + remote destinations of the backward pass); direct reads per edge
    per_edge = pg.gather is not None and pg.gather.mode == "direct"
    if world == 1:
        remote_reads = 0
    elif per_edge or pg.gather is None:
        remote_reads = int(2 * S[:, 1].max())
    else:
        remote_reads = int((S[:, 3] + S[:, 4]).max())
    rec = {
        "workload": "products_shaped_sbm_full_graph N=%d E=%d D=%d" % (n, e, d),
        "n_gpus": world, "scaling": "strong",
        "partition": partition if world > 1 else "none",
        "partition_seconds_rank0": round(part_s, 2), "plan_build_seconds": round(build_s, 2),
        "edge_cut": round(cut / max(float(S[:, 0].sum()), 1.0), 4),
        "local_edges_max": int(S[:, 0].max()), "inner_rows_max": int(S[:, 2].max()),
        "halo_rows_max": int(S[:, 3].max()), "halo_path": pg.halo_path,
        "halo_bytes_per_step": remote_reads * d * 4,
        "agg_ms_per_step": round(agg_ms, 4), "agg_gedges_per_s": round(2 * e / agg_ms / 1e6, 3),
        "agg_host_enqueue_ms": round(agg_host_ms, 4),
        "agg_local_only_ms": None if local_ms is None else round(local_ms, 4),
        "exposed_comm_ms": None if local_ms is None else round(agg_ms - local_ms, 4),
        "phase_ms": phases, "steps": steps, "warmup": warmup,
        "parity_max_rel_err": parity, "parity_bound": 1e-5,
        "gcn_epoch_ms": round(epoch_ms, 3), "gcn_epochs": epochs,
        "gcn_host_enqueue_ms": round(host_ms, 3),
        "gcn_dims": GCN_DIMS, "gcn_loss_after": round(lsum.item(), 5),
        "gcn_fwd_loss_rel_err": loss_rel, "gcn_fwd_logit_rel_err": logit_err,
        "dense": dense.mode(),
    }
    if pg.gather is not None:
        rec["halo_mode"] = pg.gather.mode
    if mode_cmp:
        rec["mode_comparison"] = mode_cmp
    return rec

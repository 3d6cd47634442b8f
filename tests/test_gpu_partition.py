"""Partitioned aggregation on the GPU kernels.

One-GPU case: both ranks' partitions are built in one process and the halo rows are copied by
hand, so the inner/halo kernels + transposes are checked against the whole-graph kernel.
Two-GPU case (skipped on a one-GPU box): the real NCCL exchange under torch.multiprocessing.
"""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _graph(n, e, seed, dev):
    g = torch.Generator().manual_seed(seed)
    src = torch.randint(0, n, (e,), generator=g)
    dst = torch.randint(0, n, (e,), generator=g)
    dst[: e // 10] = 7
    return src.to(dev), dst.to(dev)


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("d", [16, 100])
def test_partitioned_kernels_match_whole_graph(world, d):
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F
    from dgl_hack_b200.partition import Partition, PartitionedGraph, range_partition
    dev = torch.device("cuda:0")
    n, e = 3000, 60000
    src, dst = _graph(n, e, 21, dev)
    x = torch.rand(n, d, device=dev) * 2 - 1
    w = torch.rand(n, d, device=dev) * 2 - 1
    g = B.GraphIndex.from_coo(n, n, src, dst, dev)
    xg = x.clone().requires_grad_(True)
    ref = F.copy_reduce("sum", g, 0, xg, n)
    ref.backward(w)
    part = range_partition(n, dst, world)
    grad = torch.zeros_like(x)
    for r in range(world):
        p = Partition.build(n, src, dst, part, r, world)
        pg = PartitionedGraph(p, dev)
        out = pg._agg("inner", x[p.inner_global].contiguous())
        if p.n_halo:
            out = out + pg._agg("halo", x[p.halo_global].contiguous())
        # the hub row sums 6000 terms of +-1 in a different (inner, then halo) order
        torch.testing.assert_close(out, ref.detach()[p.inner_global], rtol=1e-5, atol=5e-4)
        go = w[p.inner_global].contiguous()
        grad.index_add_(0, p.inner_global, pg._agg_t("inner", go))
        if p.n_halo:
            grad.index_add_(0, p.halo_global, pg._agg_t("halo", go))
        # send lists of the other ranks cover exactly this rank's halo
        got = []
        for q in range(world):
            if q != r:
                pq = Partition.build(n, src, dst, part, q, world)
                got.append(pq.inner_global[pq.send_idx[r]])
        got = torch.cat(got) if got else torch.empty(0, dtype=torch.int64, device=dev)
        assert torch.equal(got, p.halo_global)
    torch.testing.assert_close(grad, xg.grad, rtol=1e-5, atol=5e-4)


def _nccl_worker(rank, world, port, halo, d, q):
    import torch.distributed as dist
    if ":" in halo:                      # "direct:<B200_HALO_MODE>"
        halo, os.environ["B200_HALO_MODE"] = halo.split(":")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import dgl_hack_b200 as B
        from dgl_hack_b200 import backend as F
        from dgl_hack_b200.partition import Partition, PartitionedGraph, range_partition
        n, e = 20000, 400000
        src, dst = _graph(n, e, 33, dev)
        gen = torch.Generator().manual_seed(1)
        x = (torch.rand(n, d, generator=gen) * 2 - 1).to(dev)
        w = (torch.rand(n, d, generator=gen) * 2 - 1).to(dev)
        cs = (torch.rand(n, generator=gen) + 0.5).to(dev)
        rs = (torch.rand(n, generator=gen) + 0.5).to(dev)
        g = B.GraphIndex.from_coo(n, n, src, dst, dev)
        part = range_partition(n, dst, world)
        if world > 2:
            part[part == world - 1] = 0          # the last rank owns no node at all
        p = Partition.build(n, src, dst, part, rank, world)
        pg = PartitionedGraph(p, dev, halo=halo)
        assert pg.halo_path == halo, (pg.halo_path, halo)
        res = []
        for scaled in (False, True):
            xg = x.clone().requires_grad_(True)
            ref = F.scaled_copy_sum(g, xg, cs, rs) if scaled else F.copy_reduce("sum", g, 0, xg, n)
            ref.backward(w)
            xi = x[p.inner_global].clone().requires_grad_(True)
            for _ in range(3):   # several rounds: the alternating blocks get reused
                xi.grad = None
                out = pg.copy_src_sum(xi, cs[p.inner_global], rs[p.inner_global]) if scaled \
                    else pg.copy_src_sum(xi)
                out.backward(w[p.inner_global].contiguous())
            res.append(torch.allclose(out.detach(), ref.detach()[p.inner_global], rtol=1e-5, atol=5e-4))
            res.append(torch.allclose(xi.grad, xg.grad[p.inner_global], rtol=1e-5, atol=5e-4))
        torch.cuda.synchronize()
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("halo,d", [("direct", 100), ("direct", 47), ("direct:split", 100),
                                    ("direct:pull", 100), ("direct:direct", 100), ("peer", 100),
                                    ("nccl", 47)])
def test_partitioned_multi_gpu(world, halo, d):
    """The real exchanges between processes (one per GPU): peer loads inside the SpMM
    ("direct"), peer-memory pushes ("peer") and NCCL all-to-all-v, against the whole-graph
    kernel; widths not divisible by 4 and a rank without nodes included."""
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, halo, d, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=120)
        assert pr.exitcode == 0
    assert all(all(r[1]) for r in res), res


@pytest.mark.parametrize("world,method", [(2, "range"), (4, "range"), (3, "empty_rank"),
                                          (2, "no_halo")])
@pytest.mark.parametrize("d", [100, 256, 47, 6])
@pytest.mark.parametrize("mode", ["push", "split", "pull", "direct"])
def test_direct_gather_one_process(world, method, d, mode, monkeypatch):
    """The single-pass peer-gather kernels (TAB SpMM, forward over all in-edges, backward over
    all out-edges) with every rank's staging block in THIS process (LocalBlocks): same kernels
    and codes as the multi-GPU run, checked against the whole-graph kernel incl. degree scales."""
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F
    from dgl_hack_b200.partition import (Partition, PartitionedGraph, LocalBlocks,
                                         range_partition)
    # push: the owners store the halo rows into the readers' receive buffers (side stream) while
    # the inner-source edges are reduced from x itself, halo-source edges added by a second pass;
    # split: halo rows pulled into a local buffer on a side stream while the inner-source edges
    # are reduced, halo-source edges added by a second pass; pull: pull, then one pass over two
    # local slots; direct: one pass gathering in place from every rank's block
    monkeypatch.setenv("B200_HALO_MODE", mode)
    dev = torch.device("cuda:0")
    n, e = 3000, 60000
    src, dst = _graph(n, e, 21, dev)
    if method == "no_halo":          # two disconnected halves: nothing crosses the cut
        half = n // 2
        src = torch.where(dst < half, src % half, half + src % half)
    gen = torch.Generator().manual_seed(3)
    x = (torch.rand(n, d, generator=gen) * 2 - 1).to(dev)
    w = (torch.rand(n, d, generator=gen) * 2 - 1).to(dev)
    cs = (torch.rand(n, generator=gen) + 0.5).to(dev)
    rs = (torch.rand(n, generator=gen) + 0.5).to(dev)
    g = B.GraphIndex.from_coo(n, n, src, dst, dev)
    xg = x.clone().requires_grad_(True)
    ref = F.scaled_copy_sum(g, xg, cs, rs)
    ref.backward(w)
    if method == "empty_rank":
        part = range_partition(n, dst, world - 1)
    elif method == "no_halo":
        part = (torch.arange(n, device=dev) >= n // 2).long()
    else:
        part = range_partition(n, dst, world)
    plans = [Partition.build(n, src, dst, part, r, world) for r in range(world)]
    if method == "no_halo":
        assert all(p.n_halo == 0 and p.n_ohalo == 0 for p in plans)
    blocks = LocalBlocks.group(world, plans[0].rows_max, 256, dev, ext_max=plans[0].ext_max)
    pgs = [PartitionedGraph(p, dev, halo="direct", blocks=b) for p, b in zip(plans, blocks)]
    assert all(pg.gather.mode == mode for pg in pgs)
    assert all(pg.halo_path == "direct" for pg in pgs)
    for _ in range(2):               # second round runs on the alternate blocks
        staged = [pg.gather.stage(0, x[p.inner_global].contiguous(), cs[p.inner_global])
                  for p, pg in zip(plans, pgs)]
        blocks[0].barrier()
        outs = [pg.gather.reduce(0, b, d, rs[p.inner_global].contiguous(), x)
                for p, pg, b in zip(plans, pgs, staged)]
        staged = [pg.gather.stage(1, w[p.inner_global].contiguous(), rs[p.inner_global])
                  for p, pg in zip(plans, pgs)]
        blocks[0].barrier()
        grads = [pg.gather.reduce(1, b, d, cs[p.inner_global].contiguous(), x)
                 for p, pg, b in zip(plans, pgs, staged)]
    for p, out, grad in zip(plans, outs, grads):
        assert out.shape == (p.n_inner, d)
        torch.testing.assert_close(out, ref.detach()[p.inner_global], rtol=1e-5, atol=5e-4)
        torch.testing.assert_close(grad, xg.grad[p.inner_global], rtol=1e-5, atol=5e-4)


@pytest.mark.gpu
def test_partitioned_scaled_copy_sum_world1():
    """PartitionedGraph.copy_src_sum with degree scales == backend.scaled_copy_sum on the graph."""
    import numpy as np
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F
    from dgl_hack_b200.partition import Partition, PartitionedGraph
    dev = torch.device("cuda:0")
    n, e = 500, 6000
    gen = torch.Generator().manual_seed(2)
    src = torch.randint(0, n, (e,), generator=gen)
    dst = torch.randint(0, n, (e,), generator=gen)
    part = torch.zeros(n, dtype=torch.int64)
    p = Partition.build(n, src.to(dev), dst.to(dev), part.to(dev), 0, 1)
    pg = PartitionedGraph(p, dev)
    g = B.GraphIndex.from_coo(n, n, src, dst, dev)
    x = torch.rand(n, 12, generator=gen).to(dev).requires_grad_(True)
    cs, rs = torch.rand(n, generator=gen).to(dev) + 0.1, torch.rand(n, generator=gen).to(dev) + 0.1
    go = torch.rand(n, 12, generator=gen).to(dev)
    a = pg.copy_src_sum(x, cs, rs)
    a.backward(go)
    ga, x.grad = x.grad.clone(), None
    b = F.scaled_copy_sum(g, x, cs, rs)
    b.backward(go)
    assert torch.equal(a, b) and torch.equal(ga, x.grad)

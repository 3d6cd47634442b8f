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


def _nccl_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import dgl_hack_b200 as B
        from dgl_hack_b200 import backend as F
        from dgl_hack_b200.partition import Partition, PartitionedGraph, range_partition
        n, e, d = 20000, 400000, 64
        src, dst = _graph(n, e, 33, dev)
        gen = torch.Generator().manual_seed(1)
        x = (torch.rand(n, d, generator=gen) * 2 - 1).to(dev)
        w = (torch.rand(n, d, generator=gen) * 2 - 1).to(dev)
        g = B.GraphIndex.from_coo(n, n, src, dst, dev)
        xg = x.clone().requires_grad_(True)
        ref = F.copy_reduce("sum", g, 0, xg, n)
        ref.backward(w)
        part = range_partition(n, dst, world)
        p = Partition.build(n, src, dst, part, rank, world)
        pg = PartitionedGraph(p, dev)
        xi = x[p.inner_global].clone().requires_grad_(True)
        out = pg.copy_src_sum(xi)
        out.backward(w[p.inner_global].contiguous())
        ok_f = torch.allclose(out.detach(), ref.detach()[p.inner_global], rtol=1e-5, atol=5e-4)
        ok_b = torch.allclose(xi.grad, xg.grad[p.inner_global], rtol=1e-5, atol=5e-4)
        q.put((rank, bool(ok_f), bool(ok_b)))
    finally:
        dist.destroy_process_group()


def test_partitioned_two_gpus_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    res = [q.get(timeout=300) for _ in range(2)]
    for pr in procs:
        pr.join(timeout=120)
        assert pr.exitcode == 0
    assert all(r[1] and r[2] for r in res), res


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

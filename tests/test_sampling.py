"""Mini-batch path. CPU: the to_block relabelling against the unmodified
reference (tests/golden/sampling.npz, make_golden_sampling.py). GPU: in_subgraph against the same
golden, sampler invariants / distribution, and blocks with full neighbourhoods reproducing the
full-graph SAGEConv output for the seed nodes."""
import os

import numpy as np
import pytest
import torch

import cases

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sampling.npz")
DEV = "cuda:0"
N = 300


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _graph_arrays():
    return cases.powerlaw_graph(N, N, 6000, 0.6, 11)


def test_compact_block_matches_reference_to_block(gold):
    from dgl_hack_b200.sampling import compact_block
    src_nid, src_local, dst_local = compact_block(
        torch.from_numpy(gold["in_sub/src"]), torch.from_numpy(gold["in_sub/dst"]),
        torch.from_numpy(gold["seeds"]))
    assert np.array_equal(src_nid.numpy(), gold["block/src_nid"])
    assert np.array_equal(src_local.numpy(), gold["block/src"])
    assert np.array_equal(dst_local.numpy(), gold["block/dst"])
    with pytest.raises(Exception):
        compact_block(torch.tensor([1, 2]), torch.tensor([5, 9]), torch.tensor([5, 6]))
    with pytest.raises(Exception):
        compact_block(torch.tensor([1, 2]), torch.tensor([5, 5]), torch.tensor([5, 5]))
    s, a, b = compact_block(torch.zeros(0, dtype=torch.int64), torch.zeros(0, dtype=torch.int64),
                            torch.tensor([7, 3]))
    assert s.tolist() == [7, 3] and a.numel() == 0 and b.numel() == 0


@pytest.mark.gpu
def test_in_subgraph_and_block_match_reference(gold):
    from dgl_hack_b200.graph import Graph
    from dgl_hack_b200 import sampling as S
    src, dst = _graph_arrays()
    g = Graph(src, dst, num_nodes=N, device=DEV)
    seeds = torch.from_numpy(gold["seeds"]).to(DEV)
    fr = S.in_subgraph(g, seeds)
    assert np.array_equal(fr.src.cpu().numpy(), gold["in_sub/src"])
    assert np.array_equal(fr.dst.cpu().numpy(), gold["in_sub/dst"])
    assert np.array_equal(fr.eid.cpu().numpy(), gold["in_sub/eid"])
    block = S.to_block(fr, seeds)
    assert np.array_equal(block.srcdata[S.NID].cpu().numpy(), gold["block/src_nid"])
    assert np.array_equal(block.dstdata[S.NID].cpu().numpy(), gold["block/dst_nid"])
    bs, bd = block.edges()
    assert np.array_equal(bs.cpu().numpy(), gold["block/src"])
    assert np.array_equal(bd.cpu().numpy(), gold["block/dst"])
    assert block.number_of_src_nodes() == 235 and block.number_of_dst_nodes() == 40


@pytest.mark.gpu
@pytest.mark.parametrize("fanout,replace", [(5, True), (5, False), (40, False), (100, False), (3, True)])
def test_sample_neighbors_invariants(fanout, replace):
    from dgl_hack_b200.graph import Graph
    from dgl_hack_b200 import sampling as S
    src, dst = _graph_arrays()
    g = Graph(src, dst, num_nodes=N, device=DEV)
    seeds = torch.arange(0, N, 2, device=DEV)
    fr = S.sample_neighbors(g, seeds, fanout, replace=replace, rng_seed=123)
    fs, fd, fe = fr.src.cpu().numpy(), fr.dst.cpu().numpy(), fr.eid.cpu().numpy()
    # every sampled edge is an edge of the graph, identified by its id
    assert np.array_equal(src[fe], fs) and np.array_equal(dst[fe], fd)
    deg = np.bincount(dst, minlength=N)
    got = np.bincount(fd, minlength=N)
    for v in range(N):
        if v % 2:
            assert got[v] == 0
        elif replace:
            assert got[v] == (fanout if deg[v] > 0 else 0)
        else:
            assert got[v] == min(deg[v], fanout)
    if not replace:  # distinct edges per seed
        assert np.unique(fe).size == fe.size
    # grouped by seed in seed order
    assert np.all(np.diff(fd) >= 0)
    # reproducible for a seed, different for another
    again = S.sample_neighbors(g, seeds, fanout, replace=replace, rng_seed=123)
    assert torch.equal(again.eid, fr.eid)
    other = S.sample_neighbors(g, seeds, fanout, replace=replace, rng_seed=124)
    if fanout < 40:
        assert not torch.equal(other.eid, fr.eid)


@pytest.mark.gpu
def test_sample_neighbors_is_uniform():
    """One hub with 64 in-edges: every edge is drawn about equally often, with and without
    replacement (chi-square well inside the 99.9 % quantile for 63 degrees of freedom: 103)."""
    from dgl_hack_b200.graph import Graph
    from dgl_hack_b200 import sampling as S
    src = np.arange(1, 65, dtype=np.int64)
    dst = np.zeros(64, np.int64)
    g = Graph(src, dst, num_nodes=65, device=DEV)
    seeds = torch.zeros(1, dtype=torch.int64, device=DEV)
    for replace in (True, False):
        counts = np.zeros(64)
        for t in range(2000):
            fr = S.sample_neighbors(g, seeds, 8, replace=replace, rng_seed=1000 + t)
            np.add.at(counts, fr.eid.cpu().numpy(), 1)
        expect = 2000 * 8 / 64
        chi2 = ((counts - expect) ** 2 / expect).sum()
        assert chi2 < 110, (replace, chi2)


@pytest.mark.gpu
def test_full_neighbourhood_blocks_equal_full_graph_layer():
    """examples/pytorch/graphsage/train_sampling_multi_gpu.py forward on blocks: with every
    neighbour kept (fanout >= max degree, no replacement) a 2-layer SAGE on the sampled blocks
    gives the full-graph result at the seed nodes, values and gradients."""
    import dgl_hack_b200.nn as nnb
    from dgl_hack_b200.graph import Graph
    from dgl_hack_b200 import sampling as S
    src, dst = _graph_arrays()
    g = Graph(src, dst, num_nodes=N, device=DEV)
    torch.manual_seed(0)
    l1 = nnb.SAGEConv(12, 16, "mean", activation=torch.relu).to(DEV)
    l2 = nnb.SAGEConv(16, 5, "mean").to(DEV)
    x = torch.rand(N, 12, device=DEV)
    full = l2(g, l1(g, x))
    seeds = torch.tensor([5, 17, 0, 250, 33, 34], device=DEV)
    blocks, cur = [], seeds
    for _ in range(2):
        fr = S.sample_neighbors(g, cur, 10 ** 6, replace=False)
        b = S.to_block(fr, cur)
        cur = b.srcdata[S.NID]
        blocks.insert(0, b)
    h = x[blocks[0].srcdata[S.NID]]
    for layer, b in zip((l1, l2), blocks):
        h = layer(b, (h, h[:b.number_of_dst_nodes()]))
    torch.testing.assert_close(h, full[seeds], rtol=1e-5, atol=1e-6)
    gw_full = torch.autograd.grad(full[seeds].sum(), l1.fc_neigh.weight, retain_graph=True)[0]
    gw_blk = torch.autograd.grad(h.sum(), l1.fc_neigh.weight)[0]
    torch.testing.assert_close(gw_blk, gw_full, rtol=1e-4, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("n_dst,n_e,span,seed", [(40, 300, 500, 0), (1, 1, 3, 1), (0, 0, 5, 2),
                                                 (5, 0, 9, 3), (3000, 200000, 50000, 4),
                                                 (1000, 5000, 2 ** 40, 5)])
def test_device_compact_block_equals_host_relabelling(n_dst, n_e, span, seed):
    """csrc/block.cu (hash table) against the torch relabelling on host tensors, bit for bit, with
    the destination lookup and with the sampler's known slots."""
    from dgl_hack_b200.sampling import compact_block
    rng = np.random.default_rng(seed)
    dst_nodes = torch.from_numpy(rng.choice(min(span, 10 ** 7), n_dst, replace=False).astype(np.int64)
                                 * max(span // 10 ** 7, 1))
    src = torch.from_numpy(rng.integers(0, span, n_e))
    slot = torch.from_numpy(rng.integers(0, max(n_dst, 1), n_e))
    dst = dst_nodes[slot] if n_dst else torch.zeros(0, dtype=torch.int64)
    want = compact_block(src, dst, dst_nodes)
    got = compact_block(src.to(DEV), dst.to(DEV), dst_nodes.to(DEV))
    for w, g in zip(want, got):
        assert torch.equal(w, g.cpu())
    got = compact_block(src.to(DEV), None, dst_nodes.to(DEV), slot.to(DEV))
    assert torch.equal(want[0], got[0].cpu()) and torch.equal(want[1], got[1].cpu())


@pytest.mark.gpu
def test_device_compact_block_errors():
    from dgl_hack_b200.sampling import compact_block
    t = lambda *v: torch.tensor(v, dtype=torch.int64, device=DEV)  # noqa: E731
    with pytest.raises(Exception, match="not a destination node"):
        compact_block(t(1, 2), t(5, 9), t(5, 6))
    with pytest.raises(Exception, match="must be unique"):
        compact_block(t(1, 2), t(5, 5), t(5, 5))
    with pytest.raises(Exception, match="not a destination node"):
        compact_block(t(1, 2), t(5, 1), t(5, 6))       # 1 is a source, not a destination
    with pytest.raises(Exception, match="negative"):
        compact_block(t(1, -2), t(5, 5), t(5, 6))


@pytest.mark.gpu
def test_sample_neighbors_rejects_more_than_int32_edges():
    """Repeated seeds with fanout = -1 can ask for more than 2^31 - 1 edges: reported, not wrapped."""
    from dgl_hack_b200.graph import Graph
    from dgl_hack_b200 import sampling as S
    deg = 70000
    g = Graph(np.arange(1, deg + 1, dtype=np.int64), np.zeros(deg, np.int64), num_nodes=deg + 1,
              device=DEV)
    seeds = torch.zeros(40000, dtype=torch.int64, device=DEV)
    with pytest.raises(Exception, match="2\\^31"):
        S.sample_neighbors(g, seeds, -1)
    fr = S.sample_neighbors(g, seeds[:3], -1)
    assert fr.number_of_edges() == 3 * deg
    with pytest.raises(Exception, match="out of range"):
        S.sample_neighbors(g, torch.tensor([deg + 1], device=DEV), 2)

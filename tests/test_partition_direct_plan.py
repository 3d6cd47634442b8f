"""Host logic of the single-pass partitioned aggregation (SURVEY 8e option iii) on CPU: the
(owner, row) codes, the forward graph over all in-edges and the backward graph over all out-edges
of every rank reproduce the whole-graph copy_src+sum and its gradient when every rank's block is
simply its rows of X (torch emulation of the TAB gather; the CUDA kernel is covered by -m gpu)."""
import pytest
import torch

from dgl_hack_b200.partition import Partition, SLOT_SHIFT


def _emulate(coo_rows, coo_cols, code, blocks, n_rows, d):
    slot = (code.long() >> SLOT_SHIFT)
    row = code.long() & ((1 << SLOT_SHIFT) - 1)
    flat = torch.cat(blocks)
    offs = torch.tensor([0] + [b.shape[0] for b in blocks[:-1]]).cumsum(0)
    gathered = flat[offs[slot[coo_cols]] + row[coo_cols]]
    return torch.zeros(n_rows, d, dtype=flat.dtype).index_add_(0, coo_rows, gathered)


@pytest.mark.parametrize("world,method", [(2, "random"), (3, "random"), (4, "skewed"), (3, "empty")])
def test_direct_plan_matches_whole_graph(world, method):
    n, e, d = 300, 5000, 3
    g = torch.Generator().manual_seed(world * 7 + len(method))
    src = torch.randint(0, n, (e,), generator=g)
    dst = torch.randint(0, n, (e,), generator=g)
    dst[: e // 8] = 5
    if method == "random":
        part = torch.randint(0, world, (n,), generator=g)
    elif method == "skewed":
        part = (torch.arange(n) * world // n)          # contiguous blocks
        part[:10] = world - 1
    else:                                               # one rank owns nothing
        part = torch.randint(0, world - 1, (n,), generator=g)
    x = torch.rand(n, d, generator=g, dtype=torch.float64)
    w = torch.rand(n, d, generator=g, dtype=torch.float64)
    full = torch.zeros(n, d, dtype=torch.float64).index_add_(0, dst, x[src])
    full_t = torch.zeros(n, d, dtype=torch.float64).index_add_(0, src, w[dst])
    plans = [Partition.build(n, src, dst, part, r, world) for r in range(world)]
    xb = [x[p.inner_global] for p in plans]
    wb = [w[p.inner_global] for p in plans]
    assert sum(p.n_inner for p in plans) == n
    for r, p in enumerate(plans):
        assert p.rows_max == max(q.n_inner for q in plans)
        # codes of the inner nodes address this rank's own block rows 0..n_inner-1
        own = p.fwd_code[: p.n_inner].long()
        assert torch.equal(own, (r << SLOT_SHIFT) + torch.arange(p.n_inner))
        assert torch.equal(p.bwd_code[: p.n_inner].long(), own)
        fs, fd = p.fwd_coo
        out = _emulate(fd, fs, p.fwd_code, xb, p.n_inner, d)
        assert torch.allclose(out, full[p.inner_global], rtol=1e-12, atol=1e-12)
        bs, bd = p.bwd_coo
        grad = _emulate(bs, bd, p.bwd_code, wb, p.n_inner, d)
        assert torch.allclose(grad, full_t[p.inner_global], rtol=1e-12, atol=1e-12)
        assert int(fs.numel()) == p.local_edges
        assert p.n_ohalo == int(torch.unique(dst[(part[src] == r) & (part[dst] != r)]).numel())

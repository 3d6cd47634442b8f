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


@pytest.mark.parametrize("world,method", [(2, "random"), (4, "random"), (3, "empty")])
def test_push_plan_fills_every_halo(world, method):
    """The push lists (Partition.push): executing every rank's forward list leaves, in each
    rank's receive buffer, exactly its halo rows in halo order; the backward list its remote
    destinations in the order of bwd_code. Receive offsets are derived without communication."""
    n, e, d = 300, 5000, 2
    g = torch.Generator().manual_seed(world * 11 + len(method))
    src = torch.randint(0, n, (e,), generator=g)
    dst = torch.randint(0, n, (e,), generator=g)
    part = torch.randint(0, world if method == "random" else world - 1, (n,), generator=g)
    x = torch.rand(n, d, generator=g, dtype=torch.float64)
    plans = [Partition.build(n, src, dst, part, r, world) for r in range(world)]
    mask = (1 << SLOT_SHIFT) - 1
    assert len({p.ext_max for p in plans}) == 1
    for which in (0, 1):
        n_ext = [p.n_halo if which == 0 else p.n_ohalo for p in plans]
        assert plans[0].ext_max >= max(n_ext)
        recv = [torch.full((plans[0].ext_max, d), float("nan"), dtype=torch.float64)
                for _ in range(world)]
        for r, p in enumerate(plans):
            rows, codes = (t.long() for t in p.push[which])
            assert ((codes >> SLOT_SHIFT) != r).all()
            # round-robin over the consumers: the k-th row for every consumer comes before the
            # (k+1)-th row for any consumer (no GPU's ingress takes one sender's whole list at once)
            cons = (codes >> SLOT_SHIFT).tolist()
            seen, nth = {}, []
            for c in cons:
                nth.append(seen.get(c, 0))
                seen[c] = nth[-1] + 1
            assert nth == sorted(nth)
            xb = x[p.inner_global]
            for q in range(world):
                sel = (codes >> SLOT_SHIFT) == q
                recv[q][codes[sel] & mask] = xb[rows[sel]]
        for q, p in enumerate(plans):
            code = (p.fwd_code if which == 0 else p.bwd_code)[p.n_inner:].long()
            want = torch.stack([x[plans[int(c >> SLOT_SHIFT)].inner_global[int(c & mask)]]
                                for c in code]) if code.numel() else torch.empty(0, d, dtype=x.dtype)
            assert torch.equal(recv[q][: n_ext[q]], want)
            assert torch.isnan(recv[q][n_ext[q]:]).all()

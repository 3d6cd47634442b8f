"""The bf16 feature-storage tier (DGLB200KernelCopySumBF16): explicit dtype, fp32 accumulation.
Tolerance = north_star's bf16 bound, 1e-2 relative (of sum|terms|), written here. Oracle = the
fp32 CPU oracle run on the bf16-ROUNDED inputs (so only the kernel's own accumulation + one final
rounding are measured); also exact on small-integer data whatever the order of the sum."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu
BF16_RTOL = 1e-2


def _graphs():
    yield "powerlaw", cases.powerlaw_graph(3000, 3000, 90000, 0.8, 7)      # hub rows -> segments
    yield "bipartite", cases.powerlaw_graph(500, 1200, 20000, 0.3, 8)[:2]


@pytest.mark.parametrize("d", [8, 16, 100, 602])
@pytest.mark.parametrize("reducer", ["sum", "mean"])
def test_bf16_copy_reduce_matches_oracle(d, reducer):
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F
    from oracle import oracle as O
    ns, nd, ne = 3000, 2500, 90000
    src, dst = cases.powerlaw_graph(ns, nd, ne, 0.8, 7)
    og = O.Graph(ns, nd, src, dst)
    g = B.GraphIndex.from_coo(ns, nd, src, dst, "cuda:0")
    rs = np.random.RandomState(d)
    x = torch.from_numpy(rs.uniform(-1, 1, (ns, d)).astype(np.float32)).cuda()
    xs = F.to_bf16_storage(x).requires_grad_(True)
    assert xs.dtype == torch.bfloat16 and xs.shape[1] % 8 == 0
    out = F.copy_reduce_bf16(reducer, g, xs)
    assert out.dtype == torch.bfloat16 and out.shape == (nd, xs.shape[1])
    xr = xs.detach().float().cpu().numpy()                  # the bf16-rounded inputs, exactly
    want = O.copy_reduce("sum", og, 0, xr, nd)
    mag = O.copy_reduce("sum", og, 0, np.abs(xr), nd)
    deg = np.maximum(og.in_degrees(), 1).astype(np.float32)[:, None]
    if reducer == "mean":
        want, mag = want / deg, mag / deg
    got = out.detach().float().cpu().numpy()
    assert (np.abs(got - want) <= BF16_RTOL * mag + 1e-30).all()
    assert not got[:, d:].any()                             # the zero padding stays zero
    # gradient: transposed walk of bf16 grad_out (scaled by 1/deg for mean)
    go = torch.from_numpy(rs.uniform(-1, 1, got.shape).astype(np.float32)).cuda().to(torch.bfloat16)
    out.backward(go)
    gor = go.float().cpu().numpy()
    if reducer == "mean":
        # the kernel multiplies the bf16 grad_out by the fp32 1/deg per neighbour
        gwant = O.backward_copy_reduce("sum", og, 0, xr, want, gor / deg)
        gmag = O.backward_copy_reduce("sum", og, 0, xr, want, np.abs(gor) / deg)
    else:
        gwant = O.backward_copy_reduce("sum", og, 0, xr, want, gor)
        gmag = O.backward_copy_reduce("sum", og, 0, xr, want, np.abs(gor))
    ggot = xs.grad.float().cpu().numpy()
    assert (np.abs(ggot - gwant) <= BF16_RTOL * gmag + 1e-30).all()


def test_bf16_exact_on_small_integers_and_explicit_dtype():
    import torch
    import dgl_hack_b200 as B
    from dgl_hack_b200 import backend as F, kernel as K
    from dgl_hack_b200._lib import DGLError
    n, e = 2000, 40000           # sums stay below 256: exactly representable in bf16
    rs = np.random.RandomState(3)
    src, dst = rs.randint(0, n, e), rs.randint(0, n, e)
    g = B.GraphIndex.from_coo(n, n, src, dst, "cuda:0")
    x = torch.from_numpy(rs.randint(-2, 3, (n, 24)).astype(np.float32)).cuda()
    out = F.copy_reduce_bf16("sum", g, F.to_bf16_storage(x))
    want = F.copy_reduce("sum", g, 0, x, n)
    assert torch.equal(out.float(), want)
    # fp32 tensors are rejected by the bf16 entry, bf16 tensors by the fp32 entries
    bad = torch.empty(n, 24, device="cuda")
    with pytest.raises(DGLError, match="bfloat16"):
        K.copy_sum_bf16(g, False, x, None, None, False, bad)
    with pytest.raises(DGLError):
        K.copy_reduce("sum", g, 0, x.to(torch.bfloat16), bad)
    with pytest.raises(DGLError, match="multiple of 8|D % 8"):
        F.copy_reduce_bf16("sum", g, x[:, :20].to(torch.bfloat16).contiguous())

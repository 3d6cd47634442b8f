"""DGLB200GatAttnProjBackwardNoAlloc (static device scratch, no allocation) against the existing
DGLB200GatAttnProjBackward, through the C ABI, on two calls in a row that share the scratch array.
Collected last on purpose. Passes alone in a fresh process and
after the other tests: it depends on no state but the library itself."""
import ctypes

import pytest

pytestmark = pytest.mark.gpu


def _run(entry, feat, al, ar, gl, gr):
    import torch
    from dgl_hack_b200._lib import check_call
    from dgl_hack_b200.ndarray import NDArray
    gfeat = torch.full_like(feat, float("nan"))
    gal = torch.full_like(al, float("nan")) if al is not None else None
    gar = torch.full_like(ar, float("nan")) if ar is not None else None
    nds = [None if t is None else NDArray(t) for t in (feat, al, ar, gl, gr, gfeat, gal, gar)]
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    check_call(entry(*[None if t is None else t.ptr for t in nds], stream))
    torch.cuda.synchronize()
    return gfeat, gal, gar


def _inputs(n, h, f, seed, both=True):
    import torch
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: (torch.rand(*s, generator=g) * 2 - 1).cuda()  # noqa: E731
    feat, al, gl = r(n, h, f), r(1, h, f), r(n, h, 1)
    return (feat, al, r(1, h, f), gl, r(n, h, 1)) if both else (feat, al, None, gl, None)


def test_no_alloc_matches_allocating_entry_two_calls_in_a_row():
    import torch
    from dgl_hack_b200._lib import LIB
    # second call: fewer rows and another width, so stale partials of the first call sit in the
    # shared array beyond what the second call overwrites; third: scalar (non-float4) path, one side
    cases = [_inputs(50000, 8, 16, 1), _inputs(777, 4, 32, 2), _inputs(3001, 3, 5, 3, both=False),
             _inputs(50000, 8, 16, 1)]
    for args in cases:
        want = _run(LIB.DGLB200GatAttnProjBackward, *args)
        got = _run(LIB.DGLB200GatAttnProjBackwardNoAlloc, *args)
        for w, g in zip(want, got):
            assert (w is None) == (g is None)
            if w is not None:
                assert not torch.isnan(g).any()
                # tolerance of the attn_proj tests (test_gpu_nn.py: rtol 1e-5); on a B200 the
                # blocking is the same and the results are bit-identical
                torch.testing.assert_close(g, w, rtol=1e-5, atol=1e-6)


def test_no_alloc_rejects_widths_beyond_the_static_scratch():
    from dgl_hack_b200._lib import LIB, DGLError
    args = _inputs(64, 8, 128, 4)            # H * F = 1024 > 512
    _run(LIB.DGLB200GatAttnProjBackward, *args)   # the allocating entry takes it
    with pytest.raises(DGLError, match="static scratch"):
        _run(LIB.DGLB200GatAttnProjBackwardNoAlloc, *args)

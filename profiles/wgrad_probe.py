"""Probes of the MN-major weight-gradient kernel. A[m, k] = 1000 m + k (exact in fp32 / TF32 for
the sizes used), G one-hot at (m0, n0): C[k, n0] must equal A[m0, k]; decoding every nonzero of C
as (m', k') shows which element each output position really read."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200_DENSE_WGRAD"] = "tc"
from dgl_hack_b200 import dense  # noqa: E402

dev = "cuda:0"
torch.set_printoptions(linewidth=200)
for (M, K, N) in [(16, 32, 16), (16, 128, 32)]:
    print("== M %d K %d N %d" % (M, K, N))
    a = (torch.arange(M, device=dev).view(-1, 1) * 64 + torch.arange(K, device=dev).view(1, -1) % 64 + 1).float()
    for (m0, n0) in [(0, 0), (1, 0), (2, 3), (5, 9), (9, 1)]:
        g = torch.zeros(M, N, device=dev)
        g[m0, n0] = 1.0
        c = dense.tc_weight_grad(a, g)
        torch.cuda.synchronize()
        nzc = torch.nonzero(c.abs().sum(0)).flatten().tolist()
        print("G one-hot (m0=%d, n0=%d): nonzero columns %s" % (m0, n0, nzc[:8]))
        for col in nzc[:2]:
            v = c[:min(K, 12), col]
            print("   col %d rows 0..11 -> (m', k'+1) = %s   want m'=%d k'+1=1..12" % (
                col, [(int(x) // 64, int(x) % 64) for x in v.tolist()], m0))
    # A one-hot, G dense
    gd = (torch.arange(M, device=dev).view(-1, 1) * 64 + torch.arange(N, device=dev).view(1, -1) + 1).float()
    for (m0, k0) in [(0, 0), (1, 0), (2, 5), (9, 17)]:
        a1 = torch.zeros(M, K, device=dev)
        a1[m0, k0] = 1.0
        c = dense.tc_weight_grad(a1, gd)
        torch.cuda.synchronize()
        nzr = torch.nonzero(c.abs().sum(1)).flatten().tolist()
        print("A one-hot (m0=%d, k0=%d): nonzero rows %s" % (m0, k0, nzr[:8]))
        for row in nzr[:2]:
            v = c[row, :min(N, 12)]
            print("   row %d cols 0..11 -> (m', n'+1) = %s   want m'=%d n'+1=1..12" % (
                row, [(int(x) // 64, int(x) % 64) for x in v.tolist()], m0))
    ones = dense.tc_weight_grad(torch.ones(M, K, device=dev), torch.ones(M, N, device=dev))
    print("all-ones: min %.1f max %.1f (want %d everywhere)" % (ones.min().item(), ones.max().item(), M))

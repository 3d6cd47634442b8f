"""Direct kernel-entry timing of the fused GAT ops on the arxiv-shaped graph (no autograd, outputs
preallocated), to separate GPU time from host overhead. python profiles/gat_probe.py [reps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import dgl_hack_b200 as B  # noqa: E402
from dgl_hack_b200 import kernel as K  # noqa: E402

DEV = torch.device("cuda:0")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
n, e, H, Fw = 169343, 1166243, 8, 16
src, dst = bench.make_graph_coo(n, e, 0.4, 3, DEV)
g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
print("max in-degree", int(g.in_degrees().max()), "max out-degree", int(g.out_degrees().max()))
r = lambda *s: torch.rand(*s, device=DEV) * 2 - 1  # noqa: E731
ft, el, er, gout = r(n, H, Fw), r(n, H, 1), r(n, H, 1), r(n, H, Fw)
s_, ex, ret = torch.empty(n, H, 1, device=DEV), torch.empty(e, H, 1, device=DEV), torch.empty(n, H, Fw, device=DEV)
gft, gel, ger = torch.empty_like(ft), torch.empty_like(el), torch.empty_like(er)


def timeit(fn):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


print("fwd  ms", timeit(lambda: K.fused_gat_kernel(g, ft, el, er, s_, ex, ret, 0.2)))
print("bwd  ms", timeit(lambda: K.backward_fused_gat(g, ft, el, er, s_, ex, ret, gout, gft, gel, ger, 0.2)))
x = r(n, H * Fw)
out = torch.empty(n, H * Fw, device=DEV)
print("copy_src+sum D=128 ms", timeit(lambda: K.copy_reduce("sum", g, 0, x, out)))
w = torch.rand(e, H, 1, device=DEV)
out3 = torch.empty(n, H, Fw, device=DEV)
print("u_mul_e+sum [N,8,16]x[E,8,1] ms", timeit(lambda: K.binary_op_reduce("sum", "mul", g, 0, 2, ft, w, out3)))

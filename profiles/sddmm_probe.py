"""u_dot_v D=64 and u_add_v [E,1] on the Reddit-shaped graph through the kernel entry points."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import dgl_hack_b200 as B  # noqa: E402
from dgl_hack_b200 import kernel as K  # noqa: E402

DEV = torch.device("cuda:0")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
n, e = 232965, 114615892
src, dst = bench.make_graph_coo(n, e, 0.6, 2, DEV)
g = B.GraphIndex.from_coo(n, n, src, dst, DEV)
del src, dst
x64 = torch.rand(n, 64, device=DEV)
out = torch.empty(e, device=DEV)


def timeit(fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


print("u_dot_v D=64 ms", timeit(lambda: K.binary_op_reduce("none", "dot", g, 0, 1, x64, x64, out)))
xs = torch.rand(n, 1, device=DEV)
o1 = torch.empty(e, 1, device=DEV)
print("u_add_v [E,1] ms", timeit(lambda: K.binary_op_reduce("none", "add", g, 0, 1, xs, xs, o1)))

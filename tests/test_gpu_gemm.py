"""The hand-written tcgen05 3xTF32 GEMM (csrc/gemm.cu) through the C ABI against an fp64 product:
within 1e-5 of sum|a||b| (north_star's fp32 tolerance; plain TF32 is ~3e-4) on full / partial
tiles, K tails, N not divisible by 16, both operand orders, fused bias + ReLU; and the autograd
route of dense.matmul (B200_DENSE=tf32x3). Runs in a child process under a timeout: a pipeline bug
in a warp-specialised kernel is a hang, and a hang must fail this test, not stall the suite."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tf32x3_gemm_against_fp64():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "profiles", "gemm_check.py"),
                          "--quick", "--wgrad"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, (out.stdout[-2000:], out.stderr[-2000:])
    last = json.loads(out.stdout.strip().splitlines()[-1])
    assert last["ok"] and last["worst_err_tf32x3"] <= 1e-5


def test_dense_matmul_route_gradients():
    code = r'''
import os, sys, torch
sys.path.insert(0, %r)
os.environ["B200_DENSE"] = "tf32x3"
from dgl_hack_b200 import dense
torch.manual_seed(0)
a = (torch.rand(5000, 100, device="cuda") * 2 - 1).requires_grad_(True)
w = (torch.rand(100, 256, device="cuda") * 2 - 1).requires_grad_(True)
g = torch.rand(5000, 256, device="cuda") * 2 - 1
y = dense.matmul(a, w); y.backward(g)
ga, gw = a.grad.double(), w.grad.double()
a64, w64, g64 = a.detach().double(), w.detach().double(), g.double()
def rel(x, y, s): return ((x - y).abs().max() / s).item()
e = [rel(y.detach().double(), a64 @ w64, (a64.abs() @ w64.abs()).max()),
     rel(ga, g64 @ w64.t(), (g64.abs() @ w64.abs().t()).max()),
     rel(gw, a64.t() @ g64, (a64.abs().t() @ g64.abs()).max())]
print(e); assert max(e) <= 1e-5, e
''' % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, (out.stdout[-1500:], out.stderr[-1500:])


def test_grouped_gemm_and_tensor_core_rgcn():
    """Grouped mode (ragged row groups, one weight each) vs fp64 per group; RGCN layer through the
    tensor-core path (segment sums -> grouped GEMM -> node reduce) vs the fused SIMT typed kernel
    at 16x16 and vs a float64 dense formula at 64x48, values and all gradients."""
    code = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
import dgl_hack_b200 as B
from dgl_hack_b200 import backend as F
torch.manual_seed(0)
dev = "cuda:0"
# ---- grouped GEMM: empty groups, partial tiles, a group larger than several tiles
sizes = [0, 5, 128, 1, 700, 0, 129, 37]
gp = [0] + list(np.cumsum(sizes))
a = torch.rand(gp[-1], 64, device=dev) * 2 - 1
w = torch.rand(len(sizes), 64, 48, device=dev) * 2 - 1
c = F.GroupedMatMul.apply(a, w, gp)
for g in range(len(sizes)):
    s, e = gp[g], gp[g + 1]
    if e > s:
        want = a[s:e].double() @ w[g].double()
        scale = (a[s:e].abs().double() @ w[g].abs().double()).max()
        assert ((c[s:e].double() - want).abs().max() / scale) <= 1e-5, g
# ---- RGCN layer
n, e, R = 3000, 40000, 23
rs = np.random.RandomState(1)
src, dst, ety = rs.randint(0, n, e), rs.randint(0, n, e), rs.randint(0, R, e)
dst[:3000] = 11
g = B.GraphIndex.from_coo_with_type(n, src, dst, ety, R, dev)
norm = torch.rand(e, 1, device=dev) + 0.1
def run(fn, fin, fout, seed):
    gen = torch.Generator(device=dev).manual_seed(seed)
    x = (torch.rand(n, fin, generator=gen, device=dev) * 2 - 1).requires_grad_(True)
    W = (torch.rand(R, fin, fout, generator=gen, device=dev) * 2 - 1).requires_grad_(True)
    go = torch.rand(n, fout, generator=gen, device=dev) * 2 - 1
    out = fn(x, W)
    out.backward(go)
    return out.detach(), x.grad, W.grad
tc = run(lambda x, W: F.rgcn_layer1_tc(g, x, W, norm), 16, 16, 5)
sm = run(lambda x, W: F.rgcn_layer1(g, x, W, norm), 16, 16, 5)
for p, q in zip(tc, sm):
    assert ((p - q).abs().max() / q.abs().max()) <= 1e-5, "tc vs fused SIMT"
S, D, T = (torch.from_numpy(v).to(dev) for v in (src, dst, ety))
def dense(x, W):
    m = torch.bmm(x.double()[S].unsqueeze(1), W.double()[T]).squeeze(1) * norm.double()
    return torch.zeros(n, W.shape[2], dtype=torch.float64, device=dev).index_add_(0, D, m)
tc = run(lambda x, W: F.rgcn_layer1_tc(g, x, W, norm), 64, 48, 6)
rf = run(lambda x, W: dense(x, W).float(), 64, 48, 6)
for p, q in zip(tc, rf):
    assert ((p.double() - q.double()).abs().max() / q.abs().max()) <= 2e-5, "tc vs float64"
print("ok")
''' % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=400)
    assert out.returncode == 0, (out.stdout[-1500:], out.stderr[-2500:])

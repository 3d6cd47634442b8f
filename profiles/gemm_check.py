"""Accuracy and time of the hand-written tcgen05 3xTF32 GEMM (csrc/gemm.cu, DGLB200DenseTF32x3)
against an fp64 product, torch's fp32 SIMT GEMM, plain TF32 and cuBLASLt BF16x9 (dense.py).

    python profiles/gemm_check.py [--quick] [--out gpurun_out/r2_gemm_check.json]

err = max |c - c64| / max(sum_k |a||b|) per shape (the scale north_star's 1e-5 is measured on).
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dgl_hack_b200 import dense  # noqa: E402

DEV = torch.device("cuda:0")


def timeit(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--wgrad", action="store_true",
                    help="also check the MN-major weight-gradient kernel (B200_DENSE_WGRAD=tc)")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    torch.backends.cuda.matmul.allow_tf32 = False
    shapes = [(50, 32, 16, True, False), (1000, 256, 256, True, True), (4097, 100, 256, True, False),
              (777, 256, 47, True, True), (3000, 48, 256, False, False), (129, 8, 8, False, True),
              (20000, 256, 100, False, False)]
    if not args.quick:
        shapes += [(1 << 20, 256, 256, True, False), (1 << 20, 100, 256, True, False),
                   (1 << 20, 256, 48, True, False), (2449029, 256, 256, True, False)]
    rows = []
    gen = torch.Generator(device=DEV).manual_seed(0)
    for M, K, N, kn, fuse in shapes:
        a = torch.rand(M, K, generator=gen, device=DEV) * 2 - 1
        b = torch.rand((K, N) if kn else (N, K), generator=gen, device=DEV) * 2 - 1
        bias = torch.rand(N, generator=gen, device=DEV) if fuse else None
        bm = b if kn else b.t()
        big = M >= (1 << 20)
        sl = slice(0, 4096) if big else slice(None)
        c64 = a[sl].double() @ bm.double()
        if fuse:
            c64 = torch.relu(c64 + bias.double())
        scale = (a[sl].abs().double() @ bm.abs().double()).max().item()
        got = dense.tc_gemm(a, b, kn, bias, fuse)
        torch.cuda.synchronize()
        assert got.shape == (M, N)
        err = (got[sl].double() - c64).abs().max().item() / scale
        ref32 = a[sl] @ bm
        if fuse:
            ref32 = torch.relu(ref32 + bias)
        err32 = (ref32.double() - c64).abs().max().item() / scale
        row = {"M": M, "K": K, "N": N, "b_is_kn": kn, "bias_relu": fuse,
               "err_tf32x3": err, "err_torch_fp32": err32}
        if big:
            tail = slice(M - 1000, M)     # the last, partial tile too
            c64t = a[tail].double() @ bm.double()
            row["err_tf32x3_tail"] = (got[tail].double() - c64t).abs().max().item() / scale
            row["ms_tf32x3"] = round(timeit(lambda: dense.tc_gemm(a, b, kn)), 4)
            row["ms_torch_fp32"] = round(timeit(lambda: a @ bm, reps=5), 4)
            torch.backends.cuda.matmul.allow_tf32 = True
            row["ms_torch_tf32"] = round(timeit(lambda: a @ bm, reps=5), 4)
            torch.backends.cuda.matmul.allow_tf32 = False
            try:
                row["ms_cublaslt_bf16x9"] = round(timeit(lambda: dense.gemm(a, bm.contiguous()), reps=5), 4)
            except Exception as e:  # toolkit cuBLASLt missing
                row["ms_cublaslt_bf16x9"] = "unavailable: %s" % str(e)[:80]
            row["tflops_effective"] = round(2.0 * M * K * N / row["ms_tf32x3"] / 1e9, 1)
        rows.append(row)
        print(json.dumps(row), flush=True)
        del a, b, got
    # ---- weight gradient  c[K, N] = a[M, K].T @ g[M, N]  (MN-major kernel, split over row chunks)
    wshapes = [(20, 8, 8), (5000, 100, 256), (4097, 256, 48), (100001, 256, 256), (300, 132, 20)]
    if not args.quick:
        wshapes += [(1 << 20, 256, 256), (1 << 20, 100, 256), (2449029, 256, 48)]
    if args.wgrad:
        os.environ["B200_DENSE_WGRAD"] = "tc"
    for M, K, N in (wshapes if args.wgrad else []):
        a = torch.rand(M, K, generator=gen, device=DEV) * 2 - 1
        gg = torch.rand(M, N, generator=gen, device=DEV) * 2 - 1
        got = dense.tc_weight_grad(a, gg)
        torch.cuda.synchronize()
        want = torch.zeros(K, N, dtype=torch.float64, device=DEV)
        scale = torch.zeros(K, N, dtype=torch.float64, device=DEV)
        for r0 in range(0, M, 1 << 18):      # fp64 in slabs (memory)
            sl = slice(r0, min(M, r0 + (1 << 18)))
            want += a[sl].double().t() @ gg[sl].double()
            scale += a[sl].abs().double().t() @ gg[sl].abs().double()
        err = ((got.double() - want).abs() / scale.clamp(min=1e-30)).max().item()
        err32 = (((a.t() @ gg).double() - want).abs() / scale.clamp(min=1e-30)).max().item()
        row = {"op": "weight_grad", "M": M, "K": K, "N": N, "err_tf32x3": err,
               "err_torch_fp32": err32}
        if M >= (1 << 20):
            row["ms_tf32x3"] = round(timeit(lambda: dense.tc_weight_grad(a, gg)), 4)
            row["ms_torch_fp32"] = round(timeit(lambda: a.t() @ gg, reps=5), 4)
        rows.append(row)
        print(json.dumps(row), flush=True)
        del a, gg
    worst = max(max(r["err_tf32x3"], r.get("err_tf32x3_tail", 0.0)) for r in rows)
    print(json.dumps({"worst_err_tf32x3": worst, "ok": worst <= 1e-5}))
    if args.out:
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        json.dump({"rows": rows, "worst_err_tf32x3": worst}, open(args.out, "w"), indent=1)
    sys.exit(0 if worst <= 1e-5 else 1)


if __name__ == "__main__":
    main()

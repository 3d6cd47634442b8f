"""Probe for DESIGN section 9 item 1 (dense projections on tensor cores at fp32 accuracy): the
system cuBLASLt 12.9 exposes CUBLAS_COMPUTE_32F_EMULATED_16BFX9 (fp32 GEMM emulated with nine BF16
tensor-core products); torch's bundled cuBLAS 12.8 does not. This script loads the SYSTEM
libcublasLt by absolute path, runs the GCN-epoch GEMM shapes with compute type 68 (fp32 SIMT
reference), 77 (TF32) and 78 (BF16x9), and reports time and error against an fp64 product.
Nothing in the library or the tests depends on it."""
import ctypes as C
import glob
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = torch.device("cuda:0")
CUDA_R_32F, OP_N, PREF_WS = 0, 0, 1


class Heur(C.Structure):
    _fields_ = [("algo", C.c_uint64 * 8), ("ws", C.c_size_t), ("state", C.c_int),
                ("waves", C.c_float), ("reserved", C.c_int * 4)]


def load():
    cands = sorted(glob.glob("/usr/local/cuda/lib64/libcublasLt.so.12.*"))
    lt = C.CDLL(cands[-1] if cands else "/usr/local/cuda/lib64/libcublasLt.so.12")
    lt.cublasLtCreate.argtypes = [C.POINTER(C.c_void_p)]
    lt.cublasLtMatmulDescCreate.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int]
    lt.cublasLtMatrixLayoutCreate.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_uint64,
                                              C.c_uint64, C.c_int64]
    lt.cublasLtMatmulPreferenceCreate.argtypes = [C.POINTER(C.c_void_p)]
    lt.cublasLtMatmulPreferenceSetAttribute.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
    lt.cublasLtMatmulAlgoGetHeuristic.argtypes = [C.c_void_p] * 7 + [C.c_int, C.POINTER(Heur),
                                                                     C.POINTER(C.c_int)]
    lt.cublasLtMatmul.argtypes = [C.c_void_p] * 13 + [C.c_void_p, C.c_size_t, C.c_void_p]
    lt.cublasLtGetVersion.restype = C.c_size_t
    return lt


def layout(lt, rows, cols, ld):
    h = C.c_void_p()
    assert lt.cublasLtMatrixLayoutCreate(C.byref(h), CUDA_R_32F, rows, cols, ld) == 0
    return h


def gemm_fn(lt, handle, compute, a, b, c, ws):
    """c[M,N] = a[M,K] @ b[K,N], all row-major: column-major C^T = B^T A^T."""
    M, K = a.shape
    N = b.shape[1]
    desc = C.c_void_p()
    st = lt.cublasLtMatmulDescCreate(C.byref(desc), compute, CUDA_R_32F)
    if st:
        return None, "MatmulDescCreate status %d" % st
    la, lb, lc = layout(lt, N, K, N), layout(lt, K, M, K), layout(lt, N, M, N)
    pref = C.c_void_p()
    lt.cublasLtMatmulPreferenceCreate(C.byref(pref))
    wsz = C.c_size_t(ws.numel())
    lt.cublasLtMatmulPreferenceSetAttribute(pref, PREF_WS, C.byref(wsz), 8)
    res, got = Heur(), C.c_int(0)
    st = lt.cublasLtMatmulAlgoGetHeuristic(handle, desc, la, lb, lc, lc, pref, 1, C.byref(res),
                                           C.byref(got))
    if st or got.value == 0 or res.state:
        return None, "heuristic status %d, found %d, state %d" % (st, got.value, res.state)
    alpha, beta = C.c_float(1.0), C.c_float(0.0)
    stream = C.c_void_p(torch.cuda.current_stream(DEV).cuda_stream)

    def run():
        st = lt.cublasLtMatmul(handle, desc, C.byref(alpha), b.data_ptr(), la, a.data_ptr(), lb,
                               C.byref(beta), c.data_ptr(), lc, c.data_ptr(), lc,
                               C.byref(res.algo), ws.data_ptr(), res.ws, stream)
        if st:
            raise RuntimeError("cublasLtMatmul status %d" % st)
    return run, "workspace %d B, waves %.2f" % (res.ws, res.waves)


def timeit(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    out_path = os.path.join(ROOT, "gpurun_out", "r1_bf16x9_probe.json")
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    rows = []
    torch.zeros(1, device=DEV)
    lt = load()
    rows.append({"cublasLt_version": int(lt.cublasLtGetVersion())})
    handle = C.c_void_p()
    assert lt.cublasLtCreate(C.byref(handle)) == 0
    ws = torch.empty(512 << 20, dtype=torch.uint8, device=DEV)
    gen = torch.Generator(device=DEV)
    gen.manual_seed(0)
    for M, K, N in ((1 << 20, 256, 256), (1 << 20, 100, 256), (1 << 20, 256, 48)):
        a = torch.rand(M, K, device=DEV, generator=gen) * 2 - 1
        b = torch.rand(K, N, device=DEV, generator=gen) * 2 - 1
        ref = (a[:8192].double() @ b.double())
        scale = float(ref.abs().max())
        c = torch.empty(M, N, device=DEV)
        ms = timeit(lambda: torch.matmul(a, b, out=c))
        err = float((c[:8192].double() - ref).abs().max()) / scale
        rows.append({"shape": [M, K, N], "impl": "torch.matmul fp32 (bundled cuBLAS)",
                     "ms": round(ms, 3), "tflops": round(2e-9 * M * K * N / ms, 1),
                     "max_err_over_max": err})
        print(json.dumps(rows[-1]), flush=True)
        for name, compute in (("cublasLt 12.9 COMPUTE_32F", 68), ("cublasLt 12.9 FAST_TF32", 77),
                              ("cublasLt 12.9 EMULATED_16BFX9", 78)):
            row = {"shape": [M, K, N], "impl": name}
            try:
                fn, note = gemm_fn(lt, handle, compute, a, b, c, ws)
                row["note"] = note
                if fn is not None:
                    c.zero_()
                    ms = timeit(fn)
                    row.update(ms=round(ms, 3), tflops=round(2e-9 * M * K * N / ms, 1),
                               max_err_over_max=float((c[:8192].double() - ref).abs().max()) / scale)
            except Exception as ex:  # a probe: report, do not fail
                row["error"] = str(ex)[:200]
            rows.append(row)
            print(json.dumps(row), flush=True)
            json.dump(rows, open(out_path, "w"), indent=1)
    json.dump(rows, open(out_path, "w"), indent=1)


if __name__ == "__main__":
    sys.exit(main())

"""Opt-in dense projections at fp32 accuracy on tensor cores.

The GEMMs either side of the message-passing kernels (GraphConv's `feat @ weight`,
graphconv.py:144-152) are plain library GEMMs. torch's bundled cuBLAS 12.8 runs them as fp32 SIMT
kernels; the CUDA 12.9 toolkit's cuBLASLt has `CUBLAS_COMPUTE_32F_EMULATED_16BFX9`, an fp32 GEMM
emulated with nine BF16 tensor-core products whose error is BELOW the SIMT kernel's
(profiles/r1_bf16x9_probe.json). This module loads that library by absolute path (the bundled one
shares its SONAME, so it must not be found by name), and exposes

    matmul(a, b)            autograd-aware a @ b for 2-D fp32 CUDA tensors
    gemm(a, b, ta, tb)      op(a) @ op(b), row-major, no autograd

It is off by default: set B200_DENSE=bf16x9 (read by nn/conv.py) to route GraphConv through it.
Shapes for which the emulated kernels are not offered run cuBLASLt's fp32 compute type.
"""
import ctypes as C
import glob
import os

import torch

from ._lib import DGLError

_CUDA_R_32F = 0
_COMPUTE_32F, _COMPUTE_BF16X9 = 68, 78
_DESC_TRANSA, _DESC_TRANSB, _PREF_WS = 3, 4, 1
_WORKSPACE = 32 << 20


class _Heur(C.Structure):  # cublasLtMatmulHeuristicResult_t (cublasLt.h)
    _fields_ = [("algo", C.c_uint64 * 8), ("ws", C.c_size_t), ("state", C.c_int),
                ("waves", C.c_float), ("reserved", C.c_int * 4)]


class _Lt(object):
    _inst = None

    def __init__(self):
        cands = sorted(glob.glob(os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"),
                                              "lib64", "libcublasLt.so.12.*")))
        if not cands:
            raise DGLError("B200_DENSE=bf16x9 needs the CUDA 12.9 toolkit's libcublasLt "
                           "(not found under $CUDA_HOME/lib64)")
        lt = C.CDLL(cands[-1])
        lt.cublasLtGetVersion.restype = C.c_size_t
        if lt.cublasLtGetVersion() < 120900:
            raise DGLError("libcublasLt %d has no emulated fp32 compute type" % lt.cublasLtGetVersion())
        lt.cublasLtCreate.argtypes = [C.POINTER(C.c_void_p)]
        lt.cublasLtMatmulDescCreate.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int]
        lt.cublasLtMatmulDescSetAttribute.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        lt.cublasLtMatrixLayoutCreate.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_uint64,
                                                  C.c_uint64, C.c_int64]
        lt.cublasLtMatmulPreferenceCreate.argtypes = [C.POINTER(C.c_void_p)]
        lt.cublasLtMatmulPreferenceSetAttribute.argtypes = [C.c_void_p, C.c_int, C.c_void_p,
                                                            C.c_size_t]
        lt.cublasLtMatmulAlgoGetHeuristic.argtypes = [C.c_void_p] * 7 + [
            C.c_int, C.POINTER(_Heur), C.POINTER(C.c_int)]
        lt.cublasLtMatmul.argtypes = [C.c_void_p] * 13 + [C.c_void_p, C.c_size_t, C.c_void_p]
        self.lt = lt
        self.handles, self.ws, self.plans = {}, {}, {}
        self.alpha, self.beta = C.c_float(1.0), C.c_float(0.0)

    @classmethod
    def get(cls):
        if cls._inst is None:
            cls._inst = cls()
        return cls._inst

    def _check(self, st, what):
        if st:
            raise DGLError("cuBLASLt %s failed with status %d" % (what, st))

    def _layout(self, rows, cols, ld):
        h = C.c_void_p()
        self._check(self.lt.cublasLtMatrixLayoutCreate(C.byref(h), _CUDA_R_32F, rows, cols, ld),
                    "MatrixLayoutCreate")
        return h

    def plan(self, dev, M, K, N, ta, tb):
        """Descriptor + layouts + algorithm for row-major c[M,N] = op(a) @ op(b), cached. In
        cuBLASLt's column-major terms: C^T[N,M] = op(b)^T . op(a)^T, first operand = b's buffer."""
        key = (dev, M, K, N, ta, tb)
        if key in self.plans:
            return self.plans[key]
        lt = self.lt
        if dev not in self.handles:
            h = C.c_void_p()
            self._check(lt.cublasLtCreate(C.byref(h)), "Create")
            self.handles[dev] = h
            self.ws[dev] = torch.empty(_WORKSPACE, dtype=torch.uint8, device="cuda:%d" % dev)
        # b row-major [K,N] is column-major N x K (ld N); stored transposed [N,K] it is K x N (ld K)
        la = self._layout(K, N, K) if tb else self._layout(N, K, N)
        # a row-major [M,K] is column-major K x M (ld K); stored transposed [K,M] it is M x K (ld M)
        lb = self._layout(M, K, M) if ta else self._layout(K, M, K)
        lc = self._layout(N, M, N)
        pref = C.c_void_p()
        self._check(lt.cublasLtMatmulPreferenceCreate(C.byref(pref)), "PreferenceCreate")
        wsz = C.c_size_t(_WORKSPACE)
        lt.cublasLtMatmulPreferenceSetAttribute(pref, _PREF_WS, C.byref(wsz), 8)
        chosen = None
        for compute in (_COMPUTE_BF16X9, _COMPUTE_32F):
            desc = C.c_void_p()
            if lt.cublasLtMatmulDescCreate(C.byref(desc), compute, _CUDA_R_32F):
                continue
            for attr, flag in ((_DESC_TRANSA, tb), (_DESC_TRANSB, ta)):
                op = C.c_int32(1 if flag else 0)
                self._check(lt.cublasLtMatmulDescSetAttribute(desc, attr, C.byref(op), 4),
                            "MatmulDescSetAttribute")
            res, got = _Heur(), C.c_int(0)
            st = lt.cublasLtMatmulAlgoGetHeuristic(self.handles[dev], desc, la, lb, lc, lc, pref,
                                                   1, C.byref(res), C.byref(got))
            if st == 0 and got.value > 0 and res.state == 0:
                chosen = (desc, la, lb, lc, res, compute)
                break
        if chosen is None:
            raise DGLError("cuBLASLt offers no fp32 algorithm for [%d,%d] x [%d,%d]" % (M, K, K, N))
        self.plans[key] = chosen
        return chosen

    def run(self, a, b, c, M, K, N, ta, tb):
        dev = a.device.index or 0
        desc, la, lb, lc, res, _ = self.plan(dev, M, K, N, ta, tb)
        with torch.cuda.device(dev):
            st = self.lt.cublasLtMatmul(
                self.handles[dev], desc, C.byref(self.alpha), b.data_ptr(), la, a.data_ptr(), lb,
                C.byref(self.beta), c.data_ptr(), lc, c.data_ptr(), lc, C.byref(res.algo),
                self.ws[dev].data_ptr(), res.ws,
                C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
        self._check(st, "Matmul")


def gemm(a, b, trans_a=False, trans_b=False):
    """op(a) @ op(b) for 2-D fp32 CUDA tensors (made contiguous), fp32 result, no autograd."""
    if a.dim() != 2 or b.dim() != 2 or a.dtype != torch.float32 or b.dtype != torch.float32:
        raise DGLError("dense.gemm takes 2-D float32 tensors")
    if not (a.is_cuda and b.is_cuda and a.device == b.device):
        raise DGLError("dense.gemm: operands must live on one CUDA device (there is no CPU path)")
    a, b = a.contiguous(), b.contiguous()
    M, K = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
    Kb, N = (b.shape[1], b.shape[0]) if trans_b else (b.shape[0], b.shape[1])
    if K != Kb:
        raise DGLError("dense.gemm: inner dimensions differ (%d vs %d)" % (K, Kb))
    c = torch.empty((M, N), dtype=torch.float32, device=a.device)
    if M == 0 or N == 0:
        return c
    if K == 0:
        return c.zero_()
    _Lt.get().run(a, b, c, M, K, N, bool(trans_a), bool(trans_b))
    return c


class _MatMul(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return gemm(a, b)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        ga = gemm(g, b, trans_b=True) if ctx.needs_input_grad[0] else None
        gb = gemm(a, g, trans_a=True) if ctx.needs_input_grad[1] else None
        return ga, gb


def tc_supported(m, k, n, a):
    """Shapes the library's own tensor-core GEMM (csrc/gemm.cu) takes."""
    return n <= 256 and k % 4 == 0 and k > 0 and m > 0 and a.data_ptr() % 16 == 0


def tc_gemm(a, b, b_is_kn=True, bias=None, relu=False):
    """a @ (b if b_is_kn else b.T) (+ bias) (ReLU): the hand-written tcgen05 3xTF32 kernel
    (DGLB200DenseTF32x3); no autograd. Falls back to torch for shapes it does not take."""
    from . import kernel as K
    a, b = a.contiguous(), b.contiguous()
    n = b.shape[1] if b_is_kn else b.shape[0]
    if not tc_supported(a.shape[0], a.shape[1], n, a):
        c = a @ (b if b_is_kn else b.t())
        if bias is not None:
            c = c + bias
        return torch.relu(c) if relu else c
    c = torch.empty((a.shape[0], n), dtype=torch.float32, device=a.device)
    K.dense_tf32x3(a, b, b_is_kn, bias, relu, c)
    return c


def tc_weight_grad(a, g):
    """a.T @ g ([M, K], [M, N] -> [K, N]): the own MN-major tcgen05 kernel (csrc/gemm.cu
    gemm_tn_tf32x3_kernel: 0.64 ms against 3.5 ms for torch's fp32 GEMM at [1M, 256] x [1M, 256],
    error 3e-7 of sum|terms|, profiles/r2_gemm_check_wgrad.json); B200_DENSE_WGRAD=lib = torch."""
    from . import kernel as K
    a, g = a.contiguous(), g.contiguous()
    k, n = a.shape[1], g.shape[1]
    if os.environ.get("B200_DENSE_WGRAD", "tc") != "tc" or k > 256 or n > 256 or k % 4 or n % 4 \
            or a.shape[0] == 0 or a.data_ptr() % 16 or g.data_ptr() % 16:
        return a.t() @ g
    c = torch.empty((k, n), dtype=torch.float32, device=a.device)
    K.dense_tf32x3_weight_grad(a, g, c)
    return c


class _TcMatMul(torch.autograd.Function):
    """a @ b on the own kernel; grad_a = g @ b.T on the same kernel (b read as [N', K']);
    grad_b = a.T @ g reduces over the long dimension: the MN-major kernel (tc_weight_grad)."""

    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return tc_gemm(a, b, True)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        ga = tc_gemm(g.contiguous(), b, False) if ctx.needs_input_grad[0] else None
        gb = tc_weight_grad(a, g) if ctx.needs_input_grad[1] else None
        return ga, gb


class _TcLinear(torch.autograd.Function):
    """relu?(a @ w + bias) with the bias / ReLU applied while the accumulator is read out of
    TMEM (one pass over the output instead of three)."""

    @staticmethod
    def forward(ctx, a, w, bias, relu):
        y = tc_gemm(a, w, True, bias, relu)
        ctx.relu = relu
        ctx.save_for_backward(a, w, y if relu else None)
        return y

    @staticmethod
    def backward(ctx, g):
        a, w, y = ctx.saved_tensors
        g = g.contiguous()
        if ctx.relu:
            # one pass (read g, y; write) instead of a compare kernel and a multiply
            g = torch.ops.aten.threshold_backward(g, y, 0.0)
        ga = tc_gemm(g, w, False) if ctx.needs_input_grad[0] else None
        gw = tc_weight_grad(a, g) if ctx.needs_input_grad[1] else None
        gb = g.sum(0) if ctx.needs_input_grad[2] else None
        return ga, gw, gb, None


def linear(a, w, bias=None, relu=False):
    """relu?(a @ w + bias) with gradients. B200_DENSE=tf32x3: one fused tensor-core kernel
    (csrc/gemm.cu); otherwise the composition of matmul / add / relu."""
    if mode() == "tf32x3" and a.dim() == 2 and a.is_cuda and a.dtype == torch.float32:
        return _TcLinear.apply(a, w, bias, bool(relu))
    y = matmul(a, w) if enabled() else a @ w
    if bias is not None:
        y = y + bias
    return torch.relu(y) if relu else y


def mode():
    """B200_DENSE: "tf32x3" (default) = the library's own tcgen05 kernel; "bf16x9" = cuBLASLt's
    emulated-fp32 GEMM (library baseline); "torch" = torch.matmul (cuBLAS fp32 SIMT)."""
    return os.environ.get("B200_DENSE", "tf32x3")


def matmul(a, b):
    """a @ b with gradients on tensor cores at fp32 accuracy: B200_DENSE=tf32x3 = the library's
    own tcgen05 kernel, B200_DENSE=bf16x9 = cuBLASLt's emulated-fp32 GEMM (a baseline)."""
    if mode() == "tf32x3":
        return _TcMatMul.apply(a, b)
    return _MatMul.apply(a, b)


def enabled():
    return mode() in ("bf16x9", "tf32x3")

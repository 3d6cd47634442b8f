// GAT attention projections, forward and backward (extension; SURVEY 8(f) rank 2 "fusing the step
// either side": gatconv.py:143-144 computes el = (feat_src * attn_l).sum(-1), er = (feat_dst *
// attn_r).sum(-1) with torch elementwise + reduce kernels that stream the [N, H, F] projection
// four times forward and eight times backward).
//
//   forward : out_k[n,h] = sum_f feat[n,h,f] * attn_k[h,f]            k in {l, r}, ONE pass over feat
//   backward: grad_feat[n,h,f] = sum_k g_k[n,h] * attn_k[h,f]         ONE pass over feat, ONE write
//             grad_attn_k[h,f] = sum_n g_k[n,h] * feat[n,h,f]         block partials + ordered fold
//
// Both kernels are pure streaming (HBM bound): forward reads 4*N*D bytes, backward reads 4*N*D and
// writes 4*N*D. No atomics: the reduction over nodes is two-stage with a fixed fold order.
#include <algorithm>

#include <mutex>
#include <string>

#include "common.cuh"

namespace b200 {
namespace {

// One thread per (node, head): F contiguous floats, float4 loads when F % 4 == 0 and aligned.
template <int VEC>
__global__ void __launch_bounds__(256) attn_proj_fwd_kernel(const float* __restrict__ feat,
                                                            const float* __restrict__ al,
                                                            const float* __restrict__ ar,
                                                            float* __restrict__ el,
                                                            float* __restrict__ er, int64_t n_pairs,
                                                            int H, int F) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pairs; p += stride) {
    const int h = (int)(p % H);
    const float* x = feat + p * F;
    const float* wl = al ? al + (int64_t)h * F : nullptr;
    const float* wr = ar ? ar + (int64_t)h * F : nullptr;
    float sl = 0.f, sr = 0.f;
    if (VEC == 4) {
      for (int f = 0; f < F; f += 4) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + f));
        if (wl) {
          const float4 w = __ldg(reinterpret_cast<const float4*>(wl + f));
          sl = __fadd_rn(sl, __fmul_rn(v.x, w.x)); sl = __fadd_rn(sl, __fmul_rn(v.y, w.y));
          sl = __fadd_rn(sl, __fmul_rn(v.z, w.z)); sl = __fadd_rn(sl, __fmul_rn(v.w, w.w));
        }
        if (wr) {
          const float4 w = __ldg(reinterpret_cast<const float4*>(wr + f));
          sr = __fadd_rn(sr, __fmul_rn(v.x, w.x)); sr = __fadd_rn(sr, __fmul_rn(v.y, w.y));
          sr = __fadd_rn(sr, __fmul_rn(v.z, w.z)); sr = __fadd_rn(sr, __fmul_rn(v.w, w.w));
        }
      }
    } else {
      for (int f = 0; f < F; ++f) {
        const float v = __ldg(x + f);
        if (wl) sl = __fadd_rn(sl, __fmul_rn(v, __ldg(wl + f)));
        if (wr) sr = __fadd_rn(sr, __fmul_rn(v, __ldg(wr + f)));
      }
    }
    if (el) el[p] = sl;
    if (er) er[p] = sr;
  }
}

// Backward: a block owns a contiguous range of nodes; thread t owns the columns t, t + blockDim, ..
// of the [H*F] row (coalesced row reads / writes), keeps the attn-gradient partial sums of its
// columns in registers over the whole range and stores them once to part[block][k][D].
constexpr int kBwdThreads = 128;
constexpr int kBwdCols = 8;  // columns per thread: D <= 1024

__global__ void __launch_bounds__(kBwdThreads) attn_proj_bwd_kernel(
    const float* __restrict__ feat, const float* __restrict__ al, const float* __restrict__ ar,
    const float* __restrict__ gl, const float* __restrict__ gr, float* __restrict__ gfeat,
    float* __restrict__ part, int64_t n_nodes, int rows_per_block, int H, int F) {
  const int D = H * F;
  float pl[kBwdCols], pr[kBwdCols], wl[kBwdCols], wr[kBwdCols];
  int hd[kBwdCols];
#pragma unroll
  for (int c = 0; c < kBwdCols; ++c) {
    const int col = threadIdx.x + c * kBwdThreads;
    pl[c] = pr[c] = 0.f;
    hd[c] = col < D ? col / F : 0;
    wl[c] = (al && col < D) ? __ldg(al + col) : 0.f;
    wr[c] = (ar && col < D) ? __ldg(ar + col) : 0.f;
  }
  const int64_t beg = (int64_t)blockIdx.x * rows_per_block;
  const int64_t end = min(beg + rows_per_block, n_nodes);
#pragma unroll 4
  for (int64_t n = beg; n < end; ++n) {
#pragma unroll
    for (int c = 0; c < kBwdCols; ++c) {
      const int col = threadIdx.x + c * kBwdThreads;
      if (col < D) {
        const float x = __ldg(feat + n * D + col);
        const float a = gl ? __ldg(gl + n * H + hd[c]) : 0.f;
        const float b = gr ? __ldg(gr + n * H + hd[c]) : 0.f;
        pl[c] = __fadd_rn(pl[c], __fmul_rn(a, x));
        pr[c] = __fadd_rn(pr[c], __fmul_rn(b, x));
        gfeat[n * D + col] = __fadd_rn(__fmul_rn(a, wl[c]), __fmul_rn(b, wr[c]));
      }
    }
  }
#pragma unroll
  for (int c = 0; c < kBwdCols; ++c) {
    const int col = threadIdx.x + c * kBwdThreads;
    if (col < D) {
      part[((int64_t)blockIdx.x * 2 + 0) * D + col] = pl[c];
      part[((int64_t)blockIdx.x * 2 + 1) * D + col] = pr[c];
    }
  }
}

// Vector form (F % 4 == 0): a thread owns ONE float4 column group and every (256 / LPR)-th row of
// the block's range, LPR = D / 4 lanes per row; 16-byte loads / stores, four rows in flight per
// thread. The row groups of a block are folded through shared memory in a fixed order.
constexpr int kBwdVThreads = 256;

__global__ void __launch_bounds__(kBwdVThreads) attn_proj_bwd_vec_kernel(
    const float* __restrict__ feat, const float* __restrict__ al, const float* __restrict__ ar,
    const float* __restrict__ gl, const float* __restrict__ gr, float* __restrict__ gfeat,
    float* __restrict__ part, int64_t n_nodes, int rows_per_block, int H, int F) {
  extern __shared__ float4 sh[];  // [2][groups][LPR]
  const int D = H * F, lpr = D >> 2;
  const int groups = kBwdVThreads / lpr;
  const int grp = threadIdx.x / lpr, cg = threadIdx.x - grp * lpr;
  const bool active = grp < groups;
  const int head = (cg << 2) / F;
  float4 wl = make_float4(0.f, 0.f, 0.f, 0.f), wr = wl, pl = wl, pr = wl;
  if (active && al) wl = __ldg(reinterpret_cast<const float4*>(al) + cg);
  if (active && ar) wr = __ldg(reinterpret_cast<const float4*>(ar) + cg);
  const int64_t beg = (int64_t)blockIdx.x * rows_per_block;
  const int64_t end = min(beg + rows_per_block, n_nodes);
  if (active) {
#pragma unroll 4
    for (int64_t n = beg + grp; n < end; n += groups) {
      const float4 x = __ldg(reinterpret_cast<const float4*>(feat + n * D) + cg);
      const float a = gl ? __ldg(gl + n * H + head) : 0.f;
      const float b = gr ? __ldg(gr + n * H + head) : 0.f;
      pl.x = __fadd_rn(pl.x, __fmul_rn(a, x.x)); pl.y = __fadd_rn(pl.y, __fmul_rn(a, x.y));
      pl.z = __fadd_rn(pl.z, __fmul_rn(a, x.z)); pl.w = __fadd_rn(pl.w, __fmul_rn(a, x.w));
      pr.x = __fadd_rn(pr.x, __fmul_rn(b, x.x)); pr.y = __fadd_rn(pr.y, __fmul_rn(b, x.y));
      pr.z = __fadd_rn(pr.z, __fmul_rn(b, x.z)); pr.w = __fadd_rn(pr.w, __fmul_rn(b, x.w));
      float4 g;
      g.x = __fadd_rn(__fmul_rn(a, wl.x), __fmul_rn(b, wr.x));
      g.y = __fadd_rn(__fmul_rn(a, wl.y), __fmul_rn(b, wr.y));
      g.z = __fadd_rn(__fmul_rn(a, wl.z), __fmul_rn(b, wr.z));
      g.w = __fadd_rn(__fmul_rn(a, wl.w), __fmul_rn(b, wr.w));
      reinterpret_cast<float4*>(gfeat + n * D)[cg] = g;
    }
    sh[(0 * groups + grp) * lpr + cg] = pl;
    sh[(1 * groups + grp) * lpr + cg] = pr;
  }
  __syncthreads();
  if (threadIdx.x < 2 * lpr) {
    const int k = threadIdx.x / lpr, c = threadIdx.x - k * lpr;
    float4 acc = sh[(k * groups) * lpr + c];
    for (int q = 1; q < groups; ++q) {
      const float4 v = sh[(k * groups + q) * lpr + c];
      acc.x = __fadd_rn(acc.x, v.x); acc.y = __fadd_rn(acc.y, v.y);
      acc.z = __fadd_rn(acc.z, v.z); acc.w = __fadd_rn(acc.w, v.w);
    }
    reinterpret_cast<float4*>(part + ((int64_t)blockIdx.x * 2 + k) * D)[c] = acc;
  }
}

// grad_attn_k[col] = sum of the block partials in a FIXED order: 32 row groups x 32 columns per
// CTA, row group q folds blocks q, q + 32, ... sequentially (independent loads), then the 32 group
// sums are folded in group order through shared memory. (The first version walked all partials
// with one thread per column: 295 us for 2,368 partials -- more than the streaming kernel itself.)
__global__ void __launch_bounds__(1024) attn_proj_fold_kernel(const float* __restrict__ part,
                                                              float* __restrict__ gal,
                                                              float* __restrict__ gar, int n_blocks,
                                                              int D) {
  __shared__ float sh[2][32][33];
  const int c = threadIdx.x & 31, q = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + c;
  float sl = 0.f, sr = 0.f;
  if (col < D) {
#pragma unroll 4
    for (int b = q; b < n_blocks; b += 32) {
      sl = __fadd_rn(sl, __ldg(part + ((int64_t)b * 2 + 0) * D + col));
      sr = __fadd_rn(sr, __ldg(part + ((int64_t)b * 2 + 1) * D + col));
    }
  }
  sh[0][q][c] = sl;
  sh[1][q][c] = sr;
  __syncthreads();
  if (q < 2 && col < D) {
    float acc = 0.f;
    for (int k = 0; k < 32; ++k) acc = __fadd_rn(acc, sh[q][k][c]);
    float* dst = q == 0 ? gal : gar;
    if (dst) dst[col] = acc;
  }
}

int sm_count() {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms;
}

}  // namespace

void attn_proj_forward(const float* feat, const float* al, const float* ar, float* el, float* er,
                       int64_t n, int64_t H, int64_t F, cudaStream_t s) {
  const int64_t pairs = n * H;
  if (pairs == 0) return;
  const int blocks = (int)std::min<int64_t>((pairs + 255) / 256, (int64_t)sm_count() * 32);
  const bool v4 = F % 4 == 0 && reinterpret_cast<uintptr_t>(feat) % 16 == 0 &&
                  (!al || reinterpret_cast<uintptr_t>(al) % 16 == 0) &&
                  (!ar || reinterpret_cast<uintptr_t>(ar) % 16 == 0);
  if (v4) attn_proj_fwd_kernel<4><<<blocks, 256, 0, s>>>(feat, al, ar, el, er, pairs, (int)H, (int)F);
  else attn_proj_fwd_kernel<1><<<blocks, 256, 0, s>>>(feat, al, ar, el, er, pairs, (int)H, (int)F);
  B200_CUDA(cudaGetLastError());
  note_kernel("attn_proj_fwd", 1);
}

void attn_proj_backward(const float* feat, const float* al, const float* ar, const float* gl,
                        const float* gr, float* gfeat, float* gal, float* gar, int64_t n,
                        int64_t H, int64_t F, cudaStream_t s) {
  const int64_t D = H * F;
  B200_CHECK(D <= kBwdThreads * kBwdCols, "attn_proj: H * F above 1024 is not supported");
  int n_blocks = (int)std::min<int64_t>(std::max<int64_t>(n, 1), (int64_t)sm_count() * 5);
  const int rows = (int)((n + n_blocks - 1) / std::max(n_blocks, 1));
  n_blocks = rows > 0 ? (int)((n + rows - 1) / rows) : 0;
  Scratch part((size_t)std::max(n_blocks, 1) * 2 * D * sizeof(float), s);
  auto al16 = [](const void* p) { return !p || reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  const bool vec = F % 4 == 0 && D / 4 <= kBwdVThreads / 2 && al16(feat) && al16(al) && al16(ar) &&
                   al16(gfeat);
  if (n_blocks > 0 && vec) {
    const int lpr = (int)(D / 4), groups = kBwdVThreads / lpr;
    const size_t smem = (size_t)2 * groups * lpr * sizeof(float4);
    attn_proj_bwd_vec_kernel<<<n_blocks, kBwdVThreads, smem, s>>>(
        feat, al, ar, gl, gr, gfeat, part.as<float>(), n, rows, (int)H, (int)F);
    B200_CUDA(cudaGetLastError());
  } else if (n_blocks > 0) {
    attn_proj_bwd_kernel<<<n_blocks, kBwdThreads, 0, s>>>(feat, al, ar, gl, gr, gfeat,
                                                          part.as<float>(), n, rows, (int)H, (int)F);
    B200_CUDA(cudaGetLastError());
  }
  attn_proj_fold_kernel<<<(int)((D + 31) / 32), 1024, 0, s>>>(part.as<float>(), gal, gar, n_blocks,
                                                              (int)D);
  B200_CUDA(cudaGetLastError());
  note_kernel("attn_proj_bwd", 2);
}

// ---------------------------------------------------------------- no-allocation variant
// attn_proj_backward for callers that may not allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc): the block partials live in a file-scope device array instead
// of a stream-ordered Scratch buffer. Limits: H * F <= kStaticMaxD and at most kStaticMaxBlocks
// row blocks (the B200's 148 SMs x 5 = 740 fit, so the blocking -- and with it every sum order --
// is the one attn_proj_backward uses there: results are bit-identical). The array is
// kStaticMaxBlocks x 2 x kStaticMaxD floats = 3 MiB, held by every GPU that loads the library.
// ALL calls on a device share it: two calls that overlap on the GPU (different streams) race.
// Nothing has to start at zero: the backward kernel overwrites every partial the fold reads.
constexpr int kStaticMaxBlocks = 768;
constexpr int kStaticMaxD = 512;
static __device__ __align__(256) float g_attn_static_part[(size_t)kStaticMaxBlocks * 2 * kStaticMaxD];

constexpr int kMaxDevices = 64;
static float* g_static_part_addr[kMaxDevices];  // device address of the array, per device
static std::mutex g_static_mu;

// Device address of the array on `device` (which must be current). In host code the array's name
// is only a host-side stand-in for the symbol: the kernels get the address cudaGetSymbolAddress
// returns, looked up once per device.
static float* static_part_on(int device) {
  B200_CHECK(device >= 0 && device < kMaxDevices, "attn_proj (no-alloc): device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_part_addr[device]) {
    float* ptr = nullptr;
    B200_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_attn_static_part));
    B200_CHECK(ptr != nullptr, "attn_proj (no-alloc): cudaGetSymbolAddress returned NULL");
    g_static_part_addr[device] = ptr;
  }
  return g_static_part_addr[device];
}

int64_t attn_proj_static_max_width() { return kStaticMaxD; }

void attn_proj_backward_static(const float* feat, const float* al, const float* ar,
                               const float* gl, const float* gr, float* gfeat, float* gal,
                               float* gar, int64_t n, int64_t H, int64_t F, int device,
                               cudaStream_t s) {
  const int64_t D = H * F;
  B200_CHECK(D >= 1 && D <= kStaticMaxD,
             "attn_proj (no-alloc): H * F = " + std::to_string(D) + " exceeds the static scratch (" +
                 std::to_string(kStaticMaxD) + " floats per row); use DGLB200GatAttnProjBackward");
  int n_blocks = (int)std::min<int64_t>(std::max<int64_t>(n, 1), (int64_t)sm_count() * 5);
  n_blocks = std::min(n_blocks, kStaticMaxBlocks);
  const int rows = (int)((n + n_blocks - 1) / std::max(n_blocks, 1));
  n_blocks = rows > 0 ? (int)((n + rows - 1) / rows) : 0;
  float* part = static_part_on(device);
  auto al16 = [](const void* p) { return !p || reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  const bool vec = F % 4 == 0 && D / 4 <= kBwdVThreads / 2 && al16(feat) && al16(al) && al16(ar) &&
                   al16(gfeat);
  if (n_blocks > 0 && vec) {
    const int lpr = (int)(D / 4), groups = kBwdVThreads / lpr;
    const size_t smem = (size_t)2 * groups * lpr * sizeof(float4);
    attn_proj_bwd_vec_kernel<<<n_blocks, kBwdVThreads, smem, s>>>(feat, al, ar, gl, gr, gfeat, part,
                                                                  n, rows, (int)H, (int)F);
    B200_CUDA(cudaGetLastError());
  } else if (n_blocks > 0) {
    attn_proj_bwd_kernel<<<n_blocks, kBwdThreads, 0, s>>>(feat, al, ar, gl, gr, gfeat, part, n, rows,
                                                          (int)H, (int)F);
    B200_CUDA(cudaGetLastError());
  }
  attn_proj_fold_kernel<<<(int)((D + 31) / 32), 1024, 0, s>>>(part, gal, gar, n_blocks, (int)D);
  B200_CUDA(cudaGetLastError());
  note_kernel("attn_proj_bwd_static", 2);
}

}  // namespace b200

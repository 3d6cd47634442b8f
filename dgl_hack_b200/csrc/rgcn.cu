// RGCN "hack" layers over the typed adjacency. Replaces
//   RgcnLayer0KernelImpl / RgcnLayer0BackwardKernelImpl   (src/kernel/cuda/binary_reduce_impl.cu:913-933, 982-1008)
//   RgcnLayer1KernelImpl / RgcnLayer1BackwardKernelImpl   (binary_reduce_impl.cu:1082-1117, 1157-1194)
// behind _CAPI_DGLRgcnLayer{0,1}{,Backward} (src/kernel/binary_reduce.cc:412-450).
//
//   layer 0 : ret[v,:]  = sum_e W[type(e), src(e), :] * norm[e]                 W [R, N, F]
//   layer 1 : ret[v,x]  = sum_e sum_y h[src(e),y] * W[type(e),y,x] * norm[e]    W [R, Y, X]
//
// The reference gives one CTA per node with one thread per (y,x) pair, re-reads W per edge and
// reduces over y (and into grad_W) with global atomics. Here a row's edges arrive grouped by type
// (typed CSR), so the per-type partial sum  t[y] = sum_{e in run} norm[e] * h[src,y]  is formed
// first (the HBM-bound gather, 4*Y bytes per edge) and multiplied with W[type] once per run, from
// L1/L2 (R*Y*X floats). The same kernel with W transposed gives grad_h over the typed out-CSR.
// grad_W is a type-major contraction  sum_{e of type r} norm * h[src]^T g[dst]  over fixed chunks
// of the (type, edge id) list, chunk partials folded in chunk order: no atomics, deterministic.
#include <algorithm>

#include "common.cuh"

namespace b200 {

namespace {

__device__ __forceinline__ unsigned group_mask(int G) {
  if (G == 32) return 0xffffffffu;
  const int lane = threadIdx.x & 31;
  return ((1u << G) - 1u) << (lane / G * G);
}

// ------------------------------------------------------------------ layer 0
struct L0Args {
  const int32_t *indptr, *ids, *eids, *types;
  const float *w, *norm, *gout;
  float *ret, *grad_w;
  int n_rows;
  int64_t n_w_rows;  // N of W [R, N, F]
  int F;
  int accumulate;
};

// one group of G lanes per destination row; lane owns columns c0 + k*G + lane
template <int G, int KC>
__global__ void __launch_bounds__(256) rgcn_l0_fwd_kernel(const L0Args a) {
  const int lane = threadIdx.x % G;
  const unsigned mask = group_mask(G);
  const int64_t n_groups = (int64_t)gridDim.x * blockDim.x / G;
  for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G; r < a.n_rows;
       r += n_groups) {
    const int beg = __ldg(a.indptr + r), end = __ldg(a.indptr + r + 1);
    for (int c0 = 0; c0 < a.F; c0 += G * KC) {
      float acc[KC];
#pragma unroll
      for (int k = 0; k < KC; ++k) acc[k] = 0.f;
      for (int p0 = beg; p0 < end; p0 += G) {
        long long ro = 0;
        float n = 0.f;
        if (p0 + lane < end) {
          const int p = p0 + lane;
          ro = ((long long)__ldg(a.types + p) * a.n_w_rows + __ldg(a.ids + p)) * a.F;
          n = __ldg(a.norm + __ldg(a.eids + p));
        }
        const int cnt = min(G, end - p0);
        for (int j = 0; j < cnt; ++j) {
          const long long roj = __shfl_sync(mask, ro, j, G);
          const float nj = __shfl_sync(mask, n, j, G);
#pragma unroll
          for (int k = 0; k < KC; ++k) {
            const int c = c0 + k * G + lane;
            if (c < a.F) acc[k] = __fadd_rn(acc[k], __fmul_rn(__ldg(a.w + roj + c), nj));
          }
        }
      }
#pragma unroll
      for (int k = 0; k < KC; ++k) {
        const int c = c0 + k * G + lane;
        if (c < a.F) a.ret[r * a.F + c] = acc[k];
      }
    }
  }
}

// typed out-CSR: one group per source row u; every (type, u) run is written once.
// accumulate = 0 keeps the reference's plain store (:1004): the last edge of the run wins.
template <int G, int KC>
__global__ void __launch_bounds__(256) rgcn_l0_bwd_kernel(const L0Args a) {
  const int lane = threadIdx.x % G;
  const unsigned mask = group_mask(G);
  const int64_t n_groups = (int64_t)gridDim.x * blockDim.x / G;
  for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G; r < a.n_rows;
       r += n_groups) {
    const int beg = __ldg(a.indptr + r), end = __ldg(a.indptr + r + 1);
    for (int c0 = 0; c0 < a.F; c0 += G * KC) {
      float acc[KC];
#pragma unroll
      for (int k = 0; k < KC; ++k) acc[k] = 0.f;
      int cur = -1;
      auto flush = [&]() {
        const long long ro = ((long long)cur * a.n_w_rows + r) * a.F;
#pragma unroll
        for (int k = 0; k < KC; ++k) {
          const int c = c0 + k * G + lane;
          if (c < a.F) a.grad_w[ro + c] = acc[k];
          acc[k] = 0.f;
        }
      };
      for (int p0 = beg; p0 < end; p0 += G) {
        int d = 0, t = 0;
        float n = 0.f;
        if (p0 + lane < end) {
          const int p = p0 + lane;
          d = __ldg(a.ids + p);
          t = __ldg(a.types + p);
          n = __ldg(a.norm + __ldg(a.eids + p));
        }
        const int cnt = min(G, end - p0);
        for (int j = 0; j < cnt; ++j) {
          const int dj = __shfl_sync(mask, d, j, G);
          const int tj = __shfl_sync(mask, t, j, G);
          const float nj = __shfl_sync(mask, n, j, G);
          if (tj != cur) {
            if (cur >= 0) flush();
            cur = tj;
          }
#pragma unroll
          for (int k = 0; k < KC; ++k) {
            const int c = c0 + k * G + lane;
            if (c < a.F) {
              const float v = __fmul_rn(__ldg(a.gout + (int64_t)dj * a.F + c), nj);
              acc[k] = a.accumulate ? __fadd_rn(acc[k], v) : v;
            }
          }
        }
      }
      if (cur >= 0) flush();
    }
  }
}

// ------------------------------------------------------------------ layer 1
struct L1Args {
  const int32_t *indptr, *ids, *eids, *types;
  const float *feat;  // [n_other, A] gathered rows
  const float *m;     // [R, A, B]
  const float *norm;
  float* out;         // [n_rows, B]
  int n_rows, A, B;
};

// out[r,:] = sum over type runs of (sum_{e in run} norm[e] * feat[other(e),:]) . M[type]
// group of G lanes per row; lane owns feat columns k*G+lane (k < K) and out columns k*G+lane.
template <int G, int K>
__global__ void __launch_bounds__(256) rgcn_l1_kernel(const L1Args a) {
  const int lane = threadIdx.x % G;
  const unsigned mask = group_mask(G);
  const int64_t n_groups = (int64_t)gridDim.x * blockDim.x / G;
  const int A = a.A, B = a.B;
  for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G; r < a.n_rows;
       r += n_groups) {
    const int beg = __ldg(a.indptr + r), end = __ldg(a.indptr + r + 1);
    float t[K], acc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) t[k] = acc[k] = 0.f;
    int cur = -1;
    auto flush = [&]() {
      const float* mt = a.m + (int64_t)cur * A * B;
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
        for (int l = 0; l < G; ++l) {
          const int y = ky * G + l;
          if (y >= A) break;
          const float ty = __shfl_sync(mask, t[ky], l, G);
#pragma unroll
          for (int k = 0; k < K; ++k) {
            const int x = k * G + lane;
            if (x < B) acc[k] = __fadd_rn(acc[k], __fmul_rn(ty, __ldg(mt + y * B + x)));
          }
        }
        t[ky] = 0.f;
      }
    };
    for (int p0 = beg; p0 < end; p0 += G) {
      int o = 0, ty = 0;
      float n = 0.f;
      if (p0 + lane < end) {
        const int p = p0 + lane;
        o = __ldg(a.ids + p);
        ty = __ldg(a.types + p);
        n = __ldg(a.norm + __ldg(a.eids + p));
      }
      const int cnt = min(G, end - p0);
      for (int j = 0; j < cnt; ++j) {
        const int oj = __shfl_sync(mask, o, j, G);
        const int tj = __shfl_sync(mask, ty, j, G);
        const float nj = __shfl_sync(mask, n, j, G);
        if (tj != cur) {
          if (cur >= 0) flush();
          cur = tj;
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const int y = k * G + lane;
          if (y < A) t[k] = __fadd_rn(t[k], __fmul_rn(nj, __ldg(a.feat + (int64_t)oj * A + y)));
        }
      }
    }
    if (cur >= 0) flush();
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int x = k * G + lane;
      if (x < B) a.out[r * B + x] = acc[k];
    }
  }
}

// d = s * b + c on both halves (FFMA2, one issue slot for two fp32 FMAs)
__device__ __forceinline__ float2 ffma2(float s, float2 b, float2 c) {
  unsigned long long sa, bb, cc, d;
  asm("mov.b64 %0, {%1,%1};" : "=l"(sa) : "f"(s));
  asm("mov.b64 %0, {%1,%2};" : "=l"(bb) : "f"(b.x), "f"(b.y));
  asm("mov.b64 %0, {%1,%2};" : "=l"(cc) : "f"(c.x), "f"(c.y));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(sa), "l"(bb), "l"(cc));
  float2 r;
  asm("mov.b64 {%0,%1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(d));
  return r;
}

// Same contraction with compile-time widths (A, B in {16, 32, 64}): max(A,B)/2 lanes per row,
// every lane owns two adjacent feat columns and two adjacent out columns (64-bit loads, FFMA2).
// The run sum t[] is exchanged through shared memory (one STS.64 + A/4 broadcast LDS.128 per
// lane) instead of A shuffles, and the A loads of M[type] per lane are independent and unrolled.
// Rows of one warp proceed independently (group-masked shuffles / __syncwarp): measured on B200,
// letting each row close its runs on its own (0.51 ms on the AM-shaped graph) beats advancing the
// warp's rows in lockstep with a warp-wide contraction (0.66 ms) -- the kernel is bound by the
// indptr -> ids -> feature-row latency chain, not by issue slots.
template <int A, int B>
__global__ void __launch_bounds__(256) rgcn_l1_fast_kernel(const L1Args a) {
  constexpr int GL = (A > B ? A : B) / 2;  // lanes per row
  constexpr int RPW = 32 / GL;             // rows per warp
  __shared__ __align__(16) float ts[8][RPW][A];
  const int lane = threadIdx.x % GL;
  const int grp = (threadIdx.x & 31) / GL;
  const unsigned mask = group_mask(GL);
  float* tsm = ts[threadIdx.x >> 5][grp];
  const bool has_a = 2 * lane < A, has_b = 2 * lane < B;
  const int64_t n_groups = (int64_t)gridDim.x * blockDim.x / GL;
  for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / GL; r < a.n_rows;
       r += n_groups) {
    const int beg = __ldg(a.indptr + r), end = __ldg(a.indptr + r + 1);
    float2 t = make_float2(0.f, 0.f), acc = make_float2(0.f, 0.f);
    int cur = -1;
    auto flush = [&]() {
      if (has_a) *reinterpret_cast<float2*>(tsm + 2 * lane) = t;
      t = make_float2(0.f, 0.f);
      __syncwarp(mask);
      if (has_b) {
        const float2* wrow = reinterpret_cast<const float2*>(a.m + (int64_t)cur * A * B) + lane;
#pragma unroll
        for (int y4 = 0; y4 < A / 4; ++y4) {
          const float4 tv = *reinterpret_cast<const float4*>(tsm + 4 * y4);
          acc = ffma2(tv.x, __ldg(wrow + (4 * y4 + 0) * (B / 2)), acc);
          acc = ffma2(tv.y, __ldg(wrow + (4 * y4 + 1) * (B / 2)), acc);
          acc = ffma2(tv.z, __ldg(wrow + (4 * y4 + 2) * (B / 2)), acc);
          acc = ffma2(tv.w, __ldg(wrow + (4 * y4 + 3) * (B / 2)), acc);
        }
      }
      __syncwarp(mask);
    };
    for (int p0 = beg; p0 < end; p0 += GL) {
      int o = 0, ty = 0;
      float n = 0.f;
      if (p0 + lane < end) {
        const int p = p0 + lane;
        o = __ldg(a.ids + p);
        ty = __ldg(a.types + p);
        n = __ldg(a.norm + __ldg(a.eids + p));
      }
      const int cnt = min(GL, end - p0);
      for (int j = 0; j < cnt; ++j) {
        const int oj = __shfl_sync(mask, o, j, GL);
        const int tj = __shfl_sync(mask, ty, j, GL);
        const float nj = __shfl_sync(mask, n, j, GL);
        if (tj != cur) {
          if (cur >= 0) flush();
          cur = tj;
        }
        if (has_a)
          t = ffma2(nj, __ldg(reinterpret_cast<const float2*>(a.feat + (int64_t)oj * A) + lane), t);
      }
    }
    if (cur >= 0) flush();
    if (has_b) *reinterpret_cast<float2*>(a.out + r * B + 2 * lane) = acc;
  }
}

__global__ void transpose_w_kernel(const float* __restrict__ w, float* __restrict__ wt, int R,
                                   int Y, int X) {
  const int64_t n = (int64_t)R * Y * X;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int x = i % X;
    const int y = (i / X) % Y;
    const int64_t r = i / ((int64_t)X * Y);
    wt[(r * X + x) * Y + y] = w[i];
  }
}

// ------------------------------------------------------------------ grad_W (type-major contraction)
constexpr int kGwChunk = 1024;  // edges of the (type, edge id) list per CTA
constexpr int kGwBatch = 32;    // edges staged in shared memory at a time
constexpr int kGwThreads = 256;

struct GwArgs {
  const int32_t *bt_src, *bt_dst, *bt_eid, *type_ptr;
  const float *h, *gout, *norm;
  float* partial;  // [n_chunks + R, Y*X]; slot of (chunk b, type r) = b + r
  float* grad_w;   // [R, Y, X]
  int64_t E;
  int R, Y, X;
};

template <int P>
__global__ void __launch_bounds__(kGwThreads) rgcn_gw_partial_kernel(const GwArgs a) {
  extern __shared__ float sm[];
  const int Y = a.Y, X = a.X, YX = Y * X;
  float* hs = sm;                      // [kGwBatch][Y], already scaled by norm
  float* gs = hs + kGwBatch * Y;       // [kGwBatch][X]
  __shared__ int s_src[kGwBatch], s_dst[kGwBatch];
  __shared__ float s_n[kGwBatch];
  const int tid = threadIdx.x;
  const int64_t beg = (int64_t)blockIdx.x * kGwChunk;
  const int64_t end = min(a.E, beg + kGwChunk);
  // type of the first position: last r with type_ptr[r] <= beg
  int lo = 0, hi = a.R;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(a.type_ptr + mid) <= beg) lo = mid; else hi = mid;
  }
  int cur = lo;
  int yq[P], xq[P];
  float acc[P];
#pragma unroll
  for (int k = 0; k < P; ++k) {
    const int q = k * kGwThreads + tid;
    yq[k] = q / X;
    xq[k] = q % X;
    acc[k] = 0.f;
  }
  int64_t pos = beg;
  while (pos < end) {
    while (__ldg(a.type_ptr + cur + 1) <= pos) ++cur;  // skip empty types
    const int64_t run_end = min(end, (int64_t)__ldg(a.type_ptr + cur + 1));
    for (; pos < run_end; pos += kGwBatch) {
      const int cnt = (int)min((int64_t)kGwBatch, run_end - pos);
      __syncthreads();  // previous batch consumed
      if (tid < cnt) {
        s_src[tid] = __ldg(a.bt_src + pos + tid);
        s_dst[tid] = __ldg(a.bt_dst + pos + tid);
        s_n[tid] = __ldg(a.norm + __ldg(a.bt_eid + pos + tid));
      }
      __syncthreads();
      for (int i = tid; i < cnt * Y; i += kGwThreads) {
        const int e = i / Y, y = i - e * Y;
        hs[i] = __fmul_rn(__ldg(a.h + (int64_t)s_src[e] * Y + y), s_n[e]);
      }
      for (int i = tid; i < cnt * X; i += kGwThreads) {
        const int e = i / X, x = i - e * X;
        gs[i] = __ldg(a.gout + (int64_t)s_dst[e] * X + x);
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < P; ++k) {
        if (k * kGwThreads + tid < YX) {
          float v = acc[k];
          for (int e = 0; e < cnt; ++e)
            v = __fadd_rn(v, __fmul_rn(hs[e * Y + yq[k]], gs[e * X + xq[k]]));
          acc[k] = v;
        }
      }
    }
    pos = run_end;
    float* slot = a.partial + ((int64_t)blockIdx.x + cur) * YX;
#pragma unroll
    for (int k = 0; k < P; ++k) {
      const int q = k * kGwThreads + tid;
      if (q < YX) slot[q] = acc[k];
      acc[k] = 0.f;
    }
  }
}

// X % 4 == 0: a thread owns one y and four adjacent x (one broadcast LDS + one LDS.128 + two
// FFMA2 per staged edge); Y*X/4 threads per CTA.
__global__ void __launch_bounds__(kGwThreads) rgcn_gw_partial4_kernel(const GwArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int Y = a.Y, X = a.X, X4 = X >> 2;
  float* hs = sm;                 // [kGwBatch][Y], already scaled by norm
  float* gs = hs + kGwBatch * Y;  // [kGwBatch][X]  (kGwBatch * Y * 4 bytes is a multiple of 16)
  __shared__ int s_src[kGwBatch], s_dst[kGwBatch];
  __shared__ float s_n[kGwBatch];
  const int tid = threadIdx.x, nthr = blockDim.x;
  const bool active = tid < Y * X4;
  const int y = tid / X4, x0 = (tid - y * X4) << 2;
  const int64_t beg = (int64_t)blockIdx.x * kGwChunk;
  const int64_t end = min(a.E, beg + kGwChunk);
  int lo = 0, hi = a.R;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(a.type_ptr + mid) <= beg) lo = mid; else hi = mid;
  }
  int cur = lo;
  float2 a01 = make_float2(0.f, 0.f), a23 = make_float2(0.f, 0.f);
  int64_t pos = beg;
  while (pos < end) {
    while (__ldg(a.type_ptr + cur + 1) <= pos) ++cur;  // skip empty types
    const int64_t run_end = min(end, (int64_t)__ldg(a.type_ptr + cur + 1));
    for (; pos < run_end; pos += kGwBatch) {
      const int cnt = (int)min((int64_t)kGwBatch, run_end - pos);
      __syncthreads();  // previous batch consumed
      if (tid < cnt) {
        s_src[tid] = __ldg(a.bt_src + pos + tid);
        s_dst[tid] = __ldg(a.bt_dst + pos + tid);
        s_n[tid] = __ldg(a.norm + __ldg(a.bt_eid + pos + tid));
      }
      __syncthreads();
      for (int i = tid; i < cnt * Y; i += nthr) {
        const int e = i / Y, yy = i - e * Y;
        hs[i] = __fmul_rn(__ldg(a.h + (int64_t)s_src[e] * Y + yy), s_n[e]);
      }
      for (int i = tid; i < cnt * X; i += nthr) {
        const int e = i / X, x = i - e * X;
        gs[i] = __ldg(a.gout + (int64_t)s_dst[e] * X + x);
      }
      __syncthreads();
      if (active) {
#pragma unroll 4
        for (int e = 0; e < cnt; ++e) {
          const float hv = hs[e * Y + y];
          const float4 g = *reinterpret_cast<const float4*>(gs + e * X + x0);
          a01 = ffma2(hv, make_float2(g.x, g.y), a01);
          a23 = ffma2(hv, make_float2(g.z, g.w), a23);
        }
      }
    }
    pos = run_end;
    if (active) {
      float* slot = a.partial + ((int64_t)blockIdx.x + cur) * (Y * X) + y * X + x0;
      *reinterpret_cast<float4*>(slot) = make_float4(a01.x, a01.y, a23.x, a23.y);
    }
    a01 = a23 = make_float2(0.f, 0.f);
  }
}

// one CTA per type: folds the chunk partials of that type in chunk order
__global__ void rgcn_gw_combine_kernel(const GwArgs a) {
  const int r = blockIdx.x;
  const int YX = a.Y * a.X;
  const int64_t tb = __ldg(a.type_ptr + r), te = __ldg(a.type_ptr + r + 1);
  for (int q = threadIdx.x; q < YX; q += blockDim.x) {
    float v = 0.f;
    if (te > tb) {
      const int64_t b0 = tb / kGwChunk, b1 = (te - 1) / kGwChunk;
      const float* p = a.partial + (b0 + r) * YX + q;
#pragma unroll 8
      for (int64_t b = b0; b <= b1; ++b, p += YX) v = __fadd_rn(v, *p);
    }
    a.grad_w[(int64_t)r * YX + q] = v;
  }
}

int grid_for(int64_t n_rows, int G) {
  const int per_cta = 256 / G;
  return (int)std::max<int64_t>(1, std::min<int64_t>((n_rows + per_cta - 1) / per_cta, 148 * 32));
}

int group_for(int width) { return width > 16 ? 32 : (width > 8 ? 16 : 8); }

template <int G>
void launch_l0(const L0Args& a, bool bwd, cudaStream_t s) {
  const int kc = (a.F + G - 1) / G;
  const int grid = grid_for(a.n_rows, G);
#define B200_L0(KC)                                                    \
  do {                                                                 \
    if (bwd) rgcn_l0_bwd_kernel<G, KC><<<grid, 256, 0, s>>>(a);        \
    else rgcn_l0_fwd_kernel<G, KC><<<grid, 256, 0, s>>>(a);            \
  } while (0)
  if (G < 32 || kc <= 1) B200_L0(1);
  else if (kc == 2) B200_L0(2);
  else B200_L0(4);
#undef B200_L0
}

void run_l0(const L0Args& a, bool bwd, cudaStream_t s) {
  if (a.n_rows == 0 || a.F == 0) return;
  switch (group_for(a.F)) {
    case 8: launch_l0<8>(a, bwd, s); break;
    case 16: launch_l0<16>(a, bwd, s); break;
    default: launch_l0<32>(a, bwd, s); break;
  }
  B200_CUDA(cudaGetLastError());
  note_kernel(bwd ? "rgcn_layer0_backward" : "rgcn_layer0", 1);
}

void run_l1(const L1Args& a, const char* name, cudaStream_t s) {
  if (a.n_rows == 0 || a.B == 0) return;
  const int m = std::max(a.A, a.B);
  B200_CHECK(m <= 128, "rgcn layer 1: feature widths above 128 are not supported");
  const bool aligned = (reinterpret_cast<uintptr_t>(a.feat) % 8 == 0) &&
                       (reinterpret_cast<uintptr_t>(a.m) % 8 == 0) &&
                       (reinterpret_cast<uintptr_t>(a.out) % 8 == 0);
  auto pow2w = [](int v) { return v == 16 || v == 32 || v == 64; };
  if (aligned && pow2w(a.A) && pow2w(a.B)) {
    const int grid = grid_for(a.n_rows, m / 2);
#define B200_L1F(AA, BB) \
  if (a.A == AA && a.B == BB) rgcn_l1_fast_kernel<AA, BB><<<grid, 256, 0, s>>>(a)
    B200_L1F(16, 16); B200_L1F(16, 32); B200_L1F(16, 64);
    B200_L1F(32, 16); B200_L1F(32, 32); B200_L1F(32, 64);
    B200_L1F(64, 16); B200_L1F(64, 32); B200_L1F(64, 64);
#undef B200_L1F
    B200_CUDA(cudaGetLastError());
    note_kernel(name, 1);
    return;
  }
  const int G = group_for(m);
  const int grid = grid_for(a.n_rows, G);
  if (G == 8) rgcn_l1_kernel<8, 1><<<grid, 256, 0, s>>>(a);
  else if (G == 16) rgcn_l1_kernel<16, 1><<<grid, 256, 0, s>>>(a);
  else if (m <= 32) rgcn_l1_kernel<32, 1><<<grid, 256, 0, s>>>(a);
  else if (m <= 64) rgcn_l1_kernel<32, 2><<<grid, 256, 0, s>>>(a);
  else rgcn_l1_kernel<32, 4><<<grid, 256, 0, s>>>(a);
  B200_CUDA(cudaGetLastError());
  note_kernel(name, 1);
}

}  // namespace

static void need_types(const DGLB200Graph* g) {
  B200_CHECK(g->ty.num_types > 0,
             "graph has no edge types: call DGLB200GraphSetEdgeTypes (add_edges_with_type) first");
}

void rgcn_layer0_forward(DGLB200Graph* g, const float* w, const float* norm, float* ret,
                         int64_t n_rows_w, int64_t feat, cudaStream_t s) {
  need_types(g);
  L0Args a{};
  a.indptr = g->in.indptr;
  a.ids = g->ty.in.ids;
  a.eids = g->ty.in.eids;
  a.types = g->ty.in.types;
  a.w = w;
  a.norm = norm;
  a.ret = ret;
  a.n_rows = (int)g->num_dst;
  a.n_w_rows = n_rows_w;
  a.F = (int)feat;
  run_l0(a, false, s);
}

void rgcn_layer0_backward(DGLB200Graph* g, const float* gout, const float* norm, float* grad_w,
                          int64_t n_rows_w, int64_t feat, bool accumulate, cudaStream_t s) {
  need_types(g);
  L0Args a{};
  a.indptr = g->out.indptr;
  a.ids = g->ty.out.ids;
  a.eids = g->ty.out.eids;
  a.types = g->ty.out.types;
  a.gout = gout;
  a.norm = norm;
  a.grad_w = grad_w;
  a.n_rows = (int)g->num_src;
  a.n_w_rows = n_rows_w;
  a.F = (int)feat;
  a.accumulate = accumulate ? 1 : 0;
  run_l0(a, true, s);
}

void rgcn_layer1_forward(DGLB200Graph* g, const float* h, const float* w, const float* norm,
                         float* ret, int64_t fy, int64_t fx, cudaStream_t s) {
  need_types(g);
  L1Args a{};
  a.indptr = g->in.indptr;
  a.ids = g->ty.in.ids;
  a.eids = g->ty.in.eids;
  a.types = g->ty.in.types;
  a.feat = h;
  a.m = w;
  a.norm = norm;
  a.out = ret;
  a.n_rows = (int)g->num_dst;
  a.A = (int)fy;
  a.B = (int)fx;
  run_l1(a, "rgcn_layer1", s);
}

void rgcn_layer1_backward(DGLB200Graph* g, const float* h, const float* w, const float* norm,
                          const float* gout, float* grad_h, float* grad_w, int64_t fy, int64_t fx,
                          cudaStream_t s) {
  need_types(g);
  const int R = (int)g->ty.num_types, Y = (int)fy, X = (int)fx;
  const int64_t E = g->num_edges;
  B200_CHECK((int64_t)Y * X <= 4 * kGwThreads, "rgcn layer 1 backward: in*out must be <= 1024");
  // grad_h[u,:] = sum over type runs of (sum_e norm * gout[dst,:]) . W[type]^T
  Scratch wt((size_t)std::max<int64_t>((int64_t)R * Y * X, 1) * 4, s);
  if ((int64_t)R * Y * X > 0) {
    const int blocks = (int)std::min<int64_t>(((int64_t)R * Y * X + 255) / 256, 148 * 8);
    transpose_w_kernel<<<blocks, 256, 0, s>>>(w, wt.as<float>(), R, Y, X);
    note_kernel(nullptr, 1);
  }
  L1Args a{};
  a.indptr = g->out.indptr;
  a.ids = g->ty.out.ids;
  a.eids = g->ty.out.eids;
  a.types = g->ty.out.types;
  a.feat = gout;
  a.m = wt.as<float>();
  a.norm = norm;
  a.out = grad_h;
  a.n_rows = (int)g->num_src;
  a.A = X;
  a.B = Y;
  run_l1(a, "rgcn_layer1_backward", s);
  // grad_W
  if ((int64_t)R * Y * X == 0) return;
  const int64_t n_chunks = (E + kGwChunk - 1) / kGwChunk;
  Scratch partial((size_t)std::max<int64_t>((n_chunks + R) * Y * X, 1) * 4, s);
  GwArgs ga{};
  ga.bt_src = g->ty.bt_src;
  ga.bt_dst = g->ty.bt_dst;
  ga.bt_eid = g->ty.bt_eid;
  ga.type_ptr = g->ty.type_ptr;
  ga.h = h;
  ga.gout = gout;
  ga.norm = norm;
  ga.partial = partial.as<float>();
  ga.grad_w = grad_w;
  ga.E = E;
  ga.R = R;
  ga.Y = Y;
  ga.X = X;
  if (n_chunks > 0) {
    const size_t smem = (size_t)kGwBatch * (Y + X) * 4;
    const int P = (Y * X + kGwThreads - 1) / kGwThreads;
    const int t4 = ((Y * (X / 4) + 31) / 32) * 32;
    if (X % 4 == 0 && (kGwBatch * Y) % 4 == 0 && t4 <= kGwThreads &&
        reinterpret_cast<uintptr_t>(ga.partial) % 16 == 0)
      rgcn_gw_partial4_kernel<<<(int)n_chunks, t4, smem, s>>>(ga);
    else if (P <= 1) rgcn_gw_partial_kernel<1><<<(int)n_chunks, kGwThreads, smem, s>>>(ga);
    else if (P == 2) rgcn_gw_partial_kernel<2><<<(int)n_chunks, kGwThreads, smem, s>>>(ga);
    else rgcn_gw_partial_kernel<4><<<(int)n_chunks, kGwThreads, smem, s>>>(ga);
    B200_CUDA(cudaGetLastError());
    note_kernel(nullptr, 1);
  }
  rgcn_gw_combine_kernel<<<R, std::min(1024, std::max(32, Y * X)), 0, s>>>(ga);
  B200_CUDA(cudaGetLastError());
  note_kernel(nullptr, 1);
}

}  // namespace b200

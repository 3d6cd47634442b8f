// copy_src + sum / mean with bf16 FEATURE STORAGE and fp32 accumulation (an explicit second
// precision tier, north_star's "1e-2 in bf16"; the reference itself is fp32-only,
// src/kernel/common.h:49-55). Same mapping as the fp32 SpMM (spmm_kernels.cuh): a warp (or a
// power-of-two lane group for narrow rows) owns an output row, walks its neighbours sequentially
// in CSR order -- so the result is deterministic -- and keeps fp32 accumulators in registers; the
// gathered rows are bf16 with a row pitch that is a multiple of 8 elements, so every load is
// 16 bytes of 8 features and the gathered bytes per edge are half of the fp32 kernel's.
// Heavy rows use the same segment plans (fp32 partial rows, folded in segment order).
//
//   out[r] = row_scale[r] * (1 / max(deg r, 1) if mean) * sum_{p in row r} col_scale[c(p)] * X[c(p)]
//
// over the (src, eid)-ordered in-CSR (transpose = 0) or the out-CSR (transpose = 1, the gradient).
#include <cuda_bf16.h>

#include <algorithm>

#include "common.cuh"

namespace b200 {
namespace {

struct Bf16Args {
  const int32_t *indptr, *indices;
  int n_seg, seg_len;
  const int32_t *seg_row, *seg_beg, *seg_end, *sched;
  float* partial;  // [n_seg, D] fp32
  int n_rows, lpr, D, nvec;  // D = row pitch in elements (multiple of 8), nvec = D / 8
  const __nv_bfloat16* A;
  const float *col_scale, *row_scale;
  int mean;
  __nv_bfloat16* out;
};

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);            // low half = even element
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}

__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 p = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    w[i] = *reinterpret_cast<const uint32_t*>(&p);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

template <int CH, int U, bool FULL>
__global__ void __launch_bounds__(256) spmm_bf16_kernel(const Bf16Args a) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lpr = FULL ? 32 : a.lpr;
  const int sub = FULL ? lane : (lane & (lpr - 1));
  const int64_t item = FULL ? warp : warp * (32 / lpr) + lane / lpr;
  bool row_ok = item < (int64_t)a.n_seg + a.n_rows;
  const bool is_seg = item < a.n_seg;
  int r = 0, beg = 0, deg = 0;
  if (row_ok) {
    if (is_seg) {
      r = __ldg(a.seg_row + item);
      beg = __ldg(a.seg_beg + item);
      deg = __ldg(a.seg_end + item) - beg;
    } else {
      r = (int)(item - a.n_seg);
      if (a.sched) r = __ldg(a.sched + r);
      beg = __ldg(a.indptr + r);
      deg = __ldg(a.indptr + r + 1) - beg;
      if (a.n_seg > 0 && deg > a.seg_len) {  // heavy row: served by its segments
        row_ok = false;
        deg = 0;
      }
    }
  }
  float acc[CH][8];
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[c][e] = 0.f;
  auto col_ok = [&](int c) -> bool { return sub + c * lpr < a.nvec; };
  const uint4* A4 = reinterpret_cast<const uint4*>(a.A);

  const int maxdeg = FULL ? deg : __reduce_max_sync(0xffffffffu, deg);
  for (int base = 0; base < maxdeg; base += lpr) {
    int id = 0;
    float cs = 1.f;
    if (base + sub < deg) {
      id = __ldg(a.indices + beg + base + sub);
      if (a.col_scale) cs = __ldg(a.col_scale + id);
    }
    const int cnt = min(lpr, maxdeg - base);
    for (int i = 0; i < cnt; i += U) {
      uint4 v[U][CH];
      float s[U];
      bool on[U];
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const int src = (i + k) & (lpr - 1);
        const int nb = __shfl_sync(0xffffffffu, id, src, lpr);
        s[k] = __shfl_sync(0xffffffffu, cs, src, lpr);
        on[k] = (i + k < lpr) && (base + i + k < deg);
        const uint4* row = A4 + (int64_t)nb * a.nvec + sub;
#pragma unroll
        for (int c = 0; c < CH; ++c)
          if (on[k] && col_ok(c)) v[k][c] = __ldg(row + c * lpr);
      }
#pragma unroll
      for (int k = 0; k < U; ++k) {
        if (on[k]) {
#pragma unroll
          for (int c = 0; c < CH; ++c) {
            if (col_ok(c)) {
              float f[8];
              unpack8(v[k][c], f);
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float t = a.col_scale ? __fmul_rn(f[e], s[k]) : f[e];
                acc[c][e] = __fadd_rn(acc[c][e], t);
              }
            }
          }
        }
      }
    }
  }
  if (!row_ok) return;
  if (is_seg) {
    float* p = a.partial + item * (int64_t)a.D;
#pragma unroll
    for (int c = 0; c < CH; ++c)
      if (col_ok(c)) {
        float4* q = reinterpret_cast<float4*>(p + (sub + c * lpr) * 8);
        q[0] = make_float4(acc[c][0], acc[c][1], acc[c][2], acc[c][3]);
        q[1] = make_float4(acc[c][4], acc[c][5], acc[c][6], acc[c][7]);
      }
    return;
  }
  const float dn = (float)(deg > 1 ? deg : 1);
  const float rs = a.row_scale ? __ldg(a.row_scale + r) : 1.f;
  uint4* o = reinterpret_cast<uint4*>(a.out) + (int64_t)r * a.nvec + sub;
#pragma unroll
  for (int c = 0; c < CH; ++c)
    if (col_ok(c)) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        if (a.mean) acc[c][e] = __fdiv_rn(acc[c][e], dn);
        if (a.row_scale) acc[c][e] = __fmul_rn(acc[c][e], rs);
      }
      o[c * lpr] = pack8(acc[c]);
    }
}

// out[row] = fold of the row's fp32 partial rows in segment order -> bf16
__global__ void __launch_bounds__(128)
combine_bf16_kernel(const int32_t* __restrict__ heavy_row, const int32_t* __restrict__ heavy_ptr,
                    const float* __restrict__ partial, const int32_t* __restrict__ indptr,
                    __nv_bfloat16* __restrict__ out, int D, int mean,
                    const float* __restrict__ row_scale) {
  const int h = blockIdx.x;
  const int r = heavy_row[h];
  const int p0 = heavy_ptr[h], p1 = heavy_ptr[h + 1];
  const int deg = indptr[r + 1] - indptr[r];
  const float dn = (float)(deg > 1 ? deg : 1);
  for (int col = threadIdx.x; col < D; col += blockDim.x) {
    float acc = partial[(int64_t)p0 * D + col];
    for (int k = p0 + 1; k < p1; ++k) acc = __fadd_rn(acc, partial[(int64_t)k * D + col]);
    if (mean) acc = __fdiv_rn(acc, dn);
    if (row_scale) acc = __fmul_rn(acc, row_scale[r]);
    out[(int64_t)r * D + col] = __float2bfloat16_rn(acc);
  }
}

template <int CH, int U>
void launch(const Bf16Args& a, int64_t n_items, cudaStream_t s) {
  // 64-thread CTAs by default, as the fp32 kernels (B200_SPMM_BLOCK; spmm_kernels.cuh)
  const char* be = std::getenv("B200_SPMM_BLOCK");
  const int tb = be ? std::atoi(be) : 64;
  const int T = (tb == 128 || tb == 256) ? tb : 64;
  const int wpb = T / 32;
  if (a.lpr == 32) {
    const int blocks = (int)((n_items + wpb - 1) / wpb);
    spmm_bf16_kernel<CH, U, true><<<blocks, T, 0, s>>>(a);
  } else {
    if constexpr (CH == 1) {
      const int64_t warps = (n_items + (32 / a.lpr) - 1) / (32 / a.lpr);
      spmm_bf16_kernel<1, U, false><<<(int)((warps + wpb - 1) / wpb), T, 0, s>>>(a);
    }
  }
}

}  // namespace

void spmm_bf16(DGLB200Graph* g, int transpose, const void* in, const float* col_scale,
               const float* row_scale, int mean, void* out, int64_t D, cudaStream_t s) {
  B200_CHECK(D > 0 && D % 8 == 0 && D <= 32 * 4 * 8,
             "bf16 aggregation: the row pitch must be a multiple of 8 elements, at most 1024");
  const int walk = transpose ? kWalkOut : kWalkInSorted;
  const Csr& csr = walk_csr(g, walk, s);
  if (csr.n_rows == 0) return;
  Bf16Args a;
  a.indptr = csr.indptr;
  a.indices = csr.indices;
  a.n_rows = (int)csr.n_rows;
  a.D = (int)D;
  a.nvec = (int)(D / 8);
  a.lpr = 1;
  while (a.lpr < 32 && a.lpr < a.nvec) a.lpr <<= 1;
  const int ch = a.lpr == 32 ? (a.nvec + 31) / 32 : 1;
  a.A = static_cast<const __nv_bfloat16*>(in);
  a.col_scale = col_scale;
  a.row_scale = row_scale;
  a.mean = mean;
  a.out = static_cast<__nv_bfloat16*>(out);
  const SegPlan plan = csr_segments(g, walk, segment_length(g->num_edges), s);
  Scratch partial;
  if (plan.n_seg > 0) partial.alloc((size_t)plan.n_seg * D * 4, s);
  a.n_seg = plan.n_seg;
  a.seg_len = plan.seg_len;
  a.seg_row = plan.seg_row;
  a.seg_beg = plan.seg_beg;
  a.seg_end = plan.seg_end;
  a.sched = plan.sched;
  a.partial = partial.as<float>();
  const int64_t n_items = (int64_t)plan.n_seg + csr.n_rows;
  // neighbours in flight per lane: ~1 KB per warp per step as in the fp32 kernel
  switch (ch) {
    case 1: launch<1, 4>(a, n_items, s); break;
    case 2: launch<2, 2>(a, n_items, s); break;
    case 3: launch<3, 2>(a, n_items, s); break;
    default: launch<4, 1>(a, n_items, s); break;
  }
  B200_CUDA(cudaGetLastError());
  note_kernel("spmm_bf16", 1);
  if (plan.n_seg > 0) {
    combine_bf16_kernel<<<plan.n_heavy, 128, 0, s>>>(plan.heavy_row, plan.heavy_ptr,
                                                     partial.as<float>(), csr.indptr, a.out, (int)D,
                                                     mean, row_scale);
    B200_CUDA(cudaGetLastError());
    note_kernel(nullptr, 1);
  }
}

}  // namespace b200

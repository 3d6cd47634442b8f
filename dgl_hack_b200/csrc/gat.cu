// Fused GAT attention: replaces the fork's hand-written kernels
//   gatExpLeakyReluSumKernel + gatSumProdZipDivKernel        (src/kernel/cuda/binary_reduce_impl.cu:51-113)
//   fusedGatBackwardGradFeatSrc + fusedGatBackwardGradElEr   (binary_reduce_impl.cu:123-213)
// behind _CAPI_DGLFusedGatKernel / _CAPI_DGLKernelBackwardFusedGat (src/kernel/binary_reduce.cc:
// 380-396, 529-549). Same math, no max-subtraction (as in the reference):
//   exp[e,h]  = v > 0 ? exp(v) : exp(slope * v),  v = el[src,h] + er[dst,h]
//   sum[v,h]  = sum_e exp[e,h]                                   (in-edges of v, in-CSR order)
//   ret[v,h,:] = sum_e (exp[e,h] / sum[v,h]) * feat_src[src,h,:]
// Forward is ONE kernel (one warp per destination row: phase A computes exp and the per-head sums
// with all lanes working on (edge, head) pairs and an ordered shuffle reduction; phase B is the
// weighted gather-reduce with the weights recomputed in registers). Backward is one kernel over the
// out-CSR (grad_feat_src, grad_el and the per-edge scalar t[e,h]) plus one SpMM (copy_e + sum of t
// over the in-CSR) for grad_er -- no atomics anywhere, unlike the reference's K8.
#include <algorithm>

#include "common.cuh"
#include "spmm_kernels.cuh"

namespace b200 {

struct GatArgs {
  const int32_t *indptr, *indices, *eids;
  int n_rows;
  int H, F, D, nvec, lph;  // lph = lanes per head = F / VEC
  const float *ft, *el, *er;
  float *sum, *exp, *ret;  // forward outputs (inputs in backward)
  const float* gout;
  float *grad_ft, *grad_el, *tbuf;  // backward outputs; tbuf [E, H]
  float slope;
};

__device__ __forceinline__ float leaky_exp(float v, float slope) {
  return v > 0.f ? expf(v) : expf(__fmul_rn(slope, v));
}

template <int VEC, int CH>
__global__ void __launch_bounds__(256) gat_fwd_kernel(const GatArgs a) {
  extern __shared__ float sm_sum[];  // [warps][H]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* my_sum = sm_sum + wid * a.H;
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= a.n_rows) return;  // warp-uniform; no CTA-wide sync below
  const int beg = __ldg(a.indptr + r);
  const int deg = __ldg(a.indptr + r + 1) - beg;
  const int H = a.H;

  // ---- phase A: exp[e,h] and sum[v,h]
  const int HB = H < 32 ? H : 32;  // heads per pass
  const int J = 32 / HB;           // edges per iteration
  const int j = lane / HB;
  for (int h0 = 0; h0 < H; h0 += 32) {
    const int h = h0 + lane % HB;
    const bool active = (j < J) && (h < H);
    const float er_h = active ? __ldg(a.er + r * H + h) : 0.f;
    float s = 0.f;
    for (int base = 0; base < deg; base += J) {
      float x = 0.f;
      if (active && base + j < deg) {
        const int p = beg + base + j;
        const int src = __ldg(a.indices + p);
        const int eid = __ldg(a.eids + p);
        x = leaky_exp(__fadd_rn(__ldg(a.el + (int64_t)src * H + h), er_h), a.slope);
        a.exp[(int64_t)eid * H + h] = x;
      }
      const int cnt = min(J, deg - base);
      for (int jj = 0; jj < cnt; ++jj)
        s = __fadd_rn(s, __shfl_sync(0xffffffffu, x, jj * HB + lane % HB));
    }
    if (lane < HB && h < H) {
      my_sum[h] = s;
      a.sum[r * H + h] = s;
    }
  }
  __syncwarp();

  // ---- phase B: ret[v, :] = sum_e (exp / sum) * ft[src, :]
  int col[CH], head[CH];
  bool cok[CH];
  float sum_h[CH], er_c[CH];
  Vec<VEC> acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int vc = c * 32 + lane;
    cok[c] = vc < a.nvec;
    col[c] = vc * VEC;
    head[c] = cok[c] ? col[c] / a.F : 0;
    sum_h[c] = my_sum[head[c]];
    er_c[c] = __ldg(a.er + r * H + head[c]);
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[c].v[e] = 0.f;
  }
  for (int base = 0; base < deg; base += 32) {
    const int src_l = (base + lane < deg) ? __ldg(a.indices + beg + base + lane) : 0;
    const int cnt = min(32, deg - base);
    for (int i = 0; i < cnt; ++i) {
      const int src = __shfl_sync(0xffffffffu, src_l, i);
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        if (cok[c]) {
          const float x =
              leaky_exp(__fadd_rn(__ldg(a.el + (int64_t)src * H + head[c]), er_c[c]), a.slope);
          const float w = __fdiv_rn(x, sum_h[c]);
          const Vec<VEC> f = ld_vec<VEC>(a.ft + (int64_t)src * a.D + col[c]);
#pragma unroll
          for (int e = 0; e < VEC; ++e) acc[c].v[e] = __fadd_rn(acc[c].v[e], __fmul_rn(w, f.v[e]));
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CH; ++c)
    if (cok[c]) st_vec<VEC>(a.ret + r * a.D + col[c], acc[c]);
}

// Backward over the out-CSR: row = source node u.
template <int VEC, int CH>
__global__ void __launch_bounds__(256) gat_bwd_kernel(const GatArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t u = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (u >= a.n_rows) return;
  const int beg = __ldg(a.indptr + u);
  const int deg = __ldg(a.indptr + u + 1) - beg;
  const int H = a.H, lph = a.lph;

  int col[CH], head[CH];
  bool cok[CH];
  float el_c[CH], acc_el[CH];
  Vec<VEC> ftv[CH], acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int vc = c * 32 + lane;
    cok[c] = vc < a.nvec;
    col[c] = vc * VEC;
    head[c] = cok[c] ? col[c] / a.F : 0;
    el_c[c] = __ldg(a.el + u * H + head[c]);
    acc_el[c] = 0.f;
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      acc[c].v[e] = 0.f;
      ftv[c].v[e] = 0.f;
    }
    if (cok[c]) ftv[c] = ld_vec<VEC>(a.ft + u * a.D + col[c]);
  }
  for (int base = 0; base < deg; base += 32) {
    int dst_l = 0, eid_l = 0;
    if (base + lane < deg) {
      dst_l = __ldg(a.indices + beg + base + lane);
      eid_l = __ldg(a.eids + beg + base + lane);
    }
    const int cnt = min(32, deg - base);
    for (int i = 0; i < cnt; ++i) {
      const int dst = __shfl_sync(0xffffffffu, dst_l, i);
      const int eid = __shfl_sync(0xffffffffu, eid_l, i);
      float part[CH];
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        part[c] = 0.f;
        if (cok[c]) {
          const float sm = __ldg(a.sum + (int64_t)dst * H + head[c]);
          const float ex = __ldg(a.exp + (int64_t)eid * H + head[c]);
          const float erv = __ldg(a.er + (int64_t)dst * H + head[c]);
          const Vec<VEC> go = ld_vec<VEC>(a.gout + (int64_t)dst * a.D + col[c]);
          const Vec<VEC> rt = ld_vec<VEC>(a.ret + (int64_t)dst * a.D + col[c]);
          const float w = __fdiv_rn(ex, sm);
          const float gl = (__fadd_rn(el_c[c], erv) > 0.f) ? 1.f : a.slope;
#pragma unroll
          for (int e = 0; e < VEC; ++e) {
            acc[c].v[e] = __fadd_rn(acc[c].v[e], __fmul_rn(w, go.v[e]));
            const float ge = __fdiv_rn(__fmul_rn(go.v[e], __fsub_rn(ftv[c].v[e], rt.v[e])), sm);
            part[c] = __fadd_rn(part[c], __fmul_rn(__fmul_rn(ge, ex), gl));
          }
        }
      }
      // ---- t[e,h] = sum over the head's lanes (fixed butterfly order) and chunks
      if (lph <= 32) {
#pragma unroll
        for (int c = 0; c < CH; ++c) {
          float t = part[c];
          for (int o = lph >> 1; o > 0; o >>= 1) t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, o));
          if (cok[c] && (lane & (lph - 1)) == 0) {
            a.tbuf[(int64_t)eid * H + head[c]] = t;
            acc_el[c] = __fadd_rn(acc_el[c], t);
          }
        }
      } else {
        // a head spans lph / 32 whole chunks
        const int cph = lph >> 5;
        float run = 0.f;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
          float t = part[c];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, o));
          run = (c % cph == 0) ? t : __fadd_rn(run, t);
          if ((c % cph) == cph - 1 && c * 32 < a.nvec && lane == 0) {
            const int h = (c * 32 * VEC) / a.F;
            a.tbuf[(int64_t)eid * H + h] = run;
            acc_el[c] = __fadd_rn(acc_el[c], run);
          }
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    if (cok[c]) st_vec<VEC>(a.grad_ft + u * a.D + col[c], acc[c]);
    if (lph <= 32) {
      if (cok[c] && (lane & (lph - 1)) == 0) a.grad_el[u * H + head[c]] = acc_el[c];
    } else {
      const int cph = lph >> 5;
      if ((c % cph) == cph - 1 && c * 32 < a.nvec && lane == 0)
        a.grad_el[u * H + (c * 32 * VEC) / a.F] = acc_el[c];
    }
  }
}

template <int VEC>
static void launch_gat(bool fwd, const GatArgs& a, int ch, cudaStream_t s) {
  const int blocks = (int)(((int64_t)a.n_rows * 32 + 255) / 256);
  const size_t smem = fwd ? (size_t)8 * a.H * sizeof(float) : 0;
#define B200_GAT(CHV)                                                   \
  do {                                                                  \
    if (fwd) gat_fwd_kernel<VEC, CHV><<<blocks, 256, smem, s>>>(a);     \
    else gat_bwd_kernel<VEC, CHV><<<blocks, 256, 0, s>>>(a);            \
  } while (0)
  switch (ch) {
    case 1: B200_GAT(1); break;
    case 2: B200_GAT(2); break;
    case 3: B200_GAT(3); break;
    case 4: B200_GAT(4); break;
    case 6: B200_GAT(6); break;
    default: B200_GAT(8); break;
  }
#undef B200_GAT
}

// Shape plan shared by forward and backward; throws on unsupported shapes.
static void gat_plan(int64_t H, int64_t F, const void* p0, const void* p1, int* vec, int* ch,
                     GatArgs* a) {
  B200_CHECK(H >= 1 && F >= 1, "fused GAT: empty feature shape");
  int v = 4;
  while (v > 1 && !(F % v == 0 && reinterpret_cast<uintptr_t>(p0) % (v * 4) == 0 &&
                    reinterpret_cast<uintptr_t>(p1) % (v * 4) == 0))
    v >>= 1;
  const int64_t lph = F / v;
  B200_CHECK((lph & (lph - 1)) == 0 && (lph <= 32 || lph % 32 == 0),
             "fused GAT: feature width per head must be VEC x a power of two");
  const int64_t nvec = H * F / v;
  int c = (int)((nvec + 31) / 32);
  B200_CHECK(c <= 8, "fused GAT: H * F is too wide for the fused kernel (max 1024 floats)");
  if (c == 5) c = 6;
  if (c == 7) c = 8;
  B200_CHECK(H <= 1024, "fused GAT: too many heads");
  *vec = v;
  *ch = c;
  a->H = (int)H;
  a->F = (int)F;
  a->D = (int)(H * F);
  a->nvec = (int)nvec;
  a->lph = (int)lph;
}

void fused_gat_forward(DGLB200Graph* g, const float* ft, const float* el, const float* er,
                       float* sum, float* exp, float* ret, int64_t H, int64_t F, float slope,
                       cudaStream_t s) {
  GatArgs a{};
  int vec, ch;
  gat_plan(H, F, ft, ret, &vec, &ch, &a);
  a.indptr = g->in.indptr;
  a.indices = g->in.indices;
  a.eids = g->in.eids;
  a.n_rows = (int)g->num_dst;
  a.ft = ft;
  a.el = el;
  a.er = er;
  a.sum = sum;
  a.exp = exp;
  a.ret = ret;
  a.slope = slope;
  if (a.n_rows == 0) return;
  if (vec == 4) launch_gat<4>(true, a, ch, s);
  else if (vec == 2) launch_gat<2>(true, a, ch, s);
  else launch_gat<1>(true, a, ch, s);
  note_kernel("gat_fwd", 1);
}

void fused_gat_backward(DGLB200Graph* g, const float* ft, const float* el, const float* er,
                        const float* sum, const float* exp, const float* ret, const float* gout,
                        float* grad_ft, float* grad_el, float* grad_er, int64_t H, int64_t F,
                        float slope, cudaStream_t s) {
  GatArgs a{};
  int vec, ch;
  gat_plan(H, F, ft, grad_ft, &vec, &ch, &a);
  if (reinterpret_cast<uintptr_t>(gout) % (vec * 4) != 0 ||
      reinterpret_cast<uintptr_t>(ret) % (vec * 4) != 0)
    throw ApiError{"fused GAT: grad_out / ret are not aligned like feat_src"};
  const int64_t E = g->num_edges;
  Scratch tbuf((size_t)std::max<int64_t>(E * H, 1) * 4, s);
  a.indptr = g->out.indptr;
  a.indices = g->out.indices;
  a.eids = g->out.eids;
  a.n_rows = (int)g->num_src;
  a.ft = ft;
  a.el = el;
  a.er = er;
  a.sum = const_cast<float*>(sum);
  a.exp = const_cast<float*>(exp);
  a.ret = const_cast<float*>(ret);
  a.gout = gout;
  a.grad_ft = grad_ft;
  a.grad_el = grad_el;
  a.tbuf = tbuf.as<float>();
  a.slope = slope;
  if (a.n_rows > 0) {
    if (vec == 4) launch_gat<4>(false, a, ch, s);
    else if (vec == 2) launch_gat<2>(false, a, ch, s);
    else launch_gat<1>(false, a, ch, s);
    note_kernel("gat_bwd", 1);
  }
  // grad_er[v,h] = sum over in-edges of t[e,h]: copy_e + sum
  KernelCall c;
  c.g = g;
  c.mode = kFwd;
  c.red = kSum;
  c.op = kUseLhs;
  c.lhs_tgt = kEdge;
  c.rhs_tgt = kNone;
  c.lhs = tbuf.as<float>();
  c.rhs = nullptr;
  c.out = nullptr;
  c.gout = nullptr;
  c.dst = grad_er;
  c.lhs_rows = E;
  c.rhs_rows = 0;
  c.dst_rows = g->num_dst;
  c.out_rows = g->num_dst;
  c.lhs_map = c.rhs_map = c.out_map = nullptr;
  c.mean = false;
  c.stream = s;
  Bcast bc{};
  bc.ndim = 1;
  bc.lhs_shape[0] = bc.rhs_shape[0] = bc.out_shape[0] = (int)H;
  bc.lhs_stride[0] = bc.rhs_stride[0] = bc.out_stride[0] = 1;
  bc.lhs_len = bc.rhs_len = bc.out_len = H;
  bc.data_len = 1;
  c.bc = bc;
  if (!try_spmm(c)) run_generic(c);
}

}  // namespace b200

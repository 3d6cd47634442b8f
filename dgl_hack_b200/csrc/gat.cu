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
#include <cstdlib>
#include <initializer_list>
#include <string>

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
__global__ void __launch_bounds__(256, (CH <= 2 ? 4 : 2)) gat_fwd_kernel(const GatArgs a) {
  extern __shared__ float sm_sum[];  // [warps][H]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* my_sum = sm_sum + wid * a.H;
  const int H = a.H;
  // persistent warps: rows are short (a handful of edges), so one CTA per 8 rows would spend a
  // large share of its life being launched
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  // The row descriptor and the first 32 neighbour ids of the NEXT row are fetched while the
  // current row is processed: two dependent latencies leave the per-row critical path.
  int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int nbeg = 0, ndeg = 0, nsrc = 0, neid = 0;
  if (r < a.n_rows) {
    nbeg = __ldg(a.indptr + r);
    ndeg = __ldg(a.indptr + r + 1) - nbeg;
    if (lane < ndeg) {
      nsrc = __ldg(a.indices + nbeg + lane);
      neid = __ldg(a.eids + nbeg + lane);
    }
  }
  for (; r < a.n_rows; r += n_warps) {
  __syncwarp();  // my_sum of the previous row has been consumed
  const int beg = nbeg, deg = ndeg;
  const int src0 = nsrc, eid0 = neid;  // first 32 neighbour ids: used by both phases
  if (r + n_warps < a.n_rows) {
    nbeg = __ldg(a.indptr + r + n_warps);
    ndeg = __ldg(a.indptr + r + n_warps + 1) - nbeg;
    nsrc = neid = 0;
    if (lane < ndeg) {
      nsrc = __ldg(a.indices + nbeg + lane);
      neid = __ldg(a.eids + nbeg + lane);
    }
  }
  int col[CH], head[CH];
  bool cok[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int vc = c * 32 + lane;
    cok[c] = vc < a.nvec;
    col[c] = vc * VEC;
    head[c] = cok[c] ? col[c] / a.F : 0;
  }
  // The feature rows of the first PF neighbours do not depend on the attention weights: their
  // loads are issued now and land while phase A gathers el and computes exp.
  constexpr int PF = (CH == 1) ? 4 : ((CH == 2) ? 2 : 0);
  Vec<VEC> f0[PF > 0 ? PF : 1][CH];
#pragma unroll
  for (int k = 0; k < PF; ++k) {
    const int src = __shfl_sync(0xffffffffu, src0, k);
#pragma unroll
    for (int c = 0; c < CH; ++c)
      if (k < deg && cok[c]) f0[k][c] = ld_vec<VEC>(a.ft + (int64_t)src * a.D + col[c]);
  }

  // ---- phase A: exp[e,h] and sum[v,h]
  const int HB = H < 32 ? H : 32;  // heads per pass
  const int J = 32 / HB;           // edges per iteration
  const int j = lane / HB;
  float xa[4] = {0.f, 0.f, 0.f, 0.f};  // exp of the first four phase-A iterations (H <= 32)
  for (int h0 = 0; h0 < H; h0 += 32) {
    const int h = h0 + lane % HB;
    const bool active = (j < J) && (h < H);
    const float er_h = active ? __ldg(a.er + r * H + h) : 0.f;
    float s = 0.f;
    int it = 0;
    for (int base = 0; base < deg; base += J, ++it) {
      float x = 0.f;
      int src = __shfl_sync(0xffffffffu, src0, (base + j) & 31);
      int eid = __shfl_sync(0xffffffffu, eid0, (base + j) & 31);
      if (active && base + j < deg) {
        if (base + j >= 32) {
          src = __ldg(a.indices + beg + base + j);
          eid = __ldg(a.eids + beg + base + j);
        }
        x = leaky_exp(__fadd_rn(__ldg(a.el + (int64_t)src * H + h), er_h), a.slope);
        a.exp[(int64_t)eid * H + h] = x;
      }
      if (it == 0) xa[0] = x;
      else if (it == 1) xa[1] = x;
      else if (it == 2) xa[2] = x;
      else if (it == 3) xa[3] = x;
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
  // exp[e,h] was just written by this warp (phase A): it is read back through L2 (ld.cg; the
  // __syncwarp above orders the warp's stores before these loads) instead of being recomputed --
  // one load replaces an expf per lane and edge.
  float sum_h[CH];
  Vec<VEC> acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    sum_h[c] = my_sum[head[c]];
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[c].v[e] = 0.f;
  }
  {
    float x0[PF > 0 ? PF : 1][CH];
#pragma unroll
    for (int k = 0; k < PF; ++k) {
      const int eid = __shfl_sync(0xffffffffu, eid0, k);
      if (H <= 32) {
        // edge k was handled by phase-A iteration k / J on lanes (k % J) * HB + head
        const int itk = k / J;
        const float xs = itk == 0 ? xa[0] : (itk == 1 ? xa[1] : (itk == 2 ? xa[2] : xa[3]));
#pragma unroll
        for (int c = 0; c < CH; ++c)
          x0[k][c] = __shfl_sync(0xffffffffu, xs, ((k % J) * HB + head[c]) & 31);
      } else {
#pragma unroll
        for (int c = 0; c < CH; ++c)
          if (k < deg && cok[c]) x0[k][c] = __ldcg(a.exp + (int64_t)eid * H + head[c]);
      }
    }
#pragma unroll
    for (int k = 0; k < PF; ++k) {
      if (k < deg) {
#pragma unroll
        for (int c = 0; c < CH; ++c) {
          if (cok[c]) {
            const float w = __fdiv_rn(x0[k][c], sum_h[c]);
#pragma unroll
            for (int e = 0; e < VEC; ++e)
              acc[c].v[e] = __fadd_rn(acc[c].v[e], __fmul_rn(w, f0[k][c].v[e]));
          }
        }
      }
    }
  }
  constexpr int UB = (CH <= 2) ? 2 : 1;  // edges whose loads are issued together
  for (int base = 0; base < deg; base += 32) {
    int src_l = src0, eid_l = eid0;
    if (base > 0) {
      src_l = eid_l = 0;
      if (base + lane < deg) {
        src_l = __ldg(a.indices + beg + base + lane);
        eid_l = __ldg(a.eids + beg + base + lane);
      }
    }
    const int cnt = min(32, deg - base);
    for (int i = (base == 0 ? PF : 0); i < cnt; i += UB) {
      float x[UB][CH];
      Vec<VEC> f[UB][CH];
#pragma unroll
      for (int k = 0; k < UB; ++k) {
        const int src = __shfl_sync(0xffffffffu, src_l, (i + k) & 31);
        const int eid = __shfl_sync(0xffffffffu, eid_l, (i + k) & 31);
        const bool on = i + k < cnt;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
          if (on && cok[c]) {
            x[k][c] = __ldcg(a.exp + (int64_t)eid * H + head[c]);
            f[k][c] = ld_vec<VEC>(a.ft + (int64_t)src * a.D + col[c]);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < UB; ++k) {
        if (i + k < cnt) {
#pragma unroll
          for (int c = 0; c < CH; ++c) {
            if (cok[c]) {
              const float w = __fdiv_rn(x[k][c], sum_h[c]);
#pragma unroll
              for (int e = 0; e < VEC; ++e)
                acc[c].v[e] = __fadd_rn(acc[c].v[e], __fmul_rn(w, f[k][c].v[e]));
            }
          }
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CH; ++c)
    if (cok[c]) st_vec<VEC>(a.ret + r * a.D + col[c], acc[c]);
  }
}

// Backward over the out-CSR: row = source node u.
template <int VEC, int CH>
__global__ void __launch_bounds__(256) gat_bwd_kernel(const GatArgs a) {
  const int lane = threadIdx.x & 31;
  const int H = a.H, lph = a.lph;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t u = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; u < a.n_rows;
       u += n_warps) {
  const int beg = __ldg(a.indptr + u);
  const int deg = __ldg(a.indptr + u + 1) - beg;

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
          // t[e,h] = (sum_f go_f (ft_f - ret_f)) * (exp / sum) * leaky': the reference divides
          // and scales every element before adding; factoring the three per-(edge, head) scalars
          // out of the sum leaves one division per lane instead of VEC
          float dotp = 0.f;
#pragma unroll
          for (int e = 0; e < VEC; ++e) {
            acc[c].v[e] = __fadd_rn(acc[c].v[e], __fmul_rn(w, go.v[e]));
            dotp = __fadd_rn(dotp, __fmul_rn(go.v[e], __fsub_rn(ftv[c].v[e], rt.v[e])));
          }
          part[c] = __fmul_rn(__fmul_rn(dotp, w), gl);
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
}

// ------------------------------------------------------------------ "lane = head slice" kernels
// A row is served by G = D / W lanes (W = 4 * W4 floats per lane, W divides F), so a lane's
// columns belong to ONE head and 32 / G rows share a warp. Everything per-head -- el + er, exp,
// the running sum -- is lane-local: no shuffles, no shared memory, one pass over the edges.
// Forward accumulates the unnormalised sum_e exp * ft[src] and divides by sum once at the end
// (the reference divides every term: same value up to rounding, tolerance 1e-5). Rows are
// issued in degree-class order (the SpMM schedule) so the rows of a warp have similar lengths.
struct Gat2Args {
  GatArgs g;
  const int32_t* sched;
  int n_long, n_heavy;  // with sched: rows served by a whole warp (first / last entries of sched)
  int G, Gp, lph;       // lanes per row, padded to a power of two, lanes per head
};

// Work decomposition shared by forward and backward. Short rows: one group of Gp lanes each,
// 32 / Gp rows per warp. Rows with >= kLongRow edges (and the segment plan's heavy rows): one
// WARP each, the warp's groups taking every (32 / Gp)-th edge and folding their partial results
// at the end -- the serial chain of dependent gathers of a hub row is what bounds this kernel
// on power-law graphs (measured: 462-edge row, 231 dependent ~1 us steps = 70 % of the time).
struct Gat2Item {
  bool valid, wide;
  int row;
};

__device__ __forceinline__ Gat2Item gat2_item(const Gat2Args& b, int64_t w, int grp, int rpw,
                                              int64_t short_warps) {
  Gat2Item t;
  const int n_wide = b.n_long + b.n_heavy;
  int64_t it;
  if (w < b.n_heavy) {
    t.wide = true;
    it = b.g.n_rows - b.n_heavy + w;
  } else if (w < n_wide) {
    t.wide = true;
    it = w - b.n_heavy;
  } else {
    t.wide = false;
    it = b.n_long + (w - n_wide) * rpw + grp;
  }
  t.valid = t.wide || it < b.g.n_rows - b.n_heavy;
  t.row = 0;
  if (t.valid) t.row = b.sched ? __ldg(b.sched + it) : (int)it;
  (void)short_warps;
  return t;
}

__device__ __forceinline__ unsigned gat_group_mask(int Gp) {
  if (Gp == 32) return 0xffffffffu;
  return ((1u << Gp) - 1u) << ((threadIdx.x & 31) / Gp * Gp);
}

template <int W4>
__global__ void __launch_bounds__(256) gat_fwd2_kernel(const Gat2Args b) {
  const GatArgs& a = b.g;
  const int Gp = b.Gp, rpw = 32 / Gp;
  const int lane = threadIdx.x & 31;
  const int li = lane & (Gp - 1), grp = lane / Gp;
  const bool active = li < b.G;
  const int head = active ? li / b.lph : 0;
  const bool writer = active && (li % b.lph == 0);
  const int coff = li * (W4 * 4);
  const int H = a.H;
  const int64_t n_short = a.n_rows - b.n_long - b.n_heavy;
  const int64_t short_warps = (n_short + rpw - 1) / rpw;
  const int64_t n_items = b.n_long + b.n_heavy + short_warps;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_items;
       w += n_warps) {
    const Gat2Item t = gat2_item(b, w, grp, rpw, short_warps);
    const bool wide = t.wide;  // warp-uniform
    const int r = t.row;
    const int bw = wide ? 32 : Gp;            // ids fetched per batch
    const int bl = wide ? lane : li;          // this lane's slot in the batch
    const int stride = wide ? rpw : 1, off = wide ? grp : 0;
    int beg = 0, deg = 0;
    if (t.valid) {
      beg = __ldg(a.indptr + r);
      deg = __ldg(a.indptr + r + 1) - beg;
    }
    const float er_h = (active && t.valid) ? __ldg(a.er + (int64_t)r * H + head) : 0.f;
    float s = 0.f;
    float4 acc[W4];
#pragma unroll
    for (int k = 0; k < W4; ++k) acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    // warp-uniform trip counts + compile-time full-warp shuffle masks (a run-time member mask
    // costs a MATCH/VOTE convergence sequence per shuffle); rows of a warp have similar degrees
    // (degree-class schedule), so the lockstep wastes little
    const int maxdeg = __reduce_max_sync(0xffffffffu, deg);
    for (int base = 0; base < maxdeg; base += bw) {
      int src_l = 0, eid_l = 0;
      if (base + bl < deg) {
        src_l = __ldg(a.indices + beg + base + bl);
        eid_l = __ldg(a.eids + beg + base + bl);
      }
      const int cnt = max(0, min(bw, deg - base));
      const int cmax = min(bw, maxdeg - base);
      for (int j0 = 0; j0 < cmax; j0 += 2 * stride) {
        float elv[2];
        float4 f[2][W4];
        int eid[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int j = j0 + u * stride + off;
          const int src = __shfl_sync(0xffffffffu, src_l, j & (bw - 1), bw);
          eid[u] = __shfl_sync(0xffffffffu, eid_l, j & (bw - 1), bw);
          if (active && j < cnt) {
            elv[u] = __ldg(a.el + (int64_t)src * H + head);
            const float4* fp = reinterpret_cast<const float4*>(a.ft + (int64_t)src * a.D + coff);
#pragma unroll
            for (int k = 0; k < W4; ++k) f[u][k] = __ldg(fp + k);
          }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int j = j0 + u * stride + off;
          if (active && j < cnt) {
            const float x = leaky_exp(__fadd_rn(elv[u], er_h), a.slope);
            if (writer) a.exp[(int64_t)eid[u] * H + head] = x;
            s = __fadd_rn(s, x);
#pragma unroll
            for (int k = 0; k < W4; ++k) {
              acc[k].x = __fadd_rn(acc[k].x, __fmul_rn(x, f[u][k].x));
              acc[k].y = __fadd_rn(acc[k].y, __fmul_rn(x, f[u][k].y));
              acc[k].z = __fadd_rn(acc[k].z, __fmul_rn(x, f[u][k].z));
              acc[k].w = __fadd_rn(acc[k].w, __fmul_rn(x, f[u][k].w));
            }
          }
        }
      }
    }
    if (wide) {  // fold the groups' partial sums (fixed order)
      for (int o = Gp; o < 32; o <<= 1) {
        s = __fadd_rn(s, __shfl_xor_sync(0xffffffffu, s, o));
#pragma unroll
        for (int k = 0; k < W4; ++k) {
          acc[k].x = __fadd_rn(acc[k].x, __shfl_xor_sync(0xffffffffu, acc[k].x, o));
          acc[k].y = __fadd_rn(acc[k].y, __shfl_xor_sync(0xffffffffu, acc[k].y, o));
          acc[k].z = __fadd_rn(acc[k].z, __shfl_xor_sync(0xffffffffu, acc[k].z, o));
          acc[k].w = __fadd_rn(acc[k].w, __shfl_xor_sync(0xffffffffu, acc[k].w, o));
        }
      }
    }
    if (active && t.valid && (!wide || grp == 0)) {
      if (writer) a.sum[(int64_t)r * H + head] = s;
      float4* op = reinterpret_cast<float4*>(a.ret + (int64_t)r * a.D + coff);
#pragma unroll
      for (int k = 0; k < W4; ++k) {
        float4 o = make_float4(0.f, 0.f, 0.f, 0.f);  // rows without in-edges: ret = 0
        if (deg > 0)
          o = make_float4(__fdiv_rn(acc[k].x, s), __fdiv_rn(acc[k].y, s), __fdiv_rn(acc[k].z, s),
                          __fdiv_rn(acc[k].w, s));
        op[k] = o;
      }
    }
  }
}

template <int W4>
__global__ void __launch_bounds__(256) gat_bwd2_kernel(const Gat2Args b) {
  const GatArgs& a = b.g;
  const int Gp = b.Gp, lph = b.lph, rpw = 32 / Gp;
  const int lane = threadIdx.x & 31;
  const int li = lane & (Gp - 1), grp = lane / Gp;
  const bool active = li < b.G;
  const int head = active ? li / lph : 0;
  const bool writer = active && (li % lph == 0);
  const int coff = li * (W4 * 4);
  const int H = a.H;
  const int64_t n_short = a.n_rows - b.n_long - b.n_heavy;
  const int64_t short_warps = (n_short + rpw - 1) / rpw;
  const int64_t n_items = b.n_long + b.n_heavy + short_warps;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_items;
       w += n_warps) {
    const Gat2Item t = gat2_item(b, w, grp, rpw, short_warps);
    const bool wide = t.wide;
    const int u = t.row;
    const int bw = wide ? 32 : Gp;
    const int bl = wide ? lane : li;
    const int stride = wide ? rpw : 1, off = wide ? grp : 0;
    int beg = 0, deg = 0;
    if (t.valid) {
      beg = __ldg(a.indptr + u);
      deg = __ldg(a.indptr + u + 1) - beg;
    }
    float el_h = 0.f, acc_el = 0.f;
    float4 ftv[W4], acc[W4];
#pragma unroll
    for (int k = 0; k < W4; ++k) ftv[k] = acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active && t.valid) {
      el_h = __ldg(a.el + (int64_t)u * H + head);
      const float4* fp = reinterpret_cast<const float4*>(a.ft + (int64_t)u * a.D + coff);
#pragma unroll
      for (int k = 0; k < W4; ++k) ftv[k] = __ldg(fp + k);
    }
    const int maxdeg = __reduce_max_sync(0xffffffffu, deg);  // uniform loops, see the forward
    for (int base = 0; base < maxdeg; base += bw) {
      int dst_l = 0, eid_l = 0;
      if (base + bl < deg) {
        dst_l = __ldg(a.indices + beg + base + bl);
        eid_l = __ldg(a.eids + beg + base + bl);
      }
      const int cnt = max(0, min(bw, deg - base));
      const int cmax = min(bw, maxdeg - base);
      for (int j0 = 0; j0 < cmax; j0 += stride) {
        const int j = j0 + off;
        const int dst = __shfl_sync(0xffffffffu, dst_l, j & (bw - 1), bw);
        const int eid = __shfl_sync(0xffffffffu, eid_l, j & (bw - 1), bw);
        const bool on = active && j < cnt;
        float dotp = 0.f, wgt = 0.f, gl = 0.f;
        if (on) {
          const float sm = __ldg(a.sum + (int64_t)dst * H + head);
          const float ex = __ldg(a.exp + (int64_t)eid * H + head);
          const float erv = __ldg(a.er + (int64_t)dst * H + head);
          const float4* gp = reinterpret_cast<const float4*>(a.gout + (int64_t)dst * a.D + coff);
          const float4* rp = reinterpret_cast<const float4*>(a.ret + (int64_t)dst * a.D + coff);
          float4 go[W4], rt[W4];
#pragma unroll
          for (int k = 0; k < W4; ++k) {
            go[k] = __ldg(gp + k);
            rt[k] = __ldg(rp + k);
          }
          wgt = __fdiv_rn(ex, sm);
          gl = (__fadd_rn(el_h, erv) > 0.f) ? 1.f : a.slope;
#pragma unroll
          for (int k = 0; k < W4; ++k) {
            acc[k].x = __fadd_rn(acc[k].x, __fmul_rn(wgt, go[k].x));
            acc[k].y = __fadd_rn(acc[k].y, __fmul_rn(wgt, go[k].y));
            acc[k].z = __fadd_rn(acc[k].z, __fmul_rn(wgt, go[k].z));
            acc[k].w = __fadd_rn(acc[k].w, __fmul_rn(wgt, go[k].w));
            dotp = __fadd_rn(dotp, __fmul_rn(go[k].x, __fsub_rn(ftv[k].x, rt[k].x)));
            dotp = __fadd_rn(dotp, __fmul_rn(go[k].y, __fsub_rn(ftv[k].y, rt[k].y)));
            dotp = __fadd_rn(dotp, __fmul_rn(go[k].z, __fsub_rn(ftv[k].z, rt[k].z)));
            dotp = __fadd_rn(dotp, __fmul_rn(go[k].w, __fsub_rn(ftv[k].w, rt[k].w)));
          }
        }
        // the lanes of one head sit inside one group: xor offsets < lph never leave it
        for (int o = lph >> 1; o > 0; o >>= 1)
          dotp = __fadd_rn(dotp, __shfl_xor_sync(0xffffffffu, dotp, o));
        if (on && writer) {
          const float tv = __fmul_rn(__fmul_rn(dotp, wgt), gl);
          a.tbuf[(int64_t)eid * H + head] = tv;
          acc_el = __fadd_rn(acc_el, tv);
        }
      }
    }
    if (wide) {
      for (int o = Gp; o < 32; o <<= 1) {
        acc_el = __fadd_rn(acc_el, __shfl_xor_sync(0xffffffffu, acc_el, o));
#pragma unroll
        for (int k = 0; k < W4; ++k) {
          acc[k].x = __fadd_rn(acc[k].x, __shfl_xor_sync(0xffffffffu, acc[k].x, o));
          acc[k].y = __fadd_rn(acc[k].y, __shfl_xor_sync(0xffffffffu, acc[k].y, o));
          acc[k].z = __fadd_rn(acc[k].z, __shfl_xor_sync(0xffffffffu, acc[k].z, o));
          acc[k].w = __fadd_rn(acc[k].w, __shfl_xor_sync(0xffffffffu, acc[k].w, o));
        }
      }
    }
    if (active && t.valid && (!wide || grp == 0)) {
      float4* op = reinterpret_cast<float4*>(a.grad_ft + (int64_t)u * a.D + coff);
#pragma unroll
      for (int k = 0; k < W4; ++k) op[k] = acc[k];
      if (writer) a.grad_el[(int64_t)u * H + head] = acc_el;
    }
  }
}

// Picks the lane layout; false = shape not served by the v2 kernels (old kernels take it).
static bool gat2_plan(int64_t H, int64_t F, std::initializer_list<const void*> ptrs, Gat2Args* b,
                      int* w4) {
  for (const void* p : ptrs)
    if (reinterpret_cast<uintptr_t>(p) % 16 != 0) return false;
  if (F % 4 != 0) return false;
  const int64_t D = H * F;
  int W = (F % 16 == 0) ? 16 : ((F % 8 == 0) ? 8 : 4);
  while (D / W > 32 && W < 32 && F % (2 * W) == 0) W *= 2;
  const int64_t G = D / W, lph = F / W;
  if (G > 32 || (lph & (lph - 1)) != 0) return false;
  int gp = 1;
  while (gp < G) gp <<= 1;
  b->G = (int)G;
  b->Gp = gp;
  b->lph = (int)lph;
  *w4 = W / 4;
  return true;
}

template <bool FWD>
static void launch_gat2(const Gat2Args& b, int w4, cudaStream_t s) {
  const int rpw = 32 / b.Gp;
  const int64_t n_short = b.g.n_rows - b.n_long - b.n_heavy;
  const int64_t n_items = b.n_long + b.n_heavy + (n_short + rpw - 1) / rpw;
  // threads per CTA (B200_GAT_BLOCK = 64 | 128 | 256): see spmm_block_threads()
  const char* be = std::getenv("B200_GAT_BLOCK");
  const int tb = be ? std::atoi(be) : 64;   // arxiv shape: 0.158 / 0.467 ms at 256, 0.154 / 0.453 at 64
  const int T = (tb == 128 || tb == 256) ? tb : 64;
  const int wpb = T / 32;
  const int blocks = (int)std::max<int64_t>(
      1, std::min<int64_t>((n_items + wpb - 1) / wpb, (int64_t)148 * 32 * (8 / wpb)));
#define B200_GAT2(W4V)                                                  \
  do {                                                                  \
    if (FWD) gat_fwd2_kernel<W4V><<<blocks, T, 0, s>>>(b);              \
    else gat_bwd2_kernel<W4V><<<blocks, T, 0, s>>>(b);                  \
  } while (0)
  switch (w4) {
    case 1: B200_GAT2(1); break;
    case 2: B200_GAT2(2); break;
    case 4: B200_GAT2(4); break;
    default: B200_GAT2(8); break;
  }
#undef B200_GAT2
  B200_CUDA(cudaGetLastError());
}

// B200_GAT_PATH=warp forces the warp-per-row kernels (tests cover both)
static bool old_gat_forced() {
  const char* e = std::getenv("B200_GAT_PATH");
  return e && std::string(e) == "warp";
}

template <int VEC>
static void launch_gat(bool fwd, const GatArgs& a, int ch, cudaStream_t s) {
  int blocks = (int)(((int64_t)a.n_rows * 32 + 255) / 256);
  static int cap = -1;  // B200_GAT_CTAS_PER_SM: persistent-grid size (0 = one warp per row)
  if (cap < 0) {
    const char* e = std::getenv("B200_GAT_CTAS_PER_SM");
    cap = e ? std::atoi(e) : 8;
  }
  if (cap > 0) blocks = std::min(blocks, 148 * cap);
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
  const Csr& in = in_plain(g, s);
  a.indptr = in.indptr;
  a.indices = in.indices;
  a.eids = in.eids;
  a.n_rows = (int)g->num_dst;
  a.ft = ft;
  a.el = el;
  a.er = er;
  a.sum = sum;
  a.exp = exp;
  a.ret = ret;
  a.slope = slope;
  if (a.n_rows == 0) return;
  {
    Gat2Args b{};
    int w4;
    if (!old_gat_forced() && gat2_plan(H, F, {ft, ret}, &b, &w4)) {
      b.g = a;
      const SegPlan plan = csr_segments(g, kWalkIn, segment_length(g->num_edges), s);
      b.sched = plan.sched;
      b.n_long = plan.sched ? plan.n_long : 0;
      b.n_heavy = plan.sched ? plan.n_heavy : 0;
      launch_gat2<true>(b, w4, s);
      note_kernel("gat_fwd", 1);
      return;
    }
  }
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
  Gat2Args b{};
  int w4;
  if (a.n_rows > 0 && !old_gat_forced() && gat2_plan(H, F, {ft, ret, gout, grad_ft}, &b, &w4)) {
    b.g = a;
    const SegPlan plan = csr_segments(g, kWalkOut, segment_length(g->num_edges), s);
    b.sched = plan.sched;
    b.n_long = plan.sched ? plan.n_long : 0;
    b.n_heavy = plan.sched ? plan.n_heavy : 0;
    launch_gat2<false>(b, w4, s);
    note_kernel("gat_bwd", 1);
  } else if (a.n_rows > 0) {
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

// Edge-output (SDDMM-shaped) kernels: one result row per EDGE, a single writer per row.
//
// Serves, without broadcasting or user id maps (everything else stays on generic_edge_kernel):
//   * reducer "none" forwards -- u_op_v, v_op_u, e_op_v, u_op_e, ... with op in
//     {add, sub, mul, div, dot} and copy_u / copy_v to edges (apply_edges, the pieces of
//     edge_softmax and of GAT attention: src/kernel/binary_reduce_impl.h:60-120 with
//     ReduceNone, binary_reduce_common.h:233-247);
//   * gradients of EDGE-target operands under the sum reducer (copy_e+sum, u_mul_e+sum, u_add_e ...:
//     BackwardBinaryReduce with the edge as the differentiated operand,
//     src/kernel/cuda/backward_binary_reduce_impl.cuh:19-95): grad_e = grad_out[dst] (x other).
//
// Reference: one thread block row per edge of the out-CSR, feature loop inside (minigun
// advance_all.cuh:15-91). Here a group of L = 2^k lanes (L * VEC >= row length, or 32) owns one
// edge and moves its rows with 128-bit loads. Two traversal orders, picked per call:
//   * edge-id order (COO by edge id, built lazily): output rows are written fully coalesced;
//     right when rows are short (scalars / per-head values), where the endpoint tables sit in L2;
//   * destination-major (the in-CSR sorted by (src, eid)): the dst row stays hot across a row's
//     edges and sources ascend, halving the gather traffic for wide rows.
// dot reduces every head inside the lane group with a fixed butterfly (deterministic).
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "spmm_kernels.cuh"

namespace b200 {

namespace {

enum SdOp { kSdAdd = 0, kSdSub = 1, kSdMul = 2, kSdDiv = 3, kSdDot = 4, kSdCopyP = 5, kSdNegP = 6 };

struct SdArgs {
  const int32_t *u_of, *v_of, *e_of;  // per position: src, dst, edge id (e_of NULL = position)
  int64_t n_pos;
  const float *P, *Q;
  int p_tgt, q_tgt;
  float* out;
  int row_len;  // floats per operand row
  int nvec;     // row_len / VEC
  int lpe;      // lanes per edge (power of two <= 32)
  int lph;      // dot: lanes per head = data_len / VEC (power of two <= lpe)
  int heads;    // dot: output values per edge
  int add_zero; // gradient mode: the reference accumulates 0 + g (normalises -0)
};

template <int OP>
__device__ __forceinline__ float sd_apply(float p, float q) {
  if (OP == kSdAdd) return __fadd_rn(p, q);
  if (OP == kSdSub) return __fsub_rn(p, q);
  if (OP == kSdMul || OP == kSdDot) return __fmul_rn(p, q);
  if (OP == kSdDiv) return __fdiv_rn(p, q);
  if (OP == kSdNegP) return __fmul_rn(p, -1.f);
  return p;
}

template <int VEC, int OP>
__global__ void __launch_bounds__(256) sddmm_kernel(const SdArgs a) {
  const int L = a.lpe;
  const int sub = threadIdx.x & (L - 1);
  const int64_t n_groups = (int64_t)gridDim.x * blockDim.x / L;
  constexpr bool kUnary = (OP == kSdCopyP || OP == kSdNegP);
  for (int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / L; p < a.n_pos;
       p += n_groups) {
    const int u = __ldg(a.u_of + p), v = __ldg(a.v_of + p);
    const int e = a.e_of ? __ldg(a.e_of + p) : (int)p;
    const int pid = a.p_tgt == kSrc ? u : (a.p_tgt == kDst ? v : e);
    const int qid = a.q_tgt == kSrc ? u : (a.q_tgt == kDst ? v : e);
    const float* pp = a.P + (int64_t)pid * a.row_len;
    const float* qp = kUnary ? pp : a.Q + (int64_t)qid * a.row_len;
    if (OP != kSdDot) {
      float* op = a.out + (int64_t)e * a.row_len;
      for (int vc = sub; vc < a.nvec; vc += L) {
        const Vec<VEC> x = ld_vec<VEC>(pp + vc * VEC);
        Vec<VEC> y = x, r;
        if (!kUnary) y = ld_vec<VEC>(qp + vc * VEC);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          r.v[k] = sd_apply<OP>(x.v[k], y.v[k]);
          if (a.add_zero) r.v[k] = __fadd_rn(0.f, r.v[k]);
        }
        st_vec<VEC>(op + vc * VEC, r);
      }
    } else {
      // all lanes of the group run the same number of iterations (nvec is a multiple of lph,
      // L a multiple of lph): the shuffles below see whole groups
      const unsigned mask = (L == 32) ? 0xffffffffu
                                      : (((1u << L) - 1u) << ((threadIdx.x & 31) / L * L));
      for (int vc0 = 0; vc0 < a.nvec; vc0 += L) {
        const int vc = vc0 + sub;
        float part = 0.f;
        if (vc < a.nvec) {
          const Vec<VEC> x = ld_vec<VEC>(pp + vc * VEC);
          const Vec<VEC> y = ld_vec<VEC>(qp + vc * VEC);
#pragma unroll
          for (int k = 0; k < VEC; ++k) part = __fadd_rn(part, __fmul_rn(x.v[k], y.v[k]));
        }
        for (int o = a.lph >> 1; o > 0; o >>= 1)
          part = __fadd_rn(part, __shfl_xor_sync(mask, part, o, L));
        if (vc < a.nvec && (sub & (a.lph - 1)) == 0)
          a.out[(int64_t)e * a.heads + vc / a.lph] = part;
      }
    }
  }
}

// dot with the whole operand row in one pass of the lane group (nvec <= L): two edges per
// iteration so that four row gathers are in flight per group.
template <int VEC>
__global__ void __launch_bounds__(256) sddmm_dot2_kernel(const SdArgs a) {
  const int L = a.lpe;
  const int sub = threadIdx.x & (L - 1);
  const unsigned mask = (L == 32) ? 0xffffffffu
                                  : (((1u << L) - 1u) << ((threadIdx.x & 31) / L * L));
  const int64_t n_groups = (int64_t)gridDim.x * blockDim.x / L;
  const bool lane_on = sub < a.nvec;
  for (int64_t p0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / L; p0 < a.n_pos;
       p0 += 2 * n_groups) {
    int e[2];
    Vec<VEC> x[2], y[2];
    bool on[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int64_t p = p0 + k * n_groups;
      on[k] = p < a.n_pos;
      e[k] = 0;
      if (on[k]) {
        const int u = __ldg(a.u_of + p), v = __ldg(a.v_of + p);
        e[k] = a.e_of ? __ldg(a.e_of + p) : (int)p;
        const int pid = a.p_tgt == kSrc ? u : (a.p_tgt == kDst ? v : e[k]);
        const int qid = a.q_tgt == kSrc ? u : (a.q_tgt == kDst ? v : e[k]);
        if (lane_on) {
          x[k] = ld_vec<VEC>(a.P + (int64_t)pid * a.row_len + sub * VEC);
          y[k] = ld_vec<VEC>(a.Q + (int64_t)qid * a.row_len + sub * VEC);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      float part = 0.f;
      if (on[k] && lane_on) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) part = __fadd_rn(part, __fmul_rn(x[k].v[j], y[k].v[j]));
      }
      // on[k] is uniform inside a lane group
      if (on[k]) {
        for (int o = a.lph >> 1; o > 0; o >>= 1)
          part = __fadd_rn(part, __shfl_xor_sync(mask, part, o, L));
        if (lane_on && (sub & (a.lph - 1)) == 0)
          a.out[(int64_t)e[k] * a.heads + sub / a.lph] = part;
      }
    }
  }
}

// dot over the destination-major walk: a lane group takes a CHUNK of consecutive positions, so
// the destination's row (constant along a CSR row) is loaded once per run of equal destinations
// and kept in registers; only the source rows are gathered per edge.
constexpr int kDotChunk = 64;

template <int VEC>
__global__ void __launch_bounds__(256) sddmm_dot_rows_kernel(const SdArgs a) {
  // Every loop below has a warp-uniform trip count and every shuffle names the full warp with a
  // compile-time mask: a run-time member mask makes ptxas wrap each shuffle in a
  // MATCH/REDUX/VOTE convergence sequence (measured: 58 -> instructions per edge, issue bound).
  const int L = a.lpe;
  const int sub = threadIdx.x & (L - 1);
  const int64_t n_groups = (int64_t)gridDim.x * blockDim.x / L;
  const int64_t n_chunks = (a.n_pos + kDotChunk - 1) / kDotChunk;
  const int64_t n_iter = (n_chunks + n_groups - 1) / n_groups;
  const bool lane_on = sub < a.nvec;
  // the operand indexed by the destination is the cached one
  const bool p_is_dst = a.p_tgt == kDst;
  const float* row_tab = p_is_dst ? a.P : a.Q;   // indexed by v
  const float* gat_tab = p_is_dst ? a.Q : a.P;   // indexed by u
  int64_t ck = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / L;
  for (int64_t it = 0; it < n_iter; ++it, ck += n_groups) {
    const int64_t beg = ck * kDotChunk;
    const int n = ck < n_chunks ? (int)min((int64_t)kDotChunk, a.n_pos - beg) : 0;
    int prev_v = -1;
    Vec<VEC> rv;
#pragma unroll
    for (int j = 0; j < VEC; ++j) rv.v[j] = 0.f;
    for (int b0 = 0; b0 < kDotChunk; b0 += L) {
      int u_l = 0, v_l = 0, e_l = 0;
      if (b0 + sub < n) {
        u_l = __ldcs(a.u_of + beg + b0 + sub);  // index streams: evict first, keep the rows in L2
        v_l = __ldcs(a.v_of + beg + b0 + sub);
        e_l = __ldcs(a.e_of + beg + b0 + sub);
      }
      for (int i = 0; i < L; i += 2) {
        Vec<VEC> x[2];
        int v[2], e[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const int u = __shfl_sync(0xffffffffu, u_l, i + k, L);
          v[k] = __shfl_sync(0xffffffffu, v_l, i + k, L);
          e[k] = __shfl_sync(0xffffffffu, e_l, i + k, L);
          if (b0 + i + k < n && lane_on)
            x[k] = ld_vec<VEC>(gat_tab + (int64_t)u * a.row_len + sub * VEC);
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const bool on = b0 + i + k < n;
          if (on && v[k] != prev_v) {
            prev_v = v[k];
            if (lane_on) rv = ld_vec<VEC>(row_tab + (int64_t)prev_v * a.row_len + sub * VEC);
          }
          float part = 0.f;
          if (on && lane_on) {
#pragma unroll
            for (int j = 0; j < VEC; ++j) part = __fadd_rn(part, __fmul_rn(x[k].v[j], rv.v[j]));
          }
          for (int o = a.lph >> 1; o > 0; o >>= 1)
            part = __fadd_rn(part, __shfl_xor_sync(0xffffffffu, part, o));
          if (on && lane_on && (sub & (a.lph - 1)) == 0)
            __stcs(a.out + (int64_t)e[k] * a.heads + sub / a.lph, part);
        }
      }
    }
  }
}

template <int VEC>
void launch_sd(const SdArgs& a, int op, int blocks, cudaStream_t s) {
  switch (op) {
    case kSdAdd: sddmm_kernel<VEC, kSdAdd><<<blocks, 256, 0, s>>>(a); break;
    case kSdSub: sddmm_kernel<VEC, kSdSub><<<blocks, 256, 0, s>>>(a); break;
    case kSdMul: sddmm_kernel<VEC, kSdMul><<<blocks, 256, 0, s>>>(a); break;
    case kSdDiv: sddmm_kernel<VEC, kSdDiv><<<blocks, 256, 0, s>>>(a); break;
    case kSdDot: sddmm_kernel<VEC, kSdDot><<<blocks, 256, 0, s>>>(a); break;
    case kSdCopyP: sddmm_kernel<VEC, kSdCopyP><<<blocks, 256, 0, s>>>(a); break;
    default: sddmm_kernel<VEC, kSdNegP><<<blocks, 256, 0, s>>>(a); break;
  }
}

bool same_shape(const Bcast& bc) {
  if (bc.lhs_len != bc.out_len || bc.rhs_len != bc.out_len) return false;
  for (int d = 0; d < bc.ndim; ++d)
    if (bc.lhs_shape[d] != bc.out_shape[d] || bc.rhs_shape[d] != bc.out_shape[d]) return false;
  return true;
}

}  // namespace

bool try_sddmm(const KernelCall& c) {
  DGLB200Graph* g = c.g;
  cudaStream_t s = c.stream;
  const int64_t E = g->num_edges;
  if (c.lhs_map || c.rhs_map || c.out_map) return false;
  SdArgs a{};
  int op;
  int64_t row_len;
  if (c.mode == kFwd) {
    if (c.red != kRNone) return false;
    if (c.dst_rows != E) return false;
    if (c.op == kUseLhs) {
      if (c.lhs_tgt == kNone) return false;
      op = kSdCopyP;
      a.P = c.lhs;
      a.p_tgt = c.lhs_tgt;
      a.q_tgt = c.lhs_tgt;
      row_len = c.bc.lhs_len * c.bc.data_len;
    } else {
      if (c.lhs_tgt == kNone || c.rhs_tgt == kNone) return false;
      if (!same_shape(c.bc)) return false;
      op = c.op == kAdd ? kSdAdd : c.op == kSub ? kSdSub : c.op == kMul ? kSdMul
           : c.op == kDiv ? kSdDiv : c.op == kDot ? kSdDot : -1;
      if (op < 0) return false;
      if (op != kSdDot && c.bc.data_len != 1) return false;
      a.P = c.lhs;
      a.p_tgt = c.lhs_tgt;
      a.Q = c.rhs;
      a.q_tgt = c.rhs_tgt;
      row_len = c.bc.lhs_len * c.bc.data_len;
    }
  } else {
    // gradient of an edge-target operand under sum: one contribution per edge
    if (c.red != kSum) return false;
    const bool lhs_side = (c.mode == kGradLhs);
    const int t_tgt = lhs_side ? c.lhs_tgt : c.rhs_tgt;
    if (t_tgt != kEdge || c.dst_rows != E) return false;
    if (c.op == kDot || c.op == kDiv || c.bc.data_len != 1) return false;
    if (c.op != kUseLhs && !same_shape(c.bc)) return false;
    a.P = c.gout;
    a.p_tgt = kDst;
    a.q_tgt = kDst;
    a.add_zero = 1;
    row_len = c.bc.out_len;
    if (c.op == kUseLhs || c.op == kAdd || (c.op == kSub && lhs_side)) {
      op = kSdCopyP;
    } else if (c.op == kSub) {
      op = kSdNegP;
    } else if (c.op == kMul) {
      op = kSdMul;
      a.Q = lhs_side ? c.rhs : c.lhs;
      a.q_tgt = lhs_side ? c.rhs_tgt : c.lhs_tgt;
      if (a.q_tgt == kNone) return false;
    } else {
      return false;
    }
  }
  if (row_len <= 0 || row_len >= (1 << 30)) return false;
  if (E == 0) return true;

  auto aligned = [](const void* p, int bytes) {
    return p == nullptr || (reinterpret_cast<uintptr_t>(p) % bytes) == 0;
  };
  const int64_t unit = (op == kSdDot) ? c.bc.data_len : row_len;
  int vec = 4;
  while (vec > 1 && !(unit % vec == 0 && aligned(a.P, vec * 4) && aligned(a.Q, vec * 4) &&
                      (op == kSdDot || aligned(c.dst, vec * 4))))
    vec >>= 1;
  const int64_t nvec = row_len / vec;
  int lpe = 1;
  while (lpe < 32 && lpe < nvec) lpe <<= 1;
  if (op == kSdDot) {
    const int64_t lph = c.bc.data_len / vec;
    if (lph > 32 || (lph & (lph - 1)) != 0) return false;  // generic kernel handles odd lengths
    a.lph = (int)lph;
    a.heads = (int)c.bc.out_len;
    lpe = std::max<int>(lpe, (int)lph);
  }
  a.out = c.dst;
  a.row_len = (int)row_len;
  a.nvec = (int)nvec;
  a.lpe = lpe;
  a.n_pos = E;

  // traversal order
  const bool node_pair = (a.p_tgt != kEdge) && (a.q_tgt != kEdge) && (a.p_tgt != a.q_tgt);
  bool wide = row_len * 4 >= 256;
  if (const char* w = std::getenv("B200_SDDMM_WALK")) wide = (w[0] == 'c' && w[1] == 's');  // csr|coo
  if (node_pair && wide) {
    a.u_of = g->in_s.indices;
    a.v_of = csr_rows(g, kWalkIn, s);
    a.e_of = g->in_s.eids;
  } else {
    a.u_of = graph_coo(g, 0, s);
    a.v_of = graph_coo(g, 1, s);
    a.e_of = nullptr;
  }
  const int64_t per_cta = 256 / lpe;
  const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((E + per_cta - 1) / per_cta,
                                                                (int64_t)148 * 64));
  if (op == kSdDot && nvec <= lpe && a.e_of != nullptr) {
    const int64_t n_chunks = (E + kDotChunk - 1) / kDotChunk;
    const int cblocks = (int)std::max<int64_t>(
        1, std::min<int64_t>((n_chunks + per_cta - 1) / per_cta, (int64_t)148 * 64));
    if (vec == 4) sddmm_dot_rows_kernel<4><<<cblocks, 256, 0, s>>>(a);
    else if (vec == 2) sddmm_dot_rows_kernel<2><<<cblocks, 256, 0, s>>>(a);
    else sddmm_dot_rows_kernel<1><<<cblocks, 256, 0, s>>>(a);
  } else if (op == kSdDot && nvec <= lpe) {
    if (vec == 4) sddmm_dot2_kernel<4><<<blocks, 256, 0, s>>>(a);
    else if (vec == 2) sddmm_dot2_kernel<2><<<blocks, 256, 0, s>>>(a);
    else sddmm_dot2_kernel<1><<<blocks, 256, 0, s>>>(a);
  } else if (vec == 4) launch_sd<4>(a, op, blocks, s);
  else if (vec == 2) launch_sd<2>(a, op, blocks, s);
  else launch_sd<1>(a, op, blocks, s);
  B200_CUDA(cudaGetLastError());
  note_kernel("sddmm", 1);
  return true;
}

}  // namespace b200

// Generic (every op x target x reducer x broadcast x id-mapping combination) kernels.
//
// These two kernels are the complete semantic table of the reference's BinaryReduce /
// BinaryReduceBcast / BackwardBinaryReduce / BackwardBinaryReduceBcast UDFs
// (src/kernel/cuda/binary_reduce_impl.cuh:22-143, backward_binary_reduce_impl.cuh:21-163) without
// atomics and without the per-edge binary search of minigun's advance
// (third_party/minigun/minigun/cuda/advance_all.cuh:15-91):
//
//   * generic_row_kernel  -- output rows are NODES. One lane group per CSR row walks that row's
//     edges sequentially and writes the row once (forward reduce into dst rows over the in-CSR;
//     gradients of src-target operands over the out-CSR, of dst-target operands over the in-CSR).
//   * generic_edge_kernel -- output rows are EDGES (reducer "none" forward, gradients of
//     edge-target operands): one lane group per CSR position, single writer per output row.
//
// The specialised kernels in spmm.cu / sddmm.cu take the hot shapes; anything they decline lands
// here, so this file defines correctness for the whole 26-combination x 5-reducer table.
#include <algorithm>

#include "common.cuh"

namespace b200 {

struct GArgs {
  const int32_t* indptr;
  const int32_t* indices;
  const int32_t* eids;
  const int32_t* row;  // edge kernel only
  int n_rows;
  int n_pos;
  int row_is_dst;  // 1: iterating the in-CSR (row = dst, col = src); 0: out-CSR
  int lhs_tgt, rhs_tgt, out_tgt;
  // For an EDGE-target operand: position-indexed id array (NULL = eids). For a node-target
  // operand: node id map (NULL = identity).
  const int32_t *lhs_ids, *rhs_ids, *out_ids;
  const float *lhs, *rhs, *out, *gout;
  float* dst;
  int op, red, mode;
  int n_elem;  // elements per destination row
  int need_e;  // backward: the forward edge value is required (max/min/prod)
  float scale_by_deg;  // forward mean extension (0 = off)
  Bcast bc;
};

__device__ __forceinline__ int pick_id(int tgt, const int32_t* __restrict__ ids,
                                       const int32_t* __restrict__ eids, int u, int v, int p) {
  if (tgt == kEdge) return ids ? __ldg(ids + p) : __ldg(eids + p);
  if (tgt == kNone) return 0;
  const int id = (tgt == kSrc) ? u : v;
  return ids ? __ldg(ids + id) : id;
}

// offsets (in elements, already multiplied by data_len) of output element `tx` inside one lhs /
// rhs row: Unravel + Ravel of binary_reduce_impl.cuh:63-84 hoisted out of the edge loop.
__device__ __forceinline__ void bcast_offsets(const Bcast& bc, int tx, int64_t* lo, int64_t* ro) {
  int64_t l = 0, r = 0;
#pragma unroll 1
  for (int d = 0; d < bc.ndim; ++d) {
    const int i = (tx / bc.out_stride[d]) % bc.out_shape[d];
    l += (int64_t)min(i, bc.lhs_shape[d] - 1) * bc.lhs_stride[d];
    r += (int64_t)min(i, bc.rhs_shape[d] - 1) * bc.rhs_stride[d];
  }
  *lo = l * bc.data_len;
  *ro = r * bc.data_len;
}

__global__ void __launch_bounds__(256) generic_row_kernel(const GArgs a, int lanes_per_row) {
  const int rows_per_block = blockDim.x / lanes_per_row;
  const int lane = threadIdx.x % lanes_per_row;
  const int len = (int)a.bc.data_len;
  const int64_t lhs_row_len = a.bc.lhs_len * len, rhs_row_len = a.bc.rhs_len * len;
  const int out_len = (int)a.bc.out_len;
  const int32_t* dmap = a.mode == kFwd ? a.out_ids : (a.mode == kGradLhs ? a.lhs_ids : a.rhs_ids);
  for (int64_t r = (int64_t)blockIdx.x * rows_per_block + threadIdx.x / lanes_per_row;
       r < a.n_rows; r += (int64_t)gridDim.x * rows_per_block) {
    const int beg = __ldg(a.indptr + r), end = __ldg(a.indptr + r + 1);
    // with a destination id map only rows that have edges own a slot (reference maps hold -1 or
    // end early for the others); such rows keep the value written by the fill
    if (dmap && beg == end) continue;
    const int64_t drow = dmap ? __ldg(dmap + r) : r;
    for (int j = lane; j < a.n_elem; j += lanes_per_row) {
      int tx = j, k = 0;
      if (a.mode != kFwd && len > 1) {
        tx = j / len;
        k = j - tx * len;
      }
      int64_t lo, ro;
      bcast_offsets(a.bc, tx, &lo, &ro);
      float acc = (a.mode == kFwd) ? red_identity(a.red) : 0.f;
      for (int p = beg; p < end; ++p) {
        const int c = __ldg(a.indices + p);
        const int u = a.row_is_dst ? c : (int)r;
        const int v = a.row_is_dst ? (int)r : c;
        const int lid = pick_id(a.lhs_tgt, a.lhs_ids, a.eids, u, v, p);
        const float* lp = a.lhs + (int64_t)lid * lhs_row_len + lo;
        const float* rp = lp;
        if (a.op != kUseLhs) {
          const int rid = pick_id(a.rhs_tgt, a.rhs_ids, a.eids, u, v, p);
          rp = a.rhs + (int64_t)rid * rhs_row_len + ro;
        }
        if (a.mode == kFwd) {
          acc = red_call(a.red, acc, op_call(a.op, lp, rp, len));
        } else {
          const int oid = pick_id(a.out_tgt, a.out_ids, a.eids, u, v, p);
          const float e = a.need_e ? op_call(a.op, lp, rp, len) : 0.f;
          const float go = __ldg(a.gout + (int64_t)oid * out_len + tx);
          const float ov = a.need_e ? __ldg(a.out + (int64_t)oid * out_len + tx) : 0.f;
          const float ge = __fmul_rn(go, red_bwd(a.red, e, ov));
          const float lv = __ldg(lp + k);
          const float rv = (a.op != kUseLhs) ? __ldg(rp + k) : 0.f;
          const float d = (a.mode == kGradLhs) ? op_bwd_lhs(a.op, lv, rv) : op_bwd_rhs(a.op, lv, rv);
          acc = __fadd_rn(acc, __fmul_rn(ge, d));
        }
      }
      if (a.scale_by_deg != 0.f) {
        const int deg = end - beg;
        acc = __fdiv_rn(acc, (float)(deg > 1 ? deg : 1));
      }
      a.dst[drow * a.n_elem + j] = acc;
    }
  }
}

__global__ void __launch_bounds__(256) generic_edge_kernel(const GArgs a, int lanes_per_pos) {
  const int pos_per_block = blockDim.x / lanes_per_pos;
  const int lane = threadIdx.x % lanes_per_pos;
  const int len = (int)a.bc.data_len;
  const int64_t lhs_row_len = a.bc.lhs_len * len, rhs_row_len = a.bc.rhs_len * len;
  const int out_len = (int)a.bc.out_len;
  for (int64_t p = (int64_t)blockIdx.x * pos_per_block + threadIdx.x / lanes_per_pos;
       p < a.n_pos; p += (int64_t)gridDim.x * pos_per_block) {
    const int r = __ldg(a.row + p);
    const int c = __ldg(a.indices + p);
    const int u = a.row_is_dst ? c : r;
    const int v = a.row_is_dst ? r : c;
    const int lid = pick_id(a.lhs_tgt, a.lhs_ids, a.eids, u, v, (int)p);
    const int rid = (a.op != kUseLhs) ? pick_id(a.rhs_tgt, a.rhs_ids, a.eids, u, v, (int)p) : 0;
    const int oid = pick_id(a.out_tgt, a.out_ids, a.eids, u, v, (int)p);
    const int64_t drow = (a.mode == kFwd) ? oid : (a.mode == kGradLhs ? lid : rid);
    for (int j = lane; j < a.n_elem; j += lanes_per_pos) {
      int tx = j, k = 0;
      if (a.mode != kFwd && len > 1) {
        tx = j / len;
        k = j - tx * len;
      }
      int64_t lo, ro;
      bcast_offsets(a.bc, tx, &lo, &ro);
      const float* lp = a.lhs + (int64_t)lid * lhs_row_len + lo;
      const float* rp = (a.op != kUseLhs) ? a.rhs + (int64_t)rid * rhs_row_len + ro : lp;
      float res;
      if (a.mode == kFwd) {
        res = op_call(a.op, lp, rp, len);
      } else {
        const float e = a.need_e ? op_call(a.op, lp, rp, len) : 0.f;
        const float go = __ldg(a.gout + (int64_t)oid * out_len + tx);
        const float ov = a.need_e ? __ldg(a.out + (int64_t)oid * out_len + tx) : 0.f;
        const float ge = __fmul_rn(go, red_bwd(a.red, e, ov));
        const float lv = __ldg(lp + k);
        const float rv = (a.op != kUseLhs) ? __ldg(rp + k) : 0.f;
        const float d = (a.mode == kGradLhs) ? op_bwd_lhs(a.op, lv, rv) : op_bwd_rhs(a.op, lv, rv);
        // the reference accumulates 0 + g (zero-filled buffer, one contribution)
        res = __fadd_rn(0.f, __fmul_rn(ge, d));
      }
      a.dst[drow * a.n_elem + j] = res;
    }
  }
}

// tmp[p] = map[to_other[p]]: re-index a user edge map given for the OTHER CSR's positions
__global__ void compose_map_kernel(const int32_t* __restrict__ map,
                                   const int32_t* __restrict__ to_other,
                                   int32_t* __restrict__ out, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride)
    out[k] = map[to_other[k]];
}

static int pow2_at_most(int64_t n, int cap) {
  int p = 1;
  while (p * 2 <= cap && p < n) p *= 2;
  return p;
}

void run_generic(const KernelCall& c) {
  DGLB200Graph* g = c.g;
  cudaStream_t s = c.stream;
  const int out_tgt = (c.red == kRNone) ? kEdge : kDst;
  const int dst_tgt = c.mode == kFwd ? out_tgt : (c.mode == kGradLhs ? c.lhs_tgt : c.rhs_tgt);
  const int64_t len = c.bc.data_len;
  const int64_t n_elem = (c.mode == kFwd) ? c.bc.out_len : c.bc.out_len * len;
  const int64_t E = g->num_edges;

  // Which CSR is walked. Forward maps given for EDGE targets are indexed by out-CSR position,
  // backward maps by in-CSR position (python/dgl/runtime/spmv.py:146-215; the reference walks
  // the out-CSR forward and the in-CSR backward).
  int walk;
  if (dst_tgt == kEdge) {
    walk = (c.mode == kFwd) ? kWalkOut : kWalkIn;  // the CSR the maps are indexed by
  } else if (dst_tgt == kSrc) {
    walk = kWalkOut;
  } else {
    // forward reduce: reference accumulation order (see DGLB200Graph::in_s); gradients of a
    // dst-target operand: the reference walks the in-CSR itself
    walk = (c.mode == kFwd) ? kWalkInSorted : kWalkIn;
  }
  const bool use_out_csr = (walk == kWalkOut);
  const bool maps_index_out_csr = (c.mode == kFwd);
  const Csr& csr = walk_csr(g, walk, s);

  GArgs a;
  a.indptr = csr.indptr;
  a.indices = csr.indices;
  a.eids = csr.eids;
  a.row = nullptr;
  a.n_rows = (int)csr.n_rows;
  a.n_pos = (int)E;
  a.row_is_dst = use_out_csr ? 0 : 1;
  a.lhs_tgt = c.lhs_tgt;
  a.rhs_tgt = c.rhs_tgt;
  a.out_tgt = out_tgt;
  a.lhs = c.lhs;
  a.rhs = c.rhs;
  a.out = c.out;
  a.gout = c.gout;
  a.dst = c.dst;
  a.op = c.op;
  a.red = c.red;
  a.mode = c.mode;
  a.n_elem = (int)n_elem;
  a.need_e = (c.red == kMax || c.red == kMin || c.red == kProd) ? 1 : 0;
  a.scale_by_deg = c.mean ? 1.f : 0.f;
  a.bc = c.bc;

  // id arrays; edge maps are re-indexed if they were given for the other CSR
  Scratch tmp[3];
  const int32_t* maps[3] = {c.lhs_map, c.rhs_map, c.out_map};
  const int tgts[3] = {c.lhs_tgt, c.rhs_tgt, out_tgt};
  const int32_t* ids[3];
  for (int i = 0; i < 3; ++i) {
    ids[i] = maps[i];
    if (maps[i] && tgts[i] == kEdge && maps_index_out_csr != use_out_csr && E > 0) {
      const int32_t* to_other = csr_to_other(g, walk, s);
      tmp[i].alloc((size_t)E * 4, s);
      int blocks = (int)std::min<int64_t>((E + 255) / 256, 148 * 16);
      compose_map_kernel<<<blocks, 256, 0, s>>>(maps[i], to_other, tmp[i].as<int32_t>(), E);
      note_kernel(nullptr, 1);
      ids[i] = tmp[i].as<int32_t>();
    }
  }
  a.lhs_ids = ids[0];
  a.rhs_ids = ids[1];
  a.out_ids = ids[2];

  // Identity fill (binary_reduce_impl.h:62,137,143) only where some destination row may not be
  // written by its owner: a destination id map, or more tensor rows than CSR rows / edges.
  const int32_t* dmap = c.mode == kFwd ? c.out_map : (c.mode == kGradLhs ? c.lhs_map : c.rhs_map);
  const int64_t owner_rows = (dst_tgt == kEdge) ? E : csr.n_rows;
  const float fill_v = (c.mode == kFwd) ? red_identity_host(c.red) : 0.f;
  if (dmap != nullptr || c.dst_rows != owner_rows) fill_f32(c.dst, c.dst_rows * n_elem, fill_v, s);

  if (n_elem == 0) return;
  if (dst_tgt == kEdge) {
    if (E == 0) return;
    a.row = csr_rows(g, walk, s);
    const int lanes = pow2_at_most(n_elem, 32);
    const int pos_per_block = 256 / lanes;
    const int blocks = (int)std::min<int64_t>((E + pos_per_block - 1) / pos_per_block, 148 * 64);
    generic_edge_kernel<<<blocks, 256, 0, s>>>(a, lanes);
    note_kernel("generic_edge", 1);
  } else {
    if (csr.n_rows == 0) return;
    const int lanes = pow2_at_most(n_elem, 32);
    const int rows_per_block = 256 / lanes;
    const int blocks =
        (int)std::min<int64_t>((csr.n_rows + rows_per_block - 1) / rows_per_block, 148 * 64);
    generic_row_kernel<<<blocks, 256, 0, s>>>(a, lanes);
    note_kernel("generic_row", 1);
  }
}

}  // namespace b200

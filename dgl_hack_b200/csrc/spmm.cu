// SpMM-style gather-reduce over CSR rows (the hot kernel).
//
// Replaces, for node-output reductions, minigun's edge-parallel advance + atomics
// (third_party/minigun/minigun/cuda/advance_all.cuh:38-91 with the UDF
// src/kernel/cuda/binary_reduce_impl.cuh:22-53) and the cuSPARSE csrmm2 + cuBLAS geam transpose
// special case for copy_src+sum (src/kernel/cuda/binary_reduce_sum.cu:83-143, 228-276):
//
//   out[r, :] = REDUCE_{p in row r}  A[ida(p), :]  (op)  B[idb(p), :]
//
// over the in-CSR (forward; rows = destination nodes) or the out-CSR (gradients w.r.t. a
// source-target operand; rows = source nodes, A = grad_out gathered by destination).
//
// Mapping: the feature row is cut into VEC-wide vector columns; a lane owns CH vector columns and
// keeps their accumulators in registers for the whole row, so each output element is reduced
// sequentially in CSR order (bit-reproducible, no atomics, no identity pre-fill, each output row
// written exactly once with 64/128-bit stores). A row is served by LPR lanes (a sub-warp for narrow
// features, a full warp, or several warps = "slabs" for wide features). The row's neighbour ids
// are fetched by the lanes cooperatively (one coalesced load per LPR edges) and broadcast with
// shuffles; feature loads of U consecutive neighbours are issued back to back before they are
// consumed so U*CH independent 8/16-byte loads are in flight per lane.
#include <algorithm>

#include <cstdlib>
#include <cstring>

#include "spmm_kernels.cuh"
#include "spmm_tma_kernels.cuh"

namespace b200 {

extern template void launch_vec<1>(const SpmmArgs&, int, int, bool, int, cudaStream_t);
extern template void launch_vec<2>(const SpmmArgs&, int, int, bool, int, cudaStream_t);
extern template void launch_vec<4>(const SpmmArgs&, int, int, bool, int, cudaStream_t);
extern template void launch_tma_vec<1>(const TmaArgs&, int, bool, bool, int, int, size_t, cudaStream_t);
extern template void launch_tma_vec<2>(const TmaArgs&, int, bool, bool, int, int, size_t, cudaStream_t);
extern template void launch_tma_vec<4>(const TmaArgs&, int, bool, bool, int, int, size_t, cudaStream_t);

// B200_SPMM_PATH=reg|tma forces one SpMM implementation (profiling / A-B runs); default: auto.
static int forced_path() {
  const char* e = std::getenv("B200_SPMM_PATH");
  if (!e) return 0;
  if (!std::strcmp(e, "reg")) return 1;
  if (!std::strcmp(e, "tma")) return 2;
  return 0;
}

// Edges per segment of a heavy row. B200_SPMM_SEG=<n> overrides (0 = never split rows).
// Default: the longest sequential walk may take about a quarter of the work one resident warp
// slot gets (E / (148 SMs x 32 warps x 4)), rounded up to a power of two, at least 4096 -- rows
// up to that many edges are reduced strictly sequentially (bit-exact vs the reference).
// B200_SPMM_WIDE=0 (or B200_SPMM_SEG=0, "strictly sequential") keeps every sub-warp row on one
// lane group.
static bool wide_rows_enabled() {
  const char* e = std::getenv("B200_SPMM_WIDE");
  if (e && e[0] == '0') return false;
  const char* g = std::getenv("B200_SPMM_SEG");
  return !(g && std::atol(g) <= 0);
}

int segment_length(int64_t num_edges) {
  const char* e = std::getenv("B200_SPMM_SEG");
  if (e) {
    const long v = std::atol(e);
    return v <= 0 ? 0x7fffffff : (int)std::min<long>(v, 0x7fffffff);
  }
  int len = 4096;
  while (len < 65536 && (int64_t)len * (148 * 32 * 4) < num_edges) len <<= 1;
  return len;
}

// Segment length for one SpMM call. A row (or segment) is a chain of dependent steps of U
// neighbours, ~1 us each under load (measured: a 462-edge row alone took 70 % of a 0.16 ms
// kernel on the arxiv-shaped graph), while the rest of the grid finishes in about
// bytes / 6 TB/s. The chain may cost a quarter of that: small graphs get short segments (down to
// 64 edges), big ones keep the E-based length. Rows up to the segment length are reduced
// strictly sequentially (bit-identical to the single-threaded reference); B200_SPMM_SEG pins it.
static int segment_length_for(int64_t num_edges, int64_t row_bytes, int unroll) {
  const int base = segment_length(num_edges);
  if (std::getenv("B200_SPMM_SEG")) return base;
  const double bulk_us = (double)num_edges * (double)row_bytes / 6.0e6;
  const double target = 0.25 * bulk_us * unroll;
  int len = 64;
  for (int cand : {512, 4096})
    if (cand <= target) len = cand;
  if (base <= target || len > base) len = base;
  return len;
}

// out[row] = fold of the row's partial rows in segment order (+ mean division)
__global__ void __launch_bounds__(128)
combine_partials_kernel(const int32_t* __restrict__ heavy_row,
                        const int32_t* __restrict__ heavy_ptr, const float* __restrict__ partial,
                        const int32_t* __restrict__ indptr, float* __restrict__ out,
                        const int32_t* __restrict__ out_map, int D, int red, int mean,
                        const float* __restrict__ row_scale, const float* addend, int out_ld,
                        int out_cols) {
  const int h = blockIdx.x;
  const int r = heavy_row[h];
  const int p0 = heavy_ptr[h], p1 = heavy_ptr[h + 1];
  const int64_t orow = out_map ? out_map[r] : r;
  const int deg = indptr[r + 1] - indptr[r];
  const float dn = (float)(deg > 1 ? deg : 1);
  const float rs = row_scale ? row_scale[r] : 1.f;
  for (int col = threadIdx.x; col < out_cols; col += blockDim.x) {
    float acc = partial[(int64_t)p0 * D + col];
    for (int k = p0 + 1; k < p1; ++k) acc = red_call(red, acc, partial[(int64_t)k * D + col]);
    if (addend) acc = __fadd_rn(addend[orow * out_ld + col], acc);
    if (mean) acc = __fdiv_rn(acc, dn);
    if (row_scale) acc = __fmul_rn(acc, rs);
    out[orow * out_ld + col] = acc;
  }
}

// Short rows (relation-segment graphs: ~1 edge per (relation, node) row; the halo-source pass of a
// partitioned aggregation: ~3): the general kernel spends ~330 instructions per warp on a batch of
// such rows (schedule lookup, id batches, shuffles, unrolled predication) and is issue bound at
// 46 % with DRAM at 30 % (profiles/r2_c4_ncu_summary.json). Here a lane group walks its row
// directly -- every lane reads the neighbour id itself (one broadcast transaction), four
// neighbours' rows in flight -- in the same sequential order, so results are bit-identical.
// Plain copy + sum (+ a scale per gathered row: a.B indexed by the neighbour id, row scale,
// addend); rows longer than seg_len are left to their segments.
template <int VEC, int CH, bool SCALED>
__global__ void __launch_bounds__(256) spmm_short_rows_kernel(const SpmmArgs a) {
  const int lane = threadIdx.x & 31;
  const int lpr = a.lpr;
  const int sub = lane & (lpr - 1);
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t r = warp * (32 / lpr) + lane / lpr;
  if (r >= a.n_rows) return;
  const int beg = __ldg(a.indptr + r), end = __ldg(a.indptr + r + 1);
  if (a.n_seg > 0 && end - beg > a.seg_len) return;  // heavy row: served by its segments
  const int cbase = sub * VEC;
  Vec<VEC> acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[c].v[e] = 0.f;
  auto col_ok = [&](int c) -> bool { return sub + c * lpr < a.nvec; };
  const float* base = a.A + cbase;
  // per step: four ids, then their rows AND scales issued back to back, then the sums in order
  // (fetching the next step's ids ahead was measured and dropped: it helps 25-edge rows by 10 %
  // but costs rows with 1-7 edges 20-30 % in wasted predicated loads)
  int p = beg;
  for (; p + 4 <= end; p += 4) {
    int id[4];
    Vec<VEC> v[4][CH];
    float sc[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) id[k] = __ldg(a.indices + p + k);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      sc[k] = SCALED ? __ldg(a.B + id[k]) : 1.f;
#pragma unroll
      for (int c = 0; c < CH; ++c)
        if (col_ok(c)) v[k][c] = ld_vec<VEC>(base + (int64_t)id[k] * a.D + c * lpr * VEC);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
      for (int c = 0; c < CH; ++c)
        if (col_ok(c))
#pragma unroll
          for (int e = 0; e < VEC; ++e) {
            const float val = SCALED ? __fmul_rn(v[k][c].v[e], sc[k]) : v[k][c].v[e];
            acc[c].v[e] = __fadd_rn(acc[c].v[e], val);
          }
    }
  }
  {  // up to three left: loads first, then the sums in order
    const int left = end - p;
    int id[3];
    Vec<VEC> v[3][CH];
    float sc[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      if (k < left) id[k] = __ldg(a.indices + p + k);
#pragma unroll
    for (int k = 0; k < 3; ++k)
      if (k < left) {
        sc[k] = SCALED ? __ldg(a.B + id[k]) : 1.f;
#pragma unroll
        for (int c = 0; c < CH; ++c)
          if (col_ok(c)) v[k][c] = ld_vec<VEC>(base + (int64_t)id[k] * a.D + c * lpr * VEC);
      }
#pragma unroll
    for (int k = 0; k < 3; ++k)
      if (k < left) {
#pragma unroll
        for (int c = 0; c < CH; ++c)
          if (col_ok(c))
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
              const float val = SCALED ? __fmul_rn(v[k][c].v[e], sc[k]) : v[k][c].v[e];
              acc[c].v[e] = __fadd_rn(acc[c].v[e], val);
            }
      }
  }
  float* op = a.out + r * a.out_ld + cbase;
  const float* addp = a.addend ? a.addend + r * a.out_ld + cbase : nullptr;
  const float rs = a.row_scale ? __ldg(a.row_scale + r) : 1.f;
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    if (col_ok(c)) {
      if (addp) {
        const Vec<VEC> old = ld_vec<VEC>(addp + c * lpr * VEC);
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[c].v[e] = __fadd_rn(old.v[e], acc[c].v[e]);
      }
      if (a.row_scale) {
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[c].v[e] = __fmul_rn(acc[c].v[e], rs);
      }
      if (a.out_cols == a.D) {
        st_vec<VEC>(op + c * lpr * VEC, acc[c]);
      } else {
#pragma unroll
        for (int e = 0; e < VEC; ++e)
          if (cbase + c * lpr * VEC + e < a.out_cols) op[c * lpr * VEC + e] = acc[c].v[e];
      }
    }
  }
}

// B200_SPMM_SHORT=0 keeps short-row graphs on the general kernel (A/B runs);
// B200_SPMM_SHORT_DEG=<n> moves the mean-degree threshold (default 6).
static bool short_rows_enabled() {
  const char* e = std::getenv("B200_SPMM_SHORT");
  return !(e && e[0] == '0');
}
// Measured (profiles/r2_short_rows_ab.json): whole-warp rows (D >= 100) gain up to ~7 edges per
// row (arxiv shape D=128: 0.144 -> 0.114 ms) and break even at 25; sub-warp rows (D <= 64) still
// gain at 25 edges per row (products shape D=48: 3.25 -> 2.43 ms), where the general kernel pays
// its id batches and shuffles once per lane group.
static int64_t short_rows_max_degree(int lpr) {
  const char* e = std::getenv("B200_SPMM_SHORT_DEG");
  return e ? std::max(1, std::atoi(e)) : (lpr == 32 ? 8 : 32);
}

static int round_tma_ch(int ch) {
  const int allowed[] = {1, 2, 3, 4, 5, 6, 8, 10, 12, 16};
  for (int v : allowed)
    if (ch <= v) return v;
  return -1;
}

static int round_ch(int ch) {
  const int allowed[] = {1, 2, 3, 4, 5, 6, 8, 10, 12};
  for (int v : allowed)
    if (ch <= v) return v;
  return 12;
}

__global__ void compose_ids_kernel(const int32_t* __restrict__ map,
                                   const int32_t* __restrict__ to_other,
                                   int32_t* __restrict__ out, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride)
    out[k] = map[to_other[k]];
}

// out[r, c] = in[r, c] for c < cols (row pitches ld_in / ld_out), 0 for cols <= c < ld_out:
// pads rows to a 16-byte multiple. One item = four output columns (one 16-byte store).
__global__ void repitch_rows_kernel(const float* __restrict__ in, int64_t ld_in,
                                    float* __restrict__ out, int64_t ld_out, int64_t rows,
                                    int64_t cols) {
  const int nv = (int)(ld_out / 4);
  const int64_t total = rows * nv;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / nv;
    const int c0 = (int)(i - r * nv) * 4;
    const float* src = in + r * ld_in + c0;
    float4 v;
    v.x = c0 + 0 < cols ? __ldg(src + 0) : 0.f;
    v.y = c0 + 1 < cols ? __ldg(src + 1) : 0.f;
    v.z = c0 + 2 < cols ? __ldg(src + 2) : 0.f;
    v.w = c0 + 3 < cols ? __ldg(src + 3) : 0.f;
    *reinterpret_cast<float4*>(out + r * ld_out + c0) = v;
  }
}

// ld_out % 4 == 0 and out 16-byte aligned (a Scratch buffer)
static void repitch_rows(const float* in, int64_t ld_in, float* out, int64_t ld_out, int64_t rows,
                         int64_t cols, cudaStream_t s) {
  if (rows <= 0) return;
  const int blocks = (int)std::min<int64_t>((rows * (ld_out / 4) + 255) / 256, 148 * 32);
  repitch_rows_kernel<<<blocks, 256, 0, s>>>(in, ld_in, out, ld_out, rows, cols);
  note_kernel(nullptr, 1);
}

// B200_SPMM_PAD=0 keeps odd widths on the scalar-load kernels (A/B runs).
static bool pad_enabled() {
  const char* e = std::getenv("B200_SPMM_PAD");
  return !(e && e[0] == '0');
}

// Describes how the narrow operand lines up against the D-wide one. Returns kBNone on "cannot".
static int classify_b(const Bcast& bc, bool b_is_lhs, int* b_width, int* head_div) {
  const int* bs = b_is_lhs ? bc.lhs_shape : bc.rhs_shape;
  const int* as = b_is_lhs ? bc.rhs_shape : bc.lhs_shape;
  const int64_t blen = b_is_lhs ? bc.lhs_len : bc.rhs_len;
  const int64_t alen = b_is_lhs ? bc.rhs_len : bc.lhs_len;
  if (alen != bc.out_len) return -1;
  if (blen == bc.out_len) {
    // same length: element-wise only if no dimension is broadcast
    for (int d = 0; d < bc.ndim; ++d)
      if (bs[d] != bc.out_shape[d] || as[d] != bc.out_shape[d]) return -1;
    *b_width = (int)blen;
    *head_div = 1;
    return kBFull;
  }
  if (blen == 1) {  // [.,D...] x [.,1]: one scalar per neighbour
    *b_width = 1;
    *head_div = (int)bc.out_len;
    return kBScalar;
  }
  if (bc.ndim == 2 && bs[1] == 1 && bs[0] == bc.out_shape[0] && as[0] == bc.out_shape[0] &&
      as[1] == bc.out_shape[1]) {  // [.,H,F] x [.,H,1]
    *b_width = bs[0];
    *head_div = bc.out_shape[1];
    return kBHead;
  }
  return -1;
}

bool try_spmm(const KernelCall& c) {
  DGLB200Graph* g = c.g;
  cudaStream_t s = c.stream;
  if (c.op == kDot || c.bc.data_len != 1) return false;
  const int64_t D = c.bc.out_len;
  if (D <= 0 || D > (1 << 20)) return false;
  const int64_t E = g->num_edges;

  SpmmArgs a;
  a.B = nullptr;
  a.b_src = kFromCol;
  a.b_pos_ids = nullptr;
  a.b_node_map = nullptr;
  a.b_width = 0;
  a.b_head_div = 1;
  a.b_is_lhs = 0;
  a.a_pos_ids = nullptr;
  a.a_node_map = nullptr;
  a.mean = c.mean ? 1 : 0;
  a.row_scale = c.row_scale;
  a.red = c.red;
  a.op = c.op;
  a.write_empty = 0;
  for (int i = 0; i < kMaxSlots; ++i) a.a_tab[i] = (c.a_tab && i < c.n_slots) ? c.a_tab[i] : nullptr;
  a.addend = c.addend;
  a.out_ld = a.out_cols = 0;  // set below, once D is known
  int bmode = kBNone;
  bool use_out_csr, walk_in_plain = false;
  // (tensor, target, user map) of the two gathered operands in "row/col of the walked CSR" terms
  const float *A = nullptr, *B = nullptr;
  int a_tgt = kNone, b_tgt = kNone;
  const int32_t *a_map = nullptr, *b_map = nullptr, *o_map = nullptr;
  float fill_v = 0.f;

  if (c.mode == kFwd) {
    if (c.red == kRNone) return false;
    use_out_csr = false;
    fill_v = red_identity_host(c.red);
    o_map = c.out_map;
    if (c.op == kUseLhs) {
      A = c.lhs; a_tgt = c.lhs_tgt; a_map = c.lhs_map;
    } else {
      // pick the D-wide, gatherable operand as A
      int bw, hd;
      const bool lhs_can_be_a = c.bc.lhs_len == D && c.lhs_tgt != kDst;
      const bool rhs_can_be_a = c.bc.rhs_len == D && c.rhs_tgt != kDst;
      int m = -1;
      if (lhs_can_be_a && (m = classify_b(c.bc, /*b_is_lhs=*/false, &bw, &hd)) > 0) {
        A = c.lhs; a_tgt = c.lhs_tgt; a_map = c.lhs_map;
        B = c.rhs; b_tgt = c.rhs_tgt; b_map = c.rhs_map;
        a.b_is_lhs = 0;
      } else if (rhs_can_be_a && (m = classify_b(c.bc, /*b_is_lhs=*/true, &bw, &hd)) > 0) {
        A = c.rhs; a_tgt = c.rhs_tgt; a_map = c.rhs_map;
        B = c.lhs; b_tgt = c.lhs_tgt; b_map = c.lhs_map;
        a.b_is_lhs = 1;
      } else {
        return false;
      }
      bmode = m;
      a.b_width = bw;
      a.b_head_div = hd;
    }
  } else {
    const int t_tgt = c.mode == kGradLhs ? c.lhs_tgt : c.rhs_tgt;
    if (c.red == kRNone) {
      // edge outputs (apply_edges u_add_v, e_sub_v, copies to edges ...): the gradient of a NODE
      // operand with a +-1 derivative is copy_e + sum of grad_out over the node's out-edges
      // (source operand) or in-edges (destination operand, in-CSR order as the reference walks it)
      if (t_tgt == kEdge || !(c.op == kUseLhs || c.op == kAdd || c.op == kSub)) return false;
      if (c.op == kUseLhs && c.mode == kGradRhs) return false;
      use_out_csr = (t_tgt == kSrc);
      walk_in_plain = !use_out_csr;
      o_map = c.mode == kGradLhs ? c.lhs_map : c.rhs_map;
      A = c.gout; a_tgt = kEdge; a_map = c.out_map;
      a.red = kSum;
      a.op = (c.op == kSub && c.mode == kGradRhs) ? kOpNegA : kUseLhs;
    } else {
    // gradient of a SOURCE-target operand under a sum reduction: a transposed SpMM of grad_out
    if (c.red != kSum) return false;
    if (t_tgt != kSrc) return false;
    use_out_csr = true;
    o_map = c.mode == kGradLhs ? c.lhs_map : c.rhs_map;
    A = c.gout; a_tgt = kDst; a_map = c.out_map;  // gathered by the destination of each out-edge
    if (c.op == kUseLhs || c.op == kAdd || (c.op == kSub && c.mode == kGradLhs)) {
      a.op = kUseLhs;
    } else if (c.op == kSub) {
      a.op = kOpNegA;
    } else if (c.op == kMul) {
      int bw, hd;
      // the OTHER operand multiplies grad_out; it must line up against the out feature row
      const bool other_is_lhs = (c.mode == kGradRhs);
      Bcast sw = c.bc;
      // classify with "A" = a full-width pseudo operand: pretend the differentiated operand has
      // the out shape (grad is unreduced), keep the other operand's real shape.
      if (other_is_lhs) {
        for (int d = 0; d < sw.ndim; ++d) sw.rhs_shape[d] = sw.out_shape[d];
        sw.rhs_len = sw.out_len;
      } else {
        for (int d = 0; d < sw.ndim; ++d) sw.lhs_shape[d] = sw.out_shape[d];
        sw.lhs_len = sw.out_len;
      }
      const int m = classify_b(sw, other_is_lhs, &bw, &hd);
      if (m <= 0) return false;
      bmode = m;
      a.b_width = bw;
      a.b_head_div = hd;
      B = other_is_lhs ? c.lhs : c.rhs;
      b_tgt = other_is_lhs ? c.lhs_tgt : c.rhs_tgt;
      b_map = other_is_lhs ? c.lhs_map : c.rhs_map;
      a.op = kMul;
    } else {
      return false;
    }
    }
  }

  if (c.a_tab) {  // peer-table gather: plain copy + sum through the operand's id map only
    B200_CHECK(bmode == kBNone && a.red == kSum && a.op == kUseLhs && a_map != nullptr &&
                   a_tgt != kEdge && c.n_slots >= 1 && c.n_slots <= kMaxSlots && c.a_tab[0],
               "peer-table gather supports copy + sum with a (slot, row) id map only");
  }
  // forward: reference sum order
  const int walk = use_out_csr ? kWalkOut : (walk_in_plain ? kWalkIn : kWalkInSorted);
  const Csr& csr = walk_csr(g, walk, s);
  const int row_tgt = use_out_csr ? kSrc : kDst;
  const int col_tgt = use_out_csr ? kDst : kSrc;
  const bool maps_index_out_csr = (c.mode == kFwd);
  if (a_tgt == row_tgt) return false;  // A must be gathered per edge
  if (csr.n_rows >= (1ll << 31) || E >= (1ll << 31)) return false;

  // ---- odd widths: rows whose pitch is not a multiple of 16 bytes (D = 47: 188-byte rows) can
  // only be gathered with 4-byte loads (0.41 of the HBM peak at the products shape against 0.8 at
  // D = 48). Gather from a zero-padded copy with 16-byte rows instead and drop the padding from
  // the result: two streaming passes over N x D against the E x D gather.
  if (D % 4 != 0 && D >= 8 && c.bc.ndim == 1 && (bmode == kBNone || bmode == kBScalar) &&
      !c.a_tab && !c.addend && pad_enabled() && forced_path() != 2 && E >= 4 * csr.n_cols) {
    const int64_t Dp = (D + 3) / 4 * 4;
    const int which = c.mode != kFwd ? 2 : (A == c.lhs ? 0 : 1);
    const int64_t a_rows = which == 2 ? c.out_rows : (which == 0 ? c.lhs_rows : c.rhs_rows);
    Scratch a_pad((size_t)std::max<int64_t>(a_rows, 1) * Dp * 4, s);
    repitch_rows(A, D, a_pad.as<float>(), Dp, a_rows, D, s);
    KernelCall c2 = c;
    if (which == 0) c2.lhs = a_pad.as<float>();
    else if (which == 1) c2.rhs = a_pad.as<float>();
    else c2.gout = a_pad.as<float>();
    // the result goes straight into the caller's rows (pitch D, 4-byte stores): one streaming
    // pass over the operand is all the padding costs
    c2.dst_ld = D;
    c2.dst_cols = D;
    auto widen = [&](int* shape, int64_t* len) {
      if (*len == D) { shape[0] = (int)Dp; *len = Dp; }
    };
    widen(c2.bc.lhs_shape, &c2.bc.lhs_len);
    widen(c2.bc.rhs_shape, &c2.bc.rhs_len);
    widen(c2.bc.out_shape, &c2.bc.out_len);
    if (try_spmm(c2)) return true;
  }

  // ---- vector width from divisibility and pointer alignment
  auto aligned = [](const void* p, int bytes) {
    return (reinterpret_cast<uintptr_t>(p) % bytes) == 0;
  };
  int vec = 4;
  while (vec > 1) {
    // odd-pitch output rows (dst_cols set) are written with 4-byte stores: no alignment needed
    bool ok = (D % vec == 0) && aligned(A, vec * 4) && (c.dst_cols || aligned(c.dst, vec * 4));
    if (bmode == kBFull) ok = ok && aligned(B, vec * 4);
    if (bmode == kBHead) ok = ok && (a.b_head_div % vec == 0);
    for (int i = 0; c.a_tab && i < c.n_slots; ++i) ok = ok && aligned(c.a_tab[i], vec * 4);
    if (ok) break;
    vec >>= 1;
  }
  const int nvec = (int)(D / vec);
  int lpr = 1;
  while (lpr < 32 && lpr < nvec) lpr <<= 1;
  int slabs = 1, ch = 1;
  if (lpr == 32) {
    slabs = (nvec + 32 * 12 - 1) / (32 * 12);
    ch = round_ch((nvec + 32 * slabs - 1) / (32 * slabs));
    slabs = (nvec + 32 * ch - 1) / (32 * ch);
  }

  // ---- id sources
  Scratch tmp_a, tmp_b;
  auto edge_ids = [&](const int32_t* map, Scratch& tmp) -> const int32_t* {
    if (!map) return csr.eids;
    if (maps_index_out_csr == use_out_csr || E == 0) return map;
    const int32_t* to_other = csr_to_other(g, walk, s);
    tmp.alloc((size_t)E * 4, s);
    int blocks = (int)std::min<int64_t>((E + 255) / 256, 148 * 16);
    compose_ids_kernel<<<blocks, 256, 0, s>>>(map, to_other, tmp.as<int32_t>(), E);
    note_kernel(nullptr, 1);
    return tmp.as<int32_t>();
  };
  a.A = A;
  if (a_tgt == kEdge) {
    a.a_src = kFromEid;
    a.a_pos_ids = edge_ids(a_map, tmp_a);
  } else {  // col_tgt
    a.a_src = kFromCol;
    a.a_node_map = a_map;
  }
  a.B = B;
  if (bmode != kBNone) {
    if (b_tgt == kEdge) {
      a.b_src = kFromEid;
      a.b_pos_ids = edge_ids(b_map, tmp_b);
    } else if (b_tgt == col_tgt) {
      a.b_src = kFromCol;
      a.b_node_map = b_map;
    } else {
      a.b_src = kFromRow;
      a.b_node_map = b_map;
    }
  }
  a.need_col = (a.a_src == kFromCol) || (bmode != kBNone && a.b_src == kFromCol);

  a.indptr = csr.indptr;
  a.indices = csr.indices;
  a.n_rows = (int)csr.n_rows;
  a.lpr = lpr;
  a.slabs = slabs;
  a.D = (int)D;
  a.nvec = nvec;
  a.out = c.dst;
  a.out_map = o_map;
  a.out_ld = (int)(c.dst_ld ? c.dst_ld : D);
  a.out_cols = (int)(c.dst_cols ? c.dst_cols : D);

  B200_CHECK(!c.addend || (a.red == kSum && !c.mean && o_map == nullptr && c.dst_rows == csr.n_rows),
             "an addend needs the sum reducer and one output row per reducing node");
  if (o_map != nullptr || c.dst_rows != csr.n_rows)
    fill_f32(c.dst, c.dst_rows * (int64_t)a.out_ld, fill_v, s);
  if (csr.n_rows == 0) return true;

  // ---- heavy rows (hubs of power-law graphs) are cut into segments of seg_len edges, each
  // reduced by its own warp into a partial row; combine_partials_kernel folds the partials of a
  // row in segment order. Rows up to seg_len edges keep the strictly sequential (bit-exact vs the
  // single-threaded reference) reduction; B200_SPMM_SEG=0 disables segmentation altogether.
  const int64_t warp_bytes = (lpr == 32) ? (int64_t)ch * vec * 128 : D * 4;
  const int unroll = warp_bytes >= 1024 ? 1 : (warp_bytes >= 256 ? 4 : 8);
  // whole-warp rows (lpr == 32) do not run in lockstep with other rows: their short rows may
  // keep the natural order on graphs with a local numbering (graph.cu natural_keys_kernel)
  const SegPlan plan =
      csr_segments(g, walk, segment_length_for(E, D * 4, unroll), s, /*want_natural=*/lpr == 32);
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
  auto combine = [&]() {
    if (plan.n_seg == 0) return;
    combine_partials_kernel<<<plan.n_heavy, 128, 0, s>>>(
        plan.heavy_row, plan.heavy_ptr, partial.as<float>(), csr.indptr, c.dst, o_map, (int)D,
        a.red, a.mean, a.row_scale, a.addend, a.out_ld, a.out_cols);
    note_kernel(nullptr, 1);
  };

  // ---- TMA bulk-copy staged path: full-warp rows, plain copy or scalar-per-edge weight
  {
    const int force = forced_path();
    const bool scalar_w = (bmode == kBScalar && a.b_src == kFromEid);
    const int tch = round_tma_ch((nvec + 31) / 32);
    const int64_t row_bytes = D * 4;
    bool ok = (lpr == 32) && (bmode == kBNone || scalar_w) && tch > 0 && row_bytes <= 8192 &&
              aligned(A, 16) && o_map == nullptr && c.row_scale == nullptr &&  // mapped / scaled: LDG path
              c.a_tab == nullptr && c.addend == nullptr;
    const int64_t mean_deg = E / std::max<int64_t>(csr.n_rows, 1);
    // Measured on B200 (profiles/sweep1.txt, profiles/microbench_v1.txt): the LDG path beats the bulk-copy path on
    // this gather (TMA 1-D copies of 2.4 KB rows top out near 20-30 B/clk/SM), so it is opt-in.
    (void)mean_deg;
    const bool want = (force == 2);
    if (ok && want) {
      TmaArgs t;
      t.indptr = csr.indptr;
      t.a_pos_ids = (a.a_src == kFromCol) ? csr.indices : a.a_pos_ids;
      t.a_node_map = (a.a_src == kFromCol) ? a.a_node_map : nullptr;
      t.n_seg = a.n_seg;
      t.seg_len = a.seg_len;
      t.seg_row = a.seg_row;
      t.seg_beg = a.seg_beg;
      t.seg_end = a.seg_end;
      t.sched = a.sched;
      t.partial = a.partial;
      t.n_rows = a.n_rows;
      t.D = (int)D;
      t.nvec = nvec;
      t.A = A;
      t.a_rows = (A == c.lhs) ? c.lhs_rows : ((A == c.rhs) ? c.rhs_rows : c.out_rows);
      t.W = scalar_w ? B : nullptr;
      t.w_pos_ids = scalar_w ? a.b_pos_ids : nullptr;
      t.w_is_lhs = a.b_is_lhs;
      t.op = a.op;
      t.red = a.red;
      t.mean = a.mean;
      t.out = a.out;
      t.out_map = a.out_map;
      t.slot_bytes = (int)((row_bytes + 30) & ~(int64_t)15);
      int twarps = 8;
      t.stages = 2;
      if (const char* e = std::getenv("B200_TMA_WARPS")) twarps = std::max(1, std::min(16, std::atoi(e)));
      if (const char* e = std::getenv("B200_TMA_STAGES")) t.stages = std::max(2, std::min(16, std::atoi(e)));
      while (twarps > 1 && (size_t)twarps * t.stages * (t.slot_bytes + 8) > 200 * 1024) twarps >>= 1;
      const size_t smem = (size_t)twarps * t.stages * (t.slot_bytes + 8);
      const int tblocks = (int)((n_items + twarps - 1) / twarps);
      const bool tsum = (a.red == kSum);
      if (vec == 4) launch_tma_vec<4>(t, tch, tsum, scalar_w, tblocks, twarps, smem, s);
      else if (vec == 2) launch_tma_vec<2>(t, tch, tsum, scalar_w, tblocks, twarps, smem, s);
      else launch_tma_vec<1>(t, tch, tsum, scalar_w, tblocks, twarps, smem, s);
      note_kernel("spmm_tma", 1);
      combine();
      return true;
    }
  }

  // ---- short rows: lean kernel for the rows, the general one for the heavy rows' segments only
  const bool col_scaled = bmode == kBScalar && a.op == kMul && a.b_src == kFromCol &&
                          a.b_node_map == nullptr;  // GraphConv's source norm inside the sum
  if (short_rows_enabled() && ((bmode == kBNone && a.op == kUseLhs) || col_scaled) &&
      a.red == kSum && !a.mean &&
      a.a_src == kFromCol && a.a_node_map == nullptr && c.a_tab == nullptr && o_map == nullptr &&
      c.dst_rows == csr.n_rows && vec == 4 && slabs == 1 && ch <= 2 &&
      E <= short_rows_max_degree(lpr) * csr.n_rows && forced_path() == 0) {
    if (plan.n_seg > 0) {
      SpmmArgs h = a;
      h.n_rows = 0;  // items = segments only
      h.n_wide = (lpr < 32 && wide_rows_enabled()) ? plan.n_seg : 0;
      const int64_t hw = (lpr == 32) ? (int64_t)plan.n_seg
                                     : h.n_wide + (plan.n_seg - h.n_wide + (32 / lpr) - 1) / (32 / lpr);
      launch_vec<4>(h, ch, bmode, true, (int)((hw + 7) / 8), s);
      note_kernel(nullptr, 1);
    }
    const int rpw = 32 / lpr;
    const int64_t sw = (csr.n_rows + rpw - 1) / rpw;
    const int T = spmm_block_threads();
    const int sblocks = (int)((sw * 32 + T - 1) / T);
    if (col_scaled) {
      if (ch == 1) spmm_short_rows_kernel<4, 1, true><<<sblocks, T, 0, s>>>(a);
      else spmm_short_rows_kernel<4, 2, true><<<sblocks, T, 0, s>>>(a);
    } else {
      if (ch == 1) spmm_short_rows_kernel<4, 1, false><<<sblocks, T, 0, s>>>(a);
      else spmm_short_rows_kernel<4, 2, false><<<sblocks, T, 0, s>>>(a);
    }
    note_kernel("spmm", 1);
    combine();
    return true;
  }

  // sub-warp rows: segments and (with a degree schedule) rows of >= kLongRow edges get a warp each
  a.n_wide = 0;
  if (lpr < 32 && wide_rows_enabled())
    a.n_wide = plan.n_seg + (plan.sched ? plan.n_long : 0);
  const int64_t n_warps =
      (lpr == 32) ? n_items * slabs
                  : a.n_wide + (n_items - a.n_wide + (32 / lpr) - 1) / (32 / lpr);
  const int blocks = (int)((n_warps + 7) / 8);
  const bool sum = (a.red == kSum);
  if (vec == 4) launch_vec<4>(a, ch, bmode, sum, blocks, s);
  else if (vec == 2) launch_vec<2>(a, ch, bmode, sum, blocks, s);
  else launch_vec<1>(a, ch, bmode, sum, blocks, s);
  note_kernel("spmm", 1);
  combine();
  return true;
}

}  // namespace b200

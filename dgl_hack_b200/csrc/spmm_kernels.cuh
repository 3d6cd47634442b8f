// SpMM kernel template + launchers, included by spmm_v{1,2,4}.cu (one translation unit per
// vector width so the instantiations compile in parallel). See spmm.cu for the design notes.
#pragma once
#include "common.cuh"

namespace b200 {

enum IdSrc { kFromCol = 0, kFromEid = 1, kFromRow = 2 };
// kBHead: one B value per head (column / b_head_div); kBScalar: one B value per neighbour
enum BMode { kBNone = 0, kBFull = 1, kBHead = 2, kBScalar = 3 };
constexpr int kOpNegA = 100;  // internal: value = -A (gradient of the rhs of "sub")
constexpr int kMaxSlots = 16;
constexpr int kSlotShift = 27;  // up to 2^27 rows per slot
constexpr int kRowMask = (1 << kSlotShift) - 1;

struct SpmmArgs {
  const int32_t* indptr;
  const int32_t* indices;
  // work items: [0, n_seg) = segments of heavy rows (partial results into `partial`), then
  // one item per row (heavy rows skip themselves)
  int n_seg;
  int seg_len;
  const int32_t *seg_row, *seg_beg, *seg_end;
  const int32_t* sched;  // order of the row work items (NULL = natural)
  float* partial;  // [n_seg, D]
  int n_rows;
  int lpr;    // lanes per row (power of two <= 32)
  int slabs;  // warps per row when lpr == 32
  int D;      // floats per feature row
  int nvec;   // D / VEC
  const float* A;
  int a_src;
  const int32_t* a_pos_ids;   // a_src == kFromEid: position-indexed ids
  const int32_t* a_node_map;  // a_src == kFromCol: optional node map
  const float* B;
  int b_src;
  const int32_t* b_pos_ids;
  const int32_t* b_node_map;
  int b_width;     // floats per B row
  int b_head_div;  // kBHead: B column = A column / b_head_div
  int b_is_lhs;
  int op, red, mean;
  const float* row_scale;  // optional: every finished output row is multiplied by row_scale[row]
  float* out;
  const int32_t* out_map;
  int need_col;
  // sub-warp rows only (lpr < 32): the first n_wide work items (all segments, then the rows of
  // the schedule with >= kLongRow edges) are served by a whole warp whose 32 / lpr lane groups
  // take every (32 / lpr)-th edge and fold their partial results at the end
  int n_wide;
  // peer-table gather (TAB kernels only; partitioned graphs, csrc/peer.cu): a gathered id is
  // (slot << kSlotShift) | row and reads a_tab[slot] + row * D -- slot = which GPU's staged
  // feature block (own HBM or a peer's, mapped over NVLink), row = the node's local row there
  const float* a_tab[kMaxSlots];
  // out_map is a plain row list (every listed row is written, rows without edges included)
  int write_empty;
  // optional (sum reducer): out[row] = row_scale * (addend[row] + sum), addend laid out like out
  // and allowed to BE out -- the second pass of a two-pass aggregation (partitioned graphs: the
  // halo-source edges added to the inner-source sums) without a separate add kernel
  const float* addend;
  // output rows of pitch out_ld floats of which the first out_cols are written (out_cols < D:
  // the operand was gathered from a copy padded to a 16-byte pitch, the result goes straight
  // into the caller's odd-width rows with 4-byte stores)
  int out_ld, out_cols;
};

template <int VEC>
struct Vec {
  float v[VEC];
};

template <int VEC>
__device__ __forceinline__ Vec<VEC> ld_vec(const float* __restrict__ p);
template <>
__device__ __forceinline__ Vec<1> ld_vec<1>(const float* __restrict__ p) {
  Vec<1> r;
  r.v[0] = __ldg(p);
  return r;
}
template <>
__device__ __forceinline__ Vec<2> ld_vec<2>(const float* __restrict__ p) {
  const float2 t = __ldg(reinterpret_cast<const float2*>(p));
  Vec<2> r;
  r.v[0] = t.x;
  r.v[1] = t.y;
  return r;
}
template <>
__device__ __forceinline__ Vec<4> ld_vec<4>(const float* __restrict__ p) {
  const float4 t = __ldg(reinterpret_cast<const float4*>(p));
  Vec<4> r;
  r.v[0] = t.x;
  r.v[1] = t.y;
  r.v[2] = t.z;
  r.v[3] = t.w;
  return r;
}

template <int VEC>
__device__ __forceinline__ void st_vec(float* __restrict__ p, const Vec<VEC>& x);
template <>
__device__ __forceinline__ void st_vec<1>(float* __restrict__ p, const Vec<1>& x) {
  p[0] = x.v[0];
}
template <>
__device__ __forceinline__ void st_vec<2>(float* __restrict__ p, const Vec<2>& x) {
  *reinterpret_cast<float2*>(p) = make_float2(x.v[0], x.v[1]);
}
template <>
__device__ __forceinline__ void st_vec<4>(float* __restrict__ p, const Vec<4>& x) {
  *reinterpret_cast<float4*>(p) = make_float4(x.v[0], x.v[1], x.v[2], x.v[3]);
}

__device__ __forceinline__ float spmm_value(int op, int b_is_lhs, float av, float bv) {
  if (op == kOpNegA) return __fmul_rn(av, -1.f);
  return b_is_lhs ? op_scalar(op, bv, av) : op_scalar(op, av, bv);
}

// Minimum resident CTAs the compiler should budget registers for: the gather is latency bound
// and wants 32+ resident warps whenever accumulators + one neighbour's loads fit in ~60 registers.
constexpr int min_blocks_for(int vec, int ch, int u, int bmode) {
  const int live = ch * vec * (1 + u * (bmode == kBFull ? 2 : 1)) + (bmode == kBHead ? ch * u : 0);
  // 48 registers = 5 CTAs per SM where accumulators + one neighbour fit 16 registers (D = 256:
  // the scaled variant ran at 54 registers = 4 CTAs, 8.85 ms against 7.4 ms for the plain copy)
  if (vec == 4 && live <= 16) return 5;
  return live <= 40 ? 4 : (live <= 100 ? 2 : 1);
}

// FULL: a row is served by whole warps (lpr == 32) so every column offset is a compile-time
// constant; otherwise lpr (< 32, CH == 1) lanes serve a row and a warp handles 32 / lpr rows.
// MUL: the binary op is known to be "mul" (commutative: operand order irrelevant), so the
// per-element runtime op switch disappears from the inner loop.
// TAB: gathered ids are (slot, row) pairs resolved through a_tab (see SpmmArgs).
template <int VEC, int CH, int U, bool SUM, int BMODE, bool FULL, bool MUL, bool TAB = false>
__global__ void __launch_bounds__(256, min_blocks_for(VEC, CH, U, BMODE))
spmm_kernel(const SpmmArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lpr = FULL ? 32 : a.lpr;
  const int sub = FULL ? lane : (lane & (lpr - 1));

  // ---- work item -> (row or heavy-row segment, slab)
  int64_t item;
  int slab = 0;
  bool wide = false;  // warp-uniform
  if (FULL) {
    item = warp / a.slabs;
    slab = (int)(warp - item * a.slabs);
  } else {
    wide = warp < a.n_wide;
    item = wide ? warp : a.n_wide + (warp - a.n_wide) * (32 / lpr) + lane / lpr;
  }
  // batch width / this lane's slot in the id batch / edge stride and offset of the lane group
  const int bw = (!FULL && wide) ? 32 : lpr;
  const int bl = (!FULL && wide) ? lane : sub;
  const int estride = (!FULL && wide) ? 32 / lpr : 1;
  const int eoff = (!FULL && wide) ? lane / lpr : 0;
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
      } else if (a.out_map && deg == 0 && !a.write_empty) {
        // with an output id map only rows that receive edges have a slot (the reference's
        // _build_idx_map holds -1 / ends early for the others, runtime/scheduler.py): the row
        // keeps the identity written by the fill and the map is never read for it
        row_ok = false;
      }
    }
  }

  // ---- this lane's first vector column; column c of the lane is vc0 + c * lpr
  const int vc0 = slab * (32 * CH) + sub;
  const int cbase = vc0 * VEC;  // float offset of the lane's first column inside a row
  auto col_ok = [&](int c) -> bool { return vc0 + c * lpr < a.nvec; };

  const float ident = SUM ? 0.f : red_identity(a.red);
  Vec<VEC> acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[c].v[e] = ident;

  int hc[CH];  // kBHead: B column of each of the lane's vector columns (loop invariant)
  if (BMODE == kBHead) {
#pragma unroll
    for (int c = 0; c < CH; ++c) hc[c] = (cbase + c * lpr * VEC) / a.b_head_div;
  }
  int64_t b_row_const = 0;
  if (BMODE != kBNone && a.b_src == kFromRow)
    b_row_const = (int64_t)(a.b_node_map ? __ldg(a.b_node_map + r) : r) * a.b_width;

  const int maxdeg = FULL ? deg : __reduce_max_sync(0xffffffffu, deg);
  for (int base = 0; base < maxdeg; base += bw) {
    // ---- cooperative fetch of up to bw neighbour ids per row
    const int my = base + bl;
    int pa = 0, pb = 0;
    float psc = 0.f;  // kBScalar: the neighbour's scalar, fetched with the id batch (one gather per
                      // 32 neighbours instead of a dependent 4-byte load per neighbour)
    if (my < deg) {
      const int p = beg + my;
      int cidx = 0;
      if (a.need_col) cidx = __ldg(a.indices + p);
      if (a.a_src == kFromCol) {
        pa = a.a_node_map ? __ldg(a.a_node_map + cidx) : cidx;
      } else {
        pa = __ldg(a.a_pos_ids + p);
      }
      if (BMODE != kBNone) {
        if (a.b_src == kFromCol) {
          pb = a.b_node_map ? __ldg(a.b_node_map + cidx) : cidx;
        } else if (a.b_src == kFromEid) {
          pb = __ldg(a.b_pos_ids + p);
        }
        if (BMODE == kBScalar)
          psc = __ldg(a.B + ((a.b_src == kFromRow) ? b_row_const : (int64_t)pb * a.b_width));
      }
    }
    const int cnt = min(bw, maxdeg - base);  // warp-uniform
    for (int i = 0; i < cnt; i += U * estride) {
      Vec<VEC> va[U][CH];
      Vec<VEC> vb[U][CH];
      float sb[U][CH];
      float ss[U];
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const int ei = i + k * estride + eoff;  // edge of this lane group inside the batch
        const int srcl = ei & (bw - 1);
        const int ia = __shfl_sync(0xffffffffu, pa, srcl, bw);
        int ib = 0;
        float sck = 0.f;
        if (BMODE == kBFull || BMODE == kBHead) ib = __shfl_sync(0xffffffffu, pb, srcl, bw);
        if (BMODE == kBScalar) sck = __shfl_sync(0xffffffffu, psc, srcl, bw);
        const bool on = (ei < bw) && (base + ei < deg);
        const float* ap;
        if (TAB) {
          ap = a.a_tab[(ia >> kSlotShift) & (kMaxSlots - 1)] + (int64_t)(ia & kRowMask) * a.D + cbase;
        } else {
          ap = a.A + (int64_t)ia * a.D + cbase;
        }
        const float* bp = nullptr;
        if (BMODE == kBFull || BMODE == kBHead)
          bp = a.B + ((a.b_src == kFromRow) ? b_row_const : (int64_t)ib * a.b_width);
        if (BMODE == kBScalar) ss[k] = on ? sck : 0.f;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
          if (on && col_ok(c)) {
            va[k][c] = ld_vec<VEC>(ap + c * lpr * VEC);
            if (BMODE == kBFull) vb[k][c] = ld_vec<VEC>(bp + cbase + c * lpr * VEC);
            if (BMODE == kBHead) sb[k][c] = __ldg(bp + hc[c]);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const int ei = i + k * estride + eoff;
        if ((ei < bw) && (base + ei < deg)) {
#pragma unroll
          for (int c = 0; c < CH; ++c) {
            if (col_ok(c)) {
#pragma unroll
              for (int e = 0; e < VEC; ++e) {
                float val = va[k][c].v[e];
                if (MUL) {
                  if (BMODE == kBFull) val = __fmul_rn(val, vb[k][c].v[e]);
                  if (BMODE == kBHead) val = __fmul_rn(val, sb[k][c]);
                  if (BMODE == kBScalar) val = __fmul_rn(val, ss[k]);
                } else {
                  if (BMODE == kBFull) val = spmm_value(a.op, a.b_is_lhs, val, vb[k][c].v[e]);
                  if (BMODE == kBHead) val = spmm_value(a.op, a.b_is_lhs, val, sb[k][c]);
                  if (BMODE == kBScalar) val = spmm_value(a.op, a.b_is_lhs, val, ss[k]);
                }
                if (BMODE == kBNone && a.op == kOpNegA) val = __fmul_rn(val, -1.f);
                acc[c].v[e] = SUM ? __fadd_rn(acc[c].v[e], val) : red_call(a.red, acc[c].v[e], val);
              }
            }
          }
        }
      }
    }
  }

  if (!FULL && wide) {
    // fold the lane groups' partial results (fixed butterfly order); group 0 writes
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int e = 0; e < VEC; ++e)
        for (int o = lpr; o < 32; o <<= 1) {
          const float other = __shfl_xor_sync(0xffffffffu, acc[c].v[e], o);
          acc[c].v[e] = SUM ? __fadd_rn(acc[c].v[e], other) : red_call(a.red, acc[c].v[e], other);
        }
    if (lane >= lpr) row_ok = false;
  }
  if (row_ok) {
    float* op;
    if (is_seg) {
      op = a.partial + item * a.D;
    } else {
      const int64_t orow = a.out_map ? __ldg(a.out_map + r) : r;
      op = a.out + orow * a.out_ld;
    }
    const float* addp = (a.addend && !is_seg) ? a.addend + (op - a.out) + cbase : nullptr;
    op += cbase;
    const float dn = (float)(deg > 1 ? deg : 1);
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      if (col_ok(c)) {
        if (SUM && addp) {
          const Vec<VEC> old = ld_vec<VEC>(addp + c * lpr * VEC);
#pragma unroll
          for (int e = 0; e < VEC; ++e) acc[c].v[e] = __fadd_rn(old.v[e], acc[c].v[e]);
        }
        if (a.mean && !is_seg) {
#pragma unroll
          for (int e = 0; e < VEC; ++e) acc[c].v[e] = __fdiv_rn(acc[c].v[e], dn);
        }
        if (a.row_scale && !is_seg) {
          const float rs = __ldg(a.row_scale + r);
#pragma unroll
          for (int e = 0; e < VEC; ++e) acc[c].v[e] = __fmul_rn(acc[c].v[e], rs);
        }
        if (is_seg || a.out_cols == a.D) {
          st_vec<VEC>(op + c * lpr * VEC, acc[c]);
        } else {
#pragma unroll
          for (int e = 0; e < VEC; ++e)
            if (cbase + c * lpr * VEC + e < a.out_cols) op[c * lpr * VEC + e] = acc[c].v[e];
        }
      }
    }
  }
}

// ---------------------------------------------------------------- host side
// Neighbours whose feature loads are issued back to back. Measured on B200
// (profiles/microbench): a random-row gather runs fastest with ~64 KB of loads in flight per SM;
// more (deeper unrolling) thrashes L1 and is up to 3x slower. With 32-64 resident warps that is
// ~1-2 KB per warp, i.e. U = 1 once a warp moves >= 1 KB per neighbour. Below that, 4 neighbours
// in flight (products-shaped graph, D=100: U=2 -> 4 took the kernel from 5.72 to 4.89 ms; U=8 at
// 256 B per neighbour was slower again, 4.66 -> 5.54 ms at D=47).
constexpr int unroll_for(int vec, int ch) {
  const int bytes = ch * vec * 128;  // per neighbour per warp
  return bytes >= 1024 ? 1 : (bytes >= 256 ? 4 : 8);
}

// Threads per CTA of the SpMM launches (B200_SPMM_BLOCK = 64 | 128 | 256; the kernels are
// compiled for at most 256). A CTA's registers and warp slots are held until its LAST row
// finishes, so with rows of unequal length small CTAs waste less: measured 256 -> 64 threads on
// the products-shaped SBM graph D=100 4.31 -> 3.76 ms, D=256 7.00 -> 6.68, D=48 (short-row
// kernel) 1.52 -> 1.38, config 4 0.241 -> 0.227, Reddit shape D=602 unchanged (29.0 -> 28.9).
// `blocks256` arrives as the CTA count at 256 threads.
inline int spmm_block_threads() {
  const char* e = std::getenv("B200_SPMM_BLOCK");
  const int t = e ? std::atoi(e) : 64;
  return (t == 128 || t == 256) ? t : 64;
}

template <int VEC, int CH, int U>
static void launch_shape(const SpmmArgs& a, int bmode, bool sum, int blocks256, cudaStream_t s) {
  const int T = spmm_block_threads();
  const int blocks = (int)(((int64_t)blocks256 * 256 + T - 1) / T);
  if (a.a_tab[0] != nullptr) {  // peer-table gather: plain copy + sum only (checked by the caller)
    if (a.lpr == 32) {
      spmm_kernel<VEC, CH, U, true, kBNone, true, false, true><<<blocks, T, 0, s>>>(a);
    } else {
      if constexpr (CH == 1)
        spmm_kernel<VEC, CH, U, true, kBNone, false, false, true><<<blocks, T, 0, s>>>(a);
    }
    return;
  }
#define B200_SPMM_LAUNCH(SUMV, BM, MULV)                                             \
  do {                                                                               \
    if (a.lpr == 32) {                                                               \
      spmm_kernel<VEC, CH, U, SUMV, BM, true, MULV><<<blocks, T, 0, s>>>(a);       \
    } else {                                                                         \
      if constexpr (CH == 1)                                                         \
        spmm_kernel<VEC, CH, U, SUMV, BM, false, MULV><<<blocks, T, 0, s>>>(a);    \
    }                                                                                \
  } while (0)
  // the mul specialisation exists for the sum reducer only (u_mul_e + sum and its transposes)
  const bool mul = sum && bmode != kBNone && a.op == kMul;
  if (sum) {
    if (bmode == kBNone) B200_SPMM_LAUNCH(true, kBNone, false);
    else if (bmode == kBFull) { if (mul) B200_SPMM_LAUNCH(true, kBFull, true); else B200_SPMM_LAUNCH(true, kBFull, false); }
    else if (bmode == kBHead) { if (mul) B200_SPMM_LAUNCH(true, kBHead, true); else B200_SPMM_LAUNCH(true, kBHead, false); }
    else { if (mul) B200_SPMM_LAUNCH(true, kBScalar, true); else B200_SPMM_LAUNCH(true, kBScalar, false); }
  } else {
    if (bmode == kBNone) B200_SPMM_LAUNCH(false, kBNone, false);
    else if (bmode == kBFull) B200_SPMM_LAUNCH(false, kBFull, false);
    else if (bmode == kBHead) B200_SPMM_LAUNCH(false, kBHead, false);
    else B200_SPMM_LAUNCH(false, kBScalar, false);
  }
#undef B200_SPMM_LAUNCH
}

template <int VEC>
void launch_vec(const SpmmArgs& a, int ch, int bmode, bool sum, int blocks,
                       cudaStream_t s) {
  switch (ch) {
    case 1: launch_shape<VEC, 1, unroll_for(VEC, 1)>(a, bmode, sum, blocks, s); break;
    case 2: launch_shape<VEC, 2, unroll_for(VEC, 2)>(a, bmode, sum, blocks, s); break;
    case 3: launch_shape<VEC, 3, unroll_for(VEC, 3)>(a, bmode, sum, blocks, s); break;
    case 4: launch_shape<VEC, 4, unroll_for(VEC, 4)>(a, bmode, sum, blocks, s); break;
    case 5: launch_shape<VEC, 5, unroll_for(VEC, 5)>(a, bmode, sum, blocks, s); break;
    case 6: launch_shape<VEC, 6, unroll_for(VEC, 6)>(a, bmode, sum, blocks, s); break;
    case 8: launch_shape<VEC, 8, unroll_for(VEC, 8)>(a, bmode, sum, blocks, s); break;
    case 10: launch_shape<VEC, 10, unroll_for(VEC, 10)>(a, bmode, sum, blocks, s); break;
    default: launch_shape<VEC, 12, unroll_for(VEC, 12)>(a, bmode, sum, blocks, s); break;
  }
}

}  // namespace b200

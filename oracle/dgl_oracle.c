/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of the reference's builtin
 * message-passing path. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this file's library. The product (dgl-hack_b200/) never does.
 *
 * Parity status: PINNED. Every function here is checked (tests/test_oracle_vs_ref.py, run where
 * /root/reference exists, and tests/test_oracle_golden.py everywhere) against
 *   - the known-answer index fixtures of tests/cpp/test_spmat.cc:11-126,229-281,357-421, and
 *   - outputs of the UNMODIFIED reference CPU library (oracle/_ref/libdgl.so, built by
 *     oracle/ref_build/Makefile) committed as tests/golden/*.npz by tests/golden/make_golden.py.
 *
 * Plain C, single-threaded, float arithmetic in exactly the reference's order:
 *   forward  walks the OUT-CSR rows (graph src) in order and scatter-reduces into out
 *            (src/kernel/cpu/binary_reduce_impl.h:144-173, minigun cpu/advance.h:71-117);
 *   backward walks the IN-CSR rows (graph dst) with src/dst selectors switched
 *            (src/kernel/cpu/backward_binary_reduce_impl.h:208-246).
 * Index type is int64 throughout (the reference instantiates int32 and int64 identically,
 * src/kernel/common.h:62-81).
 */
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef int64_t idx_t;

/* target codes: src/kernel/binary_reduce_common.h:39-44 */
enum { T_SRC = 0, T_DST = 1, T_EDGE = 2, T_NONE = 3 };
/* op codes (names at binary_reduce_common.h:27-33) */
enum { OP_ADD = 0, OP_SUB = 1, OP_MUL = 2, OP_DIV = 3, OP_DOT = 4, OP_USE_LHS = 5 };
/* reducer codes (binary_reduce_common.h:18-25) */
enum { R_SUM = 0, R_MAX = 1, R_MIN = 2, R_PROD = 3, R_NONE = 4 };

/* Broadcast description == BcastInfo (src/kernel/binary_reduce.h:21-35) after CalcBcastInfo
 * (src/kernel/binary_reduce.cc:96-155). The non-broadcast path
 * (cpu/binary_reduce_impl.h:29-52) is the special case ndim=1, all shapes {D}. */
typedef struct {
  int32_t ndim;
  int64_t lhs_shape[8], lhs_stride[8];
  int64_t rhs_shape[8], rhs_stride[8];
  int64_t out_shape[8], out_stride[8];
  int64_t lhs_len, rhs_len, out_len, data_len;
} orc_bcast_t;

/* ---------------------------------------------------------------- index preparation */

/* aten::COOToCSR, src/array/cpu/spmat_op_impl_coo.cc:230-283 (unsorted branch):
 * counting sort by row, stable in COO order; data[] = COO position (edge id). */
void orc_coo_to_csr(idx_t n_rows, idx_t nnz, const idx_t* row, const idx_t* col,
                    idx_t* indptr, idx_t* indices, idx_t* data) {
  idx_t i, cumsum = 0, last = 0;
  for (i = 0; i < n_rows; ++i) indptr[i] = 0;
  for (i = 0; i < nnz; ++i) indptr[row[i]]++;
  for (i = 0; i < n_rows; ++i) { idx_t t = indptr[i]; indptr[i] = cumsum; cumsum += t; }
  indptr[n_rows] = nnz;
  for (i = 0; i < nnz; ++i) {
    idx_t r = row[i];
    indices[indptr[r]] = col[i];
    data[indptr[r]] = i;
    indptr[r]++;
  }
  for (i = 0; i <= n_rows; ++i) { idx_t t = indptr[i]; indptr[i] = last; last = t; }
}

/* aten::CSRTranspose, src/array/cpu/spmat_op_impl.cc:323-369. ax may be NULL (data = position). */
void orc_csr_transpose(idx_t n_rows, idx_t n_cols, const idx_t* ap, const idx_t* aj,
                       const idx_t* ax, idx_t* bp, idx_t* bi, idx_t* bx) {
  idx_t i, j, cumsum = 0, last = 0, nnz = ap[n_rows];
  for (i = 0; i < n_cols; ++i) bp[i] = 0;
  for (j = 0; j < nnz; ++j) bp[aj[j]]++;
  for (i = 0; i < n_cols; ++i) { idx_t t = bp[i]; bp[i] = cumsum; cumsum += t; }
  bp[n_cols] = nnz;
  for (i = 0; i < n_rows; ++i) {
    for (j = ap[i]; j < ap[i + 1]; ++j) {
      idx_t dst = aj[j];
      bi[bp[dst]] = i;
      bx[bp[dst]] = ax ? ax[j] : j;
      bp[dst]++;
    }
  }
  for (i = 0; i <= n_cols; ++i) { idx_t t = bp[i]; bp[i] = last; last = t; }
}

/* aten::CSRToCOO, src/array/cpu/spmat_op_impl.cc:371-385: expand indptr to a row array. */
void orc_csr_to_coo_rows(idx_t n_rows, const idx_t* indptr, idx_t* row) {
  idx_t i, j;
  for (i = 0; i < n_rows; ++i)
    for (j = indptr[i]; j < indptr[i + 1]; ++j) row[j] = i;
}

/* CSR degrees: src/graph/immutable_graph.cc:177 (indptr[v+1]-indptr[v]). */
void orc_degrees(idx_t n_rows, const idx_t* indptr, idx_t* deg) {
  idx_t i;
  for (i = 0; i < n_rows; ++i) deg[i] = indptr[i + 1] - indptr[i];
}

/* aten::AsNumBits 64->32, src/array/array.cc (value-preserving narrowing). */
void orc_as_int32(idx_t n, const idx_t* in, int32_t* out) {
  idx_t i;
  for (i = 0; i < n; ++i) out[i] = (int32_t)in[i];
}

/* ---------------------------------------------------------------- functors */

static float op_call(int op, const float* l, const float* r, int64_t len) {
  /* src/kernel/binary_reduce_common.h:131-213 */
  switch (op) {
    case OP_ADD: return l[0] + r[0];
    case OP_SUB: return l[0] - r[0];
    case OP_MUL: return l[0] * r[0];
    case OP_DIV: return l[0] / r[0];
    case OP_USE_LHS: return l[0];
    default: {
      float o = 0;
      int64_t i;
      for (i = 0; i < len; ++i) o += l[i] * r[i];
      return o;
    }
  }
}
static float op_bwd_lhs(int op, float l, float r, float e) {
  (void)e;
  switch (op) {
    case OP_ADD: case OP_SUB: case OP_USE_LHS: return 1.f;
    case OP_MUL: case OP_DOT: return r;
    default: return 1.f / r; /* div */
  }
}
static float op_bwd_rhs(int op, float l, float r, float e) {
  (void)e;
  switch (op) {
    case OP_ADD: return 1.f;
    case OP_SUB: return -1.f;
    case OP_MUL: case OP_DOT: return l;
    case OP_USE_LHS: return 0.f;
    default: return -l / (r * r); /* div */
  }
}
static float red_zero(int red) {
  /* Zero<Reducer>, binary_reduce_common.h:447-470 */
  switch (red) {
    case R_MAX: return -FLT_MAX;
    case R_MIN: return FLT_MAX;
    case R_PROD: return 1.f;
    default: return 0.f;
  }
}
static void red_call(int red, float* addr, float v) {
  /* src/kernel/cpu/functor.h:18-72 */
  switch (red) {
    case R_SUM: *addr += v; break;
    case R_MAX: *addr = (*addr < v) ? v : *addr; break; /* std::max(*addr, v) */
    case R_MIN: *addr = (v < *addr) ? v : *addr; break; /* std::min(*addr, v) */
    case R_PROD: *addr *= v; break;
    default: *addr = v; break;
  }
}
static float red_bwd(int red, float val, float accum) {
  switch (red) {
    case R_MAX: case R_MIN: return (float)(val == accum);
    case R_PROD: return accum / val;
    default: return 1.f;
  }
}

/* cpu/binary_reduce_impl.h:55-72 */
static void unravel(int64_t idx, int ndim, const int64_t* shape, const int64_t* stride,
                    int64_t* out) {
  int d;
  for (d = 0; d < ndim; ++d) out[d] = (idx / stride[d]) % shape[d];
}
static int64_t ravel(const int64_t* idx, int ndim, const int64_t* shape, const int64_t* stride) {
  int64_t out = 0;
  int d;
  for (d = 0; d < ndim; ++d) {
    int64_t i = idx[d] < shape[d] - 1 ? idx[d] : shape[d] - 1;
    out += i * stride[d];
  }
  return out;
}

static idx_t select_id(int tgt, idx_t src, idx_t pos, idx_t dst) {
  switch (tgt) {
    case T_SRC: return src;
    case T_DST: return dst;
    case T_EDGE: return pos;
    default: return 0;
  }
}

/* ---------------------------------------------------------------- forward */

/* BinaryReduceImpl / BinaryReduceBcastImpl<kDLCPU> (src/kernel/binary_reduce_impl.h:66-114,
 * 266-322) + cpu::BinaryReduce{,Bcast}::ApplyEdge (cpu/binary_reduce_impl.h:22-113).
 * (indptr, indices, eids) is the OUT-CSR: row = graph src, indices = graph dst, eids = csr.data.
 * Mappings may be NULL; an edge-target operand with NULL mapping reads through eids
 * (cpu/binary_reduce_impl.h:157-166). out has out_rows * out_len elements and is first filled
 * with the reducer identity (binary_reduce_impl.h:62). */
void orc_binary_reduce(int red, int op, int lhs_tgt, int rhs_tgt, idx_t n_rows,
                       const idx_t* indptr, const idx_t* indices, const idx_t* eids,
                       const float* lhs, const float* rhs, float* out, int64_t out_rows,
                       const idx_t* lhs_map, const idx_t* rhs_map, const idx_t* out_map,
                       const orc_bcast_t* bc) {
  const int64_t len = bc->data_len;
  const int out_tgt = (red == R_NONE) ? T_EDGE : T_DST;
  const float zero = red_zero(red);
  int64_t i, tx, tmp[8];
  idx_t src, pos;
  for (i = 0; i < out_rows * bc->out_len; ++i) out[i] = zero;
  if (lhs_tgt == T_EDGE && !lhs_map) lhs_map = eids;
  if (rhs_tgt == T_EDGE && !rhs_map) rhs_map = eids;
  if (out_tgt == T_EDGE && !out_map) out_map = eids;
  for (src = 0; src < n_rows; ++src) {
    for (pos = indptr[src]; pos < indptr[src + 1]; ++pos) {
      const idx_t dst = indices[pos];
      idx_t lid = select_id(lhs_tgt, src, pos, dst);
      idx_t rid = select_id(rhs_tgt, src, pos, dst);
      idx_t oid = select_id(out_tgt, src, pos, dst);
      const float *lhsoff, *rhsoff;
      float* outoff;
      if (lhs_map) lid = lhs_map[lid];
      if (rhs_map) rid = rhs_map[rid];
      if (out_map) oid = out_map[oid];
      lhsoff = lhs + lid * bc->lhs_len * len;
      rhsoff = rhs ? rhs + rid * bc->rhs_len * len : lhs;
      outoff = out + oid * bc->out_len;
      for (tx = 0; tx < bc->out_len; ++tx) {
        float v;
        unravel(tx, bc->ndim, bc->out_shape, bc->out_stride, tmp);
        v = op_call(op, lhsoff + ravel(tmp, bc->ndim, bc->lhs_shape, bc->lhs_stride) * len,
                    rhsoff + ravel(tmp, bc->ndim, bc->rhs_shape, bc->rhs_stride) * len, len);
        red_call(red, outoff + tx, v);
      }
    }
  }
}

/* ---------------------------------------------------------------- backward */

/* BackwardBinaryReduce{,Bcast}Impl<kDLCPU> (binary_reduce_impl.h:166-222, 378-436) +
 * cpu::BackwardBinaryReduce{,Bcast}::ApplyEdge (cpu/backward_binary_reduce_impl.h:20-163).
 * (indptr, indices, eids) is the IN-CSR: row = graph dst, indices = graph src, eids = csr.data.
 * mode 0 = grad wrt lhs, 1 = grad wrt rhs (BACKWARD_MODE_SWITCH never picks kGradBoth).
 * grad has grad_rows * out_len * data_len elements (the UNREDUCED out feature shape) and is
 * zero-filled first. */
void orc_backward_binary_reduce(int mode, int red, int op, int lhs_tgt, int rhs_tgt,
                                idx_t n_rows, const idx_t* indptr, const idx_t* indices,
                                const idx_t* eids, const float* lhs, const float* rhs,
                                const float* out, const float* grad_out, float* grad,
                                int64_t grad_rows, const idx_t* lhs_map, const idx_t* rhs_map,
                                const idx_t* out_map, const orc_bcast_t* bc) {
  const int64_t len = bc->data_len;
  const int out_tgt = (red == R_NONE) ? T_EDGE : T_DST;
  int64_t i, tx, k, tmp[8];
  idx_t dst, pos;
  for (i = 0; i < grad_rows * bc->out_len * len; ++i) grad[i] = 0.f;
  if (lhs_tgt == T_EDGE && !lhs_map) lhs_map = eids;
  if (rhs_tgt == T_EDGE && !rhs_map) rhs_map = eids;
  if (out_tgt == T_EDGE && !out_map) out_map = eids;
  for (dst = 0; dst < n_rows; ++dst) {
    for (pos = indptr[dst]; pos < indptr[dst + 1]; ++pos) {
      const idx_t src = indices[pos];
      idx_t lid = select_id(lhs_tgt, src, pos, dst);
      idx_t rid = select_id(rhs_tgt, src, pos, dst);
      idx_t oid = select_id(out_tgt, src, pos, dst);
      const float *lhsoff, *rhsoff, *outoff, *goutoff;
      float* gradoff;
      if (lhs_map) lid = lhs_map[lid];
      if (rhs_map) rid = rhs_map[rid];
      if (out_map) oid = out_map[oid];
      lhsoff = lhs + lid * bc->lhs_len * len;
      rhsoff = rhs ? rhs + rid * bc->rhs_len * len : lhs;
      outoff = out + oid * bc->out_len;
      goutoff = grad_out + oid * bc->out_len;
      gradoff = grad + (mode == 0 ? lid : rid) * bc->out_len * len;
      for (tx = 0; tx < bc->out_len; ++tx) {
        const float *lb, *rb;
        float e, ge;
        unravel(tx, bc->ndim, bc->out_shape, bc->out_stride, tmp);
        lb = lhsoff + ravel(tmp, bc->ndim, bc->lhs_shape, bc->lhs_stride) * len;
        rb = rhsoff + ravel(tmp, bc->ndim, bc->rhs_shape, bc->rhs_stride) * len;
        e = op_call(op, lb, rb, len);
        ge = goutoff[tx] * red_bwd(red, e, outoff[tx]);
        for (k = 0; k < len; ++k) {
          const float g = (mode == 0) ? ge * op_bwd_lhs(op, lb[k], rb[k], e)
                                      : ge * op_bwd_rhs(op, lb[k], rb[k], e);
          gradoff[tx * len + k] += g;
        }
      }
    }
  }
}

/* ---------------------------------------------------------------- fused GAT (hack ops)
 * CPU restatement of the fork's CUDA-only kernels. PARITY: pinned only through the builtin
 * composition the reference itself checks against (examples/pytorch/gat/fused_gat_unit_test.py:
 * 36-52: u_add_v -> leaky_relu -> exp -> copy_e+sum -> e_div_v -> u_mul_e+sum, no
 * max-subtraction); the reference kernels cannot run without a GPU build. */
#include <math.h>

static float gat_leaky_exp(float v, float slope) {
  /* gatLeakyReluExp, src/kernel/cuda/binary_reduce_impl.cu:46-49 */
  return v > 0 ? expf(v) : expf(slope * v);
}

/* gatExpLeakyReluSumKernel + gatSumProdZipDivKernel (binary_reduce_impl.cu:51-113) over the
 * in-CSR (row = dst). exp is indexed by edge id. */
void orc_fused_gat_forward(idx_t n_dst, const idx_t* indptr, const idx_t* indices,
                           const idx_t* eids, int64_t H, int64_t F, const float* ft,
                           const float* el, const float* er, float slope, float* sum,
                           float* exp_out, float* ret) {
  idx_t v, p;
  int64_t h, f;
  for (v = 0; v < n_dst; ++v) {
    for (h = 0; h < H; ++h) {
      float s = 0.f;
      for (p = indptr[v]; p < indptr[v + 1]; ++p) {
        float t = gat_leaky_exp(el[indices[p] * H + h] + er[v * H + h], slope);
        exp_out[eids[p] * H + h] = t;
        s += t;
      }
      sum[v * H + h] = s;
    }
    for (h = 0; h < H; ++h)
      for (f = 0; f < F; ++f) {
        float s = 0.f;
        for (p = indptr[v]; p < indptr[v + 1]; ++p)
          s += exp_out[eids[p] * H + h] / sum[v * H + h] * ft[(indices[p] * H + h) * F + f];
        ret[(v * H + h) * F + f] = s;
      }
  }
}

/* fusedGatBackwardGradFeatSrc + fusedGatBackwardGradElEr (binary_reduce_impl.cu:123-213) over the
 * out-CSR (row = src). grad_el / grad_er are zero-filled here (the torch side passes zeros). */
void orc_fused_gat_backward(idx_t n_src, idx_t n_dst, const idx_t* indptr, const idx_t* indices,
                            const idx_t* eids, int64_t H, int64_t F, const float* ft,
                            const float* el, const float* er, const float* sum,
                            const float* exp_in, const float* ret, const float* gout, float slope,
                            float* grad_ft, float* grad_el, float* grad_er) {
  idx_t u, p;
  int64_t h, f, i;
  for (i = 0; i < n_src * H; ++i) grad_el[i] = 0.f;
  for (i = 0; i < n_dst * H; ++i) grad_er[i] = 0.f;
  for (u = 0; u < n_src; ++u)
    for (h = 0; h < H; ++h)
      for (f = 0; f < F; ++f) {
        float s = 0.f, s2 = 0.f;
        for (p = indptr[u]; p < indptr[u + 1]; ++p) {
          idx_t d = indices[p], e = eids[p];
          float grad_exp, tmp_sum, tmp2;
          s += exp_in[e * H + h] / sum[d * H + h] * gout[(d * H + h) * F + f];
          grad_exp = gout[(d * H + h) * F + f] *
                     (ft[(u * H + h) * F + f] - ret[(d * H + h) * F + f]) / sum[d * H + h];
          tmp_sum = el[u * H + h] + er[d * H + h];
          tmp2 = grad_exp * exp_in[e * H + h] * (tmp_sum > 0 ? 1.f : slope);
          s2 += tmp2;
          grad_er[d * H + h] += tmp2;
        }
        grad_ft[(u * H + h) * F + f] = s;
        grad_el[u * H + h] += s2;
      }
}

/* ------------------------------------------------------------------ RGCN hack ops (a15)
 * PARITY UNPINNED: these kernels are CUDA-only in the reference (and the fork's library does not
 * link them into a CPU build), nothing in its tests touches them; what follows restates the CUDA
 * source thread by thread in plain C. */
#include <stdlib.h>

typedef struct { idx_t type, pos; } orc_tkey;

static int orc_tkey_cmp(const void* a, const void* b) {
  const orc_tkey* x = (const orc_tkey*)a;
  const orc_tkey* y = (const orc_tkey*)b;
  if (x->type != y->type) return x->type < y->type ? -1 : 1;
  return x->pos < y->pos ? -1 : (x->pos > y->pos ? 1 : 0);
}

/* Graph::GetCsrSortedByEdgeType (src/graph/graph.cc:690-741): adjacency lists in insertion
 * (edge id) order (graph.cc:97-108), each row sorted by type; std::sort's order within one type
 * is unspecified, the stable order (type, edge id) is used. key/other/type are indexed by edge id;
 * key = dst for the in-adjacency (transpose = false), src for the out-adjacency. */
void orc_typed_csr(idx_t n_rows, idx_t n_edges, const idx_t* key, const idx_t* other,
                   const idx_t* type, idx_t* indptr, idx_t* ids, idx_t* eids, idx_t* types) {
  idx_t e, r, p;
  idx_t* fill;
  orc_tkey* keys;
  for (r = 0; r <= n_rows; ++r) indptr[r] = 0;
  for (e = 0; e < n_edges; ++e) indptr[key[e] + 1]++;
  for (r = 0; r < n_rows; ++r) indptr[r + 1] += indptr[r];
  fill = (idx_t*)malloc(sizeof(idx_t) * (size_t)(n_rows + 1));
  for (r = 0; r < n_rows; ++r) fill[r] = indptr[r];
  for (e = 0; e < n_edges; ++e) eids[fill[key[e]]++] = e;
  free(fill);
  keys = (orc_tkey*)malloc(sizeof(orc_tkey) * (size_t)(n_edges + 1));
  for (r = 0; r < n_rows; ++r) {
    idx_t beg = indptr[r], n = indptr[r + 1] - indptr[r];
    for (p = 0; p < n; ++p) {
      keys[p].type = type[eids[beg + p]];
      keys[p].pos = eids[beg + p];
    }
    qsort(keys, (size_t)n, sizeof(orc_tkey), orc_tkey_cmp);
    for (p = 0; p < n; ++p) eids[beg + p] = keys[p].pos;
  }
  free(keys);
  for (p = 0; p < n_edges; ++p) {
    ids[p] = other[eids[p]];
    types[p] = type[eids[p]];
  }
}

/* RgcnLayer0KernelImpl (binary_reduce_impl.cu:913-933): block = dst row, thread = column. */
void orc_rgcn_layer0(idx_t n_nodes, const idx_t* ranges, const idx_t* src_ids, const idx_t* eids,
                     const idx_t* types, int64_t n_w_rows, int64_t feat, const float* weight,
                     const float* norm, float* ret) {
  idx_t v, p;
  int64_t tx;
  for (v = 0; v < n_nodes; ++v)
    for (tx = 0; tx < feat; ++tx) {
      float agg = 0.f;
      for (p = ranges[v]; p < ranges[v + 1]; ++p) {
        float w = weight[types[p] * n_w_rows * feat + src_ids[p] * feat + tx];
        agg += w * norm[eids[p]];
      }
      ret[v * feat + tx] = agg;
    }
}

/* RgcnLayer0BackwardKernelImpl (binary_reduce_impl.cu:982-1008): block = src row of the typed
 * out-adjacency. accumulate == 0 is the source as written (plain store, the last edge of a
 * (type, src) pair wins); accumulate != 0 sums, which is the gradient of orc_rgcn_layer0.
 * grad_weight must be zero-filled by the caller (tensor.py:454). */
void orc_rgcn_layer0_backward(idx_t n_nodes, const idx_t* ranges, const idx_t* dst_ids,
                              const idx_t* eids, const idx_t* types, int64_t n_w_rows,
                              int64_t feat, const float* grad_out, const float* norm,
                              int accumulate, float* grad_weight) {
  idx_t u, p;
  int64_t tx;
  for (u = 0; u < n_nodes; ++u)
    for (tx = 0; tx < feat; ++tx)
      for (p = ranges[u]; p < ranges[u + 1]; ++p) {
        float v = grad_out[dst_ids[p] * feat + tx] * norm[eids[p]];
        float* slot = grad_weight + types[p] * n_w_rows * feat + u * feat + tx;
        if (accumulate) *slot += v; else *slot = v;
      }
}

/* RgcnLayer1KernelImpl (binary_reduce_impl.cu:1082-1117): thread = (y, x); per-thread sum over
 * the row, then atomicAdd over y (taken in ascending y here). ret is zero-filled first. */
void orc_rgcn_layer1(idx_t n_nodes, const idx_t* ranges, const idx_t* src_ids, const idx_t* eids,
                     const idx_t* types, int64_t fy, int64_t fx, const float* hidden,
                     const float* weight, const float* norm, float* ret) {
  idx_t v, p;
  int64_t y, x;
  for (v = 0; v < n_nodes; ++v) {
    for (x = 0; x < fx; ++x) ret[v * fx + x] = 0.f;
    for (y = 0; y < fy; ++y)
      for (x = 0; x < fx; ++x) {
        float agg = 0.f;
        for (p = ranges[v]; p < ranges[v + 1]; ++p) {
          float w = weight[types[p] * fy * fx + y * fx + x];
          float h = hidden[src_ids[p] * fy + y];
          agg += h * w * norm[eids[p]];
        }
        ret[v * fx + x] += agg;
      }
  }
}

/* RgcnLayer1BackwardKernelImpl (binary_reduce_impl.cu:1157-1194) over the typed out-adjacency;
 * atomics taken in (row, y, x, position) order. Outputs are zero-filled first (tensor.py:483-484). */
void orc_rgcn_layer1_backward(idx_t n_nodes, idx_t n_types, const idx_t* ranges,
                              const idx_t* dst_ids, const idx_t* eids, const idx_t* types,
                              int64_t fy, int64_t fx, const float* hidden, const float* weight,
                              const float* norm, const float* grad_out, float* grad_hidden,
                              float* grad_weight) {
  idx_t u, p;
  int64_t y, x, i;
  for (i = 0; i < n_nodes * fy; ++i) grad_hidden[i] = 0.f;
  for (i = 0; i < n_types * fy * fx; ++i) grad_weight[i] = 0.f;
  for (u = 0; u < n_nodes; ++u)
    for (y = 0; y < fy; ++y)
      for (x = 0; x < fx; ++x) {
        float h = hidden[u * fy + y];
        float agg = 0.f;
        for (p = ranges[u]; p < ranges[u + 1]; ++p) {
          float g = grad_out[dst_ids[p] * fx + x];
          float w = weight[types[p] * fy * fx + y * fx + x];
          float n = norm[eids[p]];
          agg += g * w * n;
          grad_weight[types[p] * fy * fx + y * fx + x] += g * h * n;
        }
        grad_hidden[u * fy + y] += agg;
      }
}

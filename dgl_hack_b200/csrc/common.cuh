// Shared definitions for the sm_100a message-passing kernels.
//
// Arithmetic contract (what makes results reproducible against the reference's single-threaded
// CPU path): every float op is an explicitly rounded intrinsic (__fadd_rn / __fmul_rn / __fdiv_rn)
// so nvcc never contracts mul+add into an FMA, and every reduction visits a row's edges
// sequentially in CSR order -- the same order in which minigun's CPU advance
// (third_party/minigun/minigun/cpu/advance.h:71-117) reaches a given output element when run with
// one thread. Functor semantics follow src/kernel/binary_reduce_common.h:131-213, 406-498 and
// src/kernel/cpu/functor.h:18-72.
#pragma once
#include <cuda_runtime.h>
#include "../../include/dgl_b200.h"
#include <float.h>
#include <stdint.h>

#include <string>

namespace b200 {

enum Target { kSrc = 0, kDst = 1, kEdge = 2, kNone = 3 };
enum OpCode { kAdd = 0, kSub = 1, kMul = 2, kDiv = 3, kDot = 4, kUseLhs = 5 };
enum RedCode { kSum = 0, kMax = 1, kMin = 2, kProd = 3, kRNone = 4 };
enum Mode { kFwd = 0, kGradLhs = 1, kGradRhs = 2 };

constexpr int kMaxBcastDim = 8;  // BCAST_NDIM_SWITCH tops out at 8 (binary_reduce_impl.h:238-255)

// BcastInfo (src/kernel/binary_reduce.h:21-35) with the row dimension removed. The
// non-broadcast path is ndim == 1 with all three shapes equal to {x_len}.
struct Bcast {
  int ndim;
  int lhs_shape[kMaxBcastDim], lhs_stride[kMaxBcastDim];
  int rhs_shape[kMaxBcastDim], rhs_stride[kMaxBcastDim];
  int out_shape[kMaxBcastDim], out_stride[kMaxBcastDim];
  int64_t lhs_len, rhs_len, out_len, data_len;
};

// ---------------------------------------------------------------- functors
__device__ __forceinline__ float red_identity(int red) {
  switch (red) {
    case kMax: return -FLT_MAX;  // std::numeric_limits<float>::lowest()
    case kMin: return FLT_MAX;
    case kProd: return 1.f;
    default: return 0.f;
  }
}
inline float red_identity_host(int red) {
  switch (red) {
    case kMax: return -FLT_MAX;
    case kMin: return FLT_MAX;
    case kProd: return 1.f;
    default: return 0.f;
  }
}

__device__ __forceinline__ float red_call(int red, float acc, float v) {
  switch (red) {
    case kSum: return __fadd_rn(acc, v);
    case kMax: return (acc < v) ? v : acc;  // std::max(acc, v)
    case kMin: return (v < acc) ? v : acc;  // std::min(acc, v)
    case kProd: return __fmul_rn(acc, v);
    default: return v;
  }
}

__device__ __forceinline__ float red_bwd(int red, float val, float accum) {
  switch (red) {
    case kMax:
    case kMin: return (val == accum) ? 1.f : 0.f;
    case kProd: return __fdiv_rn(accum, val);
    default: return 1.f;
  }
}

// scalar binary op on single elements (everything except dot)
__device__ __forceinline__ float op_scalar(int op, float l, float r) {
  switch (op) {
    case kAdd: return __fadd_rn(l, r);
    case kSub: return __fsub_rn(l, r);
    case kMul: return __fmul_rn(l, r);
    case kDiv: return __fdiv_rn(l, r);
    default: return l;  // use_lhs
  }
}

__device__ __forceinline__ float op_call(int op, const float* __restrict__ l,
                                         const float* __restrict__ r, int len) {
  if (op == kDot) {
    float o = 0.f;
    for (int i = 0; i < len; ++i) o = __fadd_rn(o, __fmul_rn(__ldg(l + i), __ldg(r + i)));
    return o;
  }
  if (op == kUseLhs) return __ldg(l);
  return op_scalar(op, __ldg(l), __ldg(r));
}

__device__ __forceinline__ float op_bwd_lhs(int op, float l, float r) {
  switch (op) {
    case kMul:
    case kDot: return r;
    case kDiv: return __fdiv_rn(1.f, r);
    default: return 1.f;  // add, sub, use_lhs
  }
}

__device__ __forceinline__ float op_bwd_rhs(int op, float l, float r) {
  switch (op) {
    case kAdd: return 1.f;
    case kSub: return -1.f;
    case kMul:
    case kDot: return l;
    case kUseLhs: return 0.f;
    default: return __fdiv_rn(-l, __fmul_rn(r, r));  // div
  }
}

// ---------------------------------------------------------------- host helpers
void set_last_error(const std::string& msg);
void note_kernel(const char* family, int launches);

struct ApiError {
  std::string msg;
};

#define B200_CHECK(cond, message)                       \
  do {                                                  \
    if (!(cond)) throw ::b200::ApiError{std::string(message)}; \
  } while (0)

#define B200_CUDA(expr)                                                                   \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess)                                                                \
      throw ::b200::ApiError{std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + \
                             __FILE__ + ":" + std::to_string(__LINE__)};                   \
  } while (0)

inline int ceil_div(int64_t a, int64_t b) { return static_cast<int>((a + b - 1) / b); }

// Stream-ordered scratch allocation (the reference's WorkspacePool role,
// src/runtime/workspace_pool.cc): freed on the same stream, never synchronises.
struct Scratch {
  void* ptr = nullptr;
  cudaStream_t stream = nullptr;
  Scratch() {}
  Scratch(size_t bytes, cudaStream_t s) { alloc(bytes, s); }
  void alloc(size_t bytes, cudaStream_t s) {
    stream = s;
    if (bytes == 0) bytes = 4;
    B200_CUDA(cudaMallocAsync(&ptr, bytes, s));
  }
  ~Scratch() {
    if (ptr) cudaFreeAsync(ptr, stream);
  }
  Scratch(const Scratch&) = delete;
  Scratch& operator=(const Scratch&) = delete;
  template <typename T>
  T* as() { return static_cast<T*>(ptr); }
};

// ---------------------------------------------------------------- device graph index
struct Csr {
  int32_t* indptr = nullptr;   // [n_rows + 1]
  int32_t* indices = nullptr;  // [nnz] column (= the other endpoint)
  int32_t* eids = nullptr;     // [nnz] edge id (csr.data)
  int32_t* row = nullptr;      // [nnz] row of each position (lazily built COO expansion)
  int32_t* to_other = nullptr; // [nnz] position of the same edge in the other CSR (lazy)
  int64_t n_rows = 0, n_cols = 0;
  // Segment plans (lazy, one slot per distinct segment length in use): rows with more than
  // seg_len edges are cut into segments of seg_len consecutive positions so that no single warp
  // walks a hub row alone.
  struct Plan {
    int32_t seg_len = 0;          // 0 = slot empty
    int32_t n_seg = 0;            // total segments of heavy rows
    int32_t n_heavy = 0;          // heavy rows
    int32_t n_long = 0;           // non-heavy rows with >= kLongRow edges (a prefix of sched)
    int32_t* seg_row = nullptr;   // [n_seg] row of the segment
    int32_t* seg_beg = nullptr;   // [n_seg] first CSR position
    int32_t* seg_end = nullptr;   // [n_seg] one past the last CSR position
    int32_t* heavy_row = nullptr; // [n_heavy]
    int32_t* heavy_ptr = nullptr; // [n_heavy + 1] segment range of each heavy row
    int32_t* sched = nullptr;     // [n_rows] rows by descending degree class (NULL = natural order)
    // the same with every row below kLongRow edges in ONE class (natural order); built on the
    // first request of a one-row-per-warp kernel, and only for graphs whose numbering is local
    int32_t* sched_nat = nullptr;
    int8_t nat_state = 0;         // 0 = not probed, 1 = local (sched_nat valid), -1 = not local
  };
  static constexpr int kPlanSlots = 4;
  Plan plan[kPlanSlots];
  int plan_next = 0;  // slot to evict when all are taken
};

constexpr int kLongRow = 64;

struct SegPlan {
  int seg_len, n_seg, n_heavy;
  int n_long;  // with sched: its first n_long rows have >= kLongRow edges, its last n_heavy are heavy
  const int32_t *seg_row, *seg_beg, *seg_end, *heavy_row, *heavy_ptr;
  const int32_t* sched;  // row order for the non-segment work items (may be NULL)
};

// Typed adjacency (Graph::GetCsrSortedByEdgeType, src/graph/graph.cc:690-741): every row of the
// in-/out-CSR re-ordered by (edge type, edge id); indptr is shared with the untyped CSR.
struct TypedCsr {
  int32_t* ids = nullptr;    // [E] other endpoint
  int32_t* eids = nullptr;   // [E]
  int32_t* types = nullptr;  // [E]
};
struct TypeIndex {
  int64_t num_types = 0;      // 0 = no edge types attached
  TypedCsr in, out;
  // type-major edge list (type, edge id): the contraction order of the weight gradients
  int32_t *bt_src = nullptr, *bt_dst = nullptr, *bt_eid = nullptr;
  int32_t* type_ptr = nullptr;  // [num_types + 1]
};

}  // namespace b200

struct DGLB200Graph {
  int64_t num_src = 0, num_dst = 0, num_edges = 0;
  int device = 0;
  b200::Csr in;   // rows = dst, cols = src
  b200::Csr out;  // rows = src, cols = dst
  // The in-CSR with every row re-ordered by (src, edge id) = Transpose(out-CSR). This is the
  // order in which the reference's forward pass (out-CSR walk, one thread) reaches the
  // contributions of a destination row, so reducing over it sequentially reproduces the
  // reference's floating-point sums bit for bit; ascending sources also gather with better
  // locality. indptr / row / heavy_first alias `in` (same degrees); to_other -> out-CSR position.
  b200::Csr in_s;
  b200::TypeIndex ty;
  // endpoints by edge id (lazy; the edge-output kernels walk edges in id order)
  int32_t *coo_src = nullptr, *coo_dst = nullptr;
  // Id validation of graphs built from DEVICE arrays: out-of-range ids are clamped to 0 while the
  // index is built (no out-of-bounds write can happen) and recorded in this device flag, which
  // is read at the first host synchronisation the handle goes through anyway (first kernel plan,
  // CopyCSR, Degrees). Host arrays are checked before the constructor returns.
  int32_t* bad_ids = nullptr;
  bool ids_checked = true;
  // relation-segment graphs only (DGLB200GraphCreateRelationSegments): destination node and
  // relation of every segment (= destination row of this graph)
  int32_t *seg_dst = nullptr, *seg_type = nullptr;
};

namespace b200 {

// graph.cu
enum Walk { kWalkIn = 0, kWalkOut = 1, kWalkInSorted = 2 };
// The plain in-CSR's indices / eids (edge-id order inside a row) are NOT resident for graphs
// built from COO: the hot paths walk in_s and out only, so the handle keeps 4 x E ids (the
// reference's two CSRs) and rebuilds these two arrays from the out-CSR the first time something
// asks for them (GetCSR / CopyCSR, dst-operand gradients, id-mapped calls, the fused GAT
// kernels, the sampler). in.indptr is always there.
const Csr& in_plain(DGLB200Graph* g, cudaStream_t s);
inline const Csr& walk_csr(DGLB200Graph* g, int walk, cudaStream_t s) {
  return walk == kWalkOut ? g->out : (walk == kWalkInSorted ? g->in_s : in_plain(g, s));
}
const int32_t* csr_rows(DGLB200Graph* g, int walk, cudaStream_t s);      // row-of-position
// position in the "other" CSR: in -> out, out -> in, in_sorted -> out
const int32_t* csr_to_other(DGLB200Graph* g, int walk, cudaStream_t s);
// Segment plan of a CSR for segments of `seg_len` positions (built once; synchronises then).
// want_natural: the caller's kernel gives every row its own warp (no lockstep between the rows of
// a warp), so short rows may keep their natural order if the node numbering is local.
SegPlan csr_segments(DGLB200Graph* g, int walk, int seg_len, cudaStream_t s,
                     bool want_natural = false);
int segment_length(int64_t num_edges);  // spmm.cu: the segment length every plan is built with
// Reads the deferred id-validation flag (call only where the stream is synchronised anyway).
void ensure_valid_ids(DGLB200Graph* g, cudaStream_t s);
void exclusive_scan_i32(int32_t* data, int64_t n, cudaStream_t s);
void fill_f32(float* p, int64_t n, float v, cudaStream_t s);

// generic.cu / spmm.cu / sddmm.cu
struct KernelCall {
  DGLB200Graph* g;
  int mode;  // Mode
  int red, op;
  int lhs_tgt, rhs_tgt;
  const float *lhs, *rhs, *out, *gout;  // out/gout only in backward
  float* dst;                           // forward: out tensor; backward: grad tensor
  int64_t lhs_rows, rhs_rows, dst_rows, out_rows;
  const int32_t *lhs_map, *rhs_map, *out_map;  // user mappings (NULL = none)
  Bcast bc;
  bool mean;  // extension: divide the row sum by max(deg, 1)
  const float* row_scale = nullptr;  // extension (SpMM only): multiply finished row r by row_scale[r]
  // extension (SpMM only, plain copy + sum): the gathered operand lives in n_slots blocks
  // (own HBM and peer GPUs' mapped over NVLink); the operand's id map yields (slot, row) codes
  const float* const* a_tab = nullptr;
  int n_slots = 0;
  // SpMM only, sum reducer: dst = row_scale * (addend + sum); addend may alias dst
  const float* addend = nullptr;
  // SpMM only: dst rows have a pitch of dst_ld floats and dst_cols valid columns (0 = the
  // feature width); set by the odd-width route, which gathers from a padded copy
  int64_t dst_ld = 0, dst_cols = 0;
  cudaStream_t stream;
};

void run_generic(const KernelCall& c);
// graph.cu: source (which = 0) / destination (1) of every edge id, built on first use
const int32_t* graph_coo(DGLB200Graph* g, int which, cudaStream_t s);
// gat.cu
void fused_gat_forward(DGLB200Graph* g, const float* ft, const float* el, const float* er,
                       float* sum, float* exp, float* ret, int64_t H, int64_t F, float slope,
                       cudaStream_t s);
void fused_gat_backward(DGLB200Graph* g, const float* ft, const float* el, const float* er,
                        const float* sum, const float* exp, const float* ret, const float* gout,
                        float* grad_ft, float* grad_el, float* grad_er, int64_t H, int64_t F,
                        float slope, cudaStream_t s);
// kway.cu (host): multilevel k-way partitioner; returns the edge cut
int64_t symmetrize_coo(int64_t nvtxs, int64_t num_edges, const int64_t* src, const int64_t* dst,
                       int64_t* xadj, int64_t* adjncy);
int64_t kway_partition(int64_t nvtxs, const int64_t* xadj, const int64_t* adjncy,
                       const int64_t* vwgt, int k, uint64_t seed, double ubfactor, int64_t* part);
// attn.cu
void attn_proj_forward(const float* feat, const float* al, const float* ar, float* el, float* er,
                       int64_t n, int64_t H, int64_t F, cudaStream_t s);
void attn_proj_backward(const float* feat, const float* al, const float* ar, const float* gl,
                        const float* gr, float* gfeat, float* gal, float* gar, int64_t n,
                        int64_t H, int64_t F, cudaStream_t s);
// attn.cu, no-allocation variant: block partials in a file-scope device array (3 MiB per GPU)
void attn_proj_backward_static(const float* feat, const float* al, const float* ar,
                               const float* gl, const float* gr, float* gfeat, float* gal,
                               float* gar, int64_t n, int64_t H, int64_t F, int device,
                               cudaStream_t s);
int64_t attn_proj_static_max_width();
// gemm.cu: fp32 dense product on the tensor cores (tcgen05, 3xTF32 error compensation)
bool gemm_tf32x3_supported(int64_t M, int64_t N, int64_t K, const float* A, int64_t lda);
void gemm_tf32x3(const float* A, int64_t lda, const float* B, int b_is_kn, const float* bias,
                 int relu, float* C, int64_t ldc, int64_t M, int64_t N, int64_t K, int groups,
                 const int64_t* group_ptr, cudaStream_t s);
bool gemm_tn_tf32x3_supported(int64_t M, int64_t Kf, int64_t N, const float* A, const float* G);
void gemm_tn_tf32x3(const float* A, const float* G, float* C, int64_t M, int64_t Kf, int64_t N,
                    cudaStream_t s);
// spmm_bf16.cu: copy + sum / mean over bf16 feature rows (pitch D elements, D % 8 == 0)
void spmm_bf16(DGLB200Graph* g, int transpose, const void* in, const float* col_scale,
               const float* row_scale, int mean, void* out, int64_t D, cudaStream_t s);
// sample.cu
void sample_neighbors(DGLB200Graph* g, const int64_t* seeds, int64_t n_seeds, int fanout,
                      bool replace, uint64_t rng_seed, const int64_t* offsets, int64_t* out_src,
                      int64_t* out_dst, int64_t* out_eid, int64_t* out_slot, cudaStream_t s);
void sample_counts(DGLB200Graph* g, const int64_t* seeds, int64_t n_seeds, int fanout, bool replace,
                   int64_t* offsets, cudaStream_t s);
// block.cu
void compact_block(const int64_t* dst_nodes, int64_t n_dst, const int64_t* src, const int64_t* dst,
                   int64_t n_edges, int64_t* src_nid, int64_t* src_local, int64_t* dst_local,
                   int64_t* counts, cudaStream_t s);
// softmax.cu
void edge_softmax_forward(DGLB200Graph* g, const float* z, float* out, int64_t H, cudaStream_t s);
void edge_softmax_backward(DGLB200Graph* g, const float* out, const float* gout, float* gz,
                           int64_t H, cudaStream_t s);
// rgcn.cu
void rgcn_layer0_forward(DGLB200Graph* g, const float* w, const float* norm, float* ret,
                         int64_t n_rows_w, int64_t feat, cudaStream_t s);
void rgcn_layer0_backward(DGLB200Graph* g, const float* gout, const float* norm, float* grad_w,
                          int64_t n_rows_w, int64_t feat, bool accumulate, cudaStream_t s);
void rgcn_layer1_forward(DGLB200Graph* g, const float* h, const float* w, const float* norm,
                         float* ret, int64_t fy, int64_t fx, cudaStream_t s);
void rgcn_layer1_backward(DGLB200Graph* g, const float* h, const float* w, const float* norm,
                          const float* gout, float* grad_h, float* grad_w, int64_t fy, int64_t fx,
                          cudaStream_t s);
bool try_spmm(const KernelCall& c);  // returns false if the call is not an SpMM-shaped case
bool try_sddmm(const KernelCall& c);

}  // namespace b200

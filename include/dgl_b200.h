/*
 * dgl_b200.h -- C ABI of the B200-native (sm_100a) replacement for DGL 0.4's builtin
 * message-passing kernels (reference: ydwu4/dgl-hack, src/kernel BinaryOpReduce / CopyReduce and
 * their backward kernels, plus the CSR/COO index preparation they consume).
 *
 * Every entry point is `extern "C"`, takes plain pointers and sizes, returns 0 on success and
 * -1 on failure with the message available from DGLB200GetLastError() (thread-local), i.e. the
 * reference's API_BEGIN/API_END + DGLGetLastError convention (src/runtime/runtime_base.h:13-31,
 * src/runtime/c_runtime_api.cc:138-146). Nothing here calls exit() or throws across the ABI.
 *
 * Tensors cross as `DGLB200Tensor`, which is layout-identical to the `DLTensor` the reference's
 * PackedFuncs receive (third_party/dlpack/include/dlpack/dlpack.h, DLPACK_VERSION 020): a
 * DLTensor* obtained from the reference's NDArray / torch DLPack capsule can be passed as is.
 * Optional id-mapping tensors follow the reference's convention: NULL or a 1-D array with
 * shape[0]==0 means "not given" (aten::IsNullArray, include/dgl/array.h:102-104).
 *
 * All kernels are enqueued on `stream` (a cudaStream_t passed as void*; NULL = legacy default
 * stream, the reference's CUDAThreadEntry default, src/runtime/cuda/cuda_common.h:54). No entry
 * point synchronises the host unless its comment says so.
 *
 * There is no CPU path: every compute entry point fails with an error if its tensors are not on
 * a CUDA device.
 */
#ifndef DGL_B200_H_
#define DGL_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- DLTensor-compatible tensor view (dlpack.h: DLContext, DLDataType, DLTensor) ---- */
typedef struct {
  int32_t device_type; /* 1 = kDLCPU, 2 = kDLGPU, 3 = kDLCPUPinned */
  int32_t device_id;
} DGLB200Context;

typedef struct {
  uint8_t code; /* 0 = kDLInt, 1 = kDLUInt, 2 = kDLFloat */
  uint8_t bits;
  uint16_t lanes;
} DGLB200DataType;

typedef struct {
  void* data;
  DGLB200Context ctx;
  int32_t ndim;
  DGLB200DataType dtype;
  int64_t* shape;
  int64_t* strides; /* NULL = compact row-major (the only layout accepted, as in the reference) */
  uint64_t byte_offset;
} DGLB200Tensor;

/* Target codes: src/kernel/binary_reduce_common.h:39-44 == python/dgl/function/base.py:7-21 */
enum { DGLB200_SRC = 0, DGLB200_DST = 1, DGLB200_EDGE = 2, DGLB200_NONE = 3 };

/* Opaque device-resident graph index: in-CSR + out-CSR (int32 indptr / indices / edge ids) of a
 * (num_src x num_dst) graph, i.e. what CSRWrapper::GetInCSRMatrix / GetOutCSRMatrix hand to the
 * reference kernels (src/kernel/csr_interface.h:23-29, include/dgl/array.h:299-318). */
typedef struct DGLB200Graph* DGLB200GraphHandle;

/* ---- errors / library info ---- */
/* Replaces DGLGetLastError (include/dgl/runtime/c_runtime_api.h:166). */
const char* DGLB200GetLastError(void);
/* Returns the library ABI version (bumped on any signature change). */
int DGLB200Version(void);

/* ---- index preparation (replaces the single-threaded CPU path
 *      GraphIndex.get_immutable_gidx(ctx) -> ImmutableGraph::AsNumBits / CopyTo
 *      (python/dgl/graph_index.py:671-686, src/graph/immutable_graph.cc:548-590), i.e.
 *      aten::COOToCSR (src/array/cpu/spmat_op_impl_coo.cc:230-283) + aten::CSRTranspose
 *      (src/array/cpu/spmat_op_impl.cc:323-369) + narrowing to int32 + H2D copy) ---- */

/* Builds both CSRs ON THE DEVICE from a COO edge list (edge id = position in the list).
 * src/dst: int64 arrays of length num_edges, host or device memory (detected).
 * Results are bit-identical to the reference: in-CSR = stable counting sort by dst,
 * out-CSR = stable transpose of the in-CSR. `device` is the CUDA ordinal. */
int DGLB200GraphCreateFromCOO(int64_t num_src, int64_t num_dst, int64_t num_edges,
                              const int64_t* src, const int64_t* dst, int device, void* stream,
                              DGLB200GraphHandle* out);

/* Wraps CSR arrays that already exist (e.g. the outputs of the reference's
 * graph_index._CAPI_DGLGraphGetAdj(g, transpose, "csr"), src/graph/graph_apis.cc:285-292):
 * arrays may be int32 or int64 (`idx_bits`), host or device; they are copied to `device` as
 * int32. Replaces ImmutableGraph::CopyTo / UnitGraph::CopyTo (immutable_graph.cc:548-559,
 * unit_graph.cc:1146-1161). */
int DGLB200GraphCreateFromCSR(int64_t num_src, int64_t num_dst, int64_t num_edges, int idx_bits,
                              const void* in_indptr, const void* in_indices, const void* in_eids,
                              const void* out_indptr, const void* out_indices,
                              const void* out_eids, int device, void* stream,
                              DGLB200GraphHandle* out);

int DGLB200GraphFree(DGLB200GraphHandle g);

/* Bytes of index arrays the handle holds right now (CSR arrays and their lazily built
 * by-products, endpoints by edge id, typed adjacency; segment plans excluded). A graph built
 * from COO keeps 4 x E int32 ids + the two indptr arrays -- what the reference's in-CSR +
 * out-CSR hold (include/dgl/immutable_graph.h:242-246: indptr / indices / edge_ids per CSR) --
 * and rebuilds the plain in-CSR order, row expansions and position maps on demand. */
int DGLB200GraphIndexBytes(DGLB200GraphHandle g, int64_t* bytes);

/* num_src, num_dst, num_edges, device -> info[4]. */
int DGLB200GraphInfo(DGLB200GraphHandle g, int64_t* info);

/* Device pointers (int32) to one CSR: transpose=0 -> in-CSR (rows = dst), 1 -> out-CSR
 * (rows = src); same meaning as _CAPI_DGLGraphGetAdj(g, transpose, "csr"). */
int DGLB200GraphGetCSR(DGLB200GraphHandle g, int transpose, const int32_t** indptr,
                       const int32_t** indices, const int32_t** eids);

/* Copies one CSR (same `transpose` meaning) into caller buffers (host or device memory):
 * indptr [rows+1], indices [E], eids [E], int32. Synchronises `stream` when a buffer is on the
 * host. Any of the three may be NULL. */
int DGLB200GraphCopyCSR(DGLB200GraphHandle g, int transpose, int32_t* indptr, int32_t* indices,
                        int32_t* eids, void* stream);

/* Copies in-degrees (which=0) or out-degrees (which=1) as int64 into `out` (host or device),
 * = indptr[v+1]-indptr[v] (src/graph/immutable_graph.cc:177). Synchronises `stream`. */
int DGLB200GraphDegrees(DGLB200GraphHandle g, int which, int64_t* out, void* stream);

/* Standalone index ops on device int32/int64 arrays, bit-exact with the reference's CPU
 * functions (test surface of tests/cpp/test_spmat.cc). All pointers are DEVICE pointers.
 * COOToCSR: aten::COOToCSR (spmat_op_impl_coo.cc:230-283); data[] = COO position. */
int DGLB200COOToCSR(int64_t num_rows, int64_t nnz, const int32_t* row, const int32_t* col,
                    int32_t* indptr, int32_t* indices, int32_t* data, void* stream);
/* CSRTranspose: aten::CSRTranspose (spmat_op_impl.cc:323-369); data may be NULL (= position). */
int DGLB200CSRTranspose(int64_t num_rows, int64_t num_cols, int64_t nnz, const int32_t* indptr,
                        const int32_t* indices, const int32_t* data, int32_t* t_indptr,
                        int32_t* t_indices, int32_t* t_data, void* stream);
/* CSRToCOO row expansion: aten::CSRToCOO (spmat_op_impl.cc:371-385). */
int DGLB200CSRToCOORows(int64_t num_rows, int64_t nnz, const int32_t* indptr, int32_t* row,
                        void* stream);
/* AsNumBits 64 -> 32 (src/array/array.cc AsNumBits; ImmutableGraph::AsNumBits
 * immutable_graph.cc:574-590). */
int DGLB200AsInt32(int64_t n, const int64_t* in, int32_t* out, void* stream);

/* ---- kernels: one entry point per reference PackedFunc, same positional meaning ---- */

/* kernel._CAPI_DGLKernelInferBinaryFeatureShape(op, lhs, rhs) (src/kernel/binary_reduce.cc:274-293).
 * Pure host function (no GPU needed). Writes the broadcast output feature shape (for "dot" it
 * includes the reduced last dim, exactly as the reference) into out_shape[0..*out_ndim). */
int DGLB200KernelInferBinaryFeatureShape(const char* op, const DGLB200Tensor* lhs,
                                         const DGLB200Tensor* rhs, int32_t* out_ndim,
                                         int64_t* out_shape /* capacity >= 16 */);

/* kernel._CAPI_DGLKernelBinaryOpReduce(reducer, op, graph, lhs_tgt, rhs_tgt, lhs, rhs, out,
 *   lhs_map, rhs_map, out_map)  (src/kernel/binary_reduce.cc:295-378).
 * reducer in {"sum","max","min","prod","none"} ("mean" is rejected like the reference,
 * binary_reduce_impl.h:95-98); op in {"add","sub","mul","div","dot"}. `out` is caller-allocated
 * and fully overwritten (rows without in-edges receive the reducer identity: 0, -FLT_MAX,
 * FLT_MAX, 1 -- binary_reduce_common.h:447-470). */
int DGLB200KernelBinaryOpReduce(const char* reducer, const char* op, DGLB200GraphHandle graph,
                                int lhs_target, int rhs_target, const DGLB200Tensor* lhs_data,
                                const DGLB200Tensor* rhs_data, DGLB200Tensor* out_data,
                                const DGLB200Tensor* lhs_mapping,
                                const DGLB200Tensor* rhs_mapping,
                                const DGLB200Tensor* out_mapping, void* stream);

/* kernel._CAPI_DGLKernelBackwardLhsBinaryOpReduce(reducer, op, graph, lhs_tgt, rhs_tgt, lhs_map,
 *   rhs_map, out_map, lhs, rhs, out, grad_out, grad_lhs)  (binary_reduce.cc:452-527).
 * grad_lhs has the UNREDUCED out feature shape (rows = lhs rows), as in the reference. */
int DGLB200KernelBackwardLhsBinaryOpReduce(
    const char* reducer, const char* op, DGLB200GraphHandle graph, int lhs_target, int rhs_target,
    const DGLB200Tensor* lhs_mapping, const DGLB200Tensor* rhs_mapping,
    const DGLB200Tensor* out_mapping, const DGLB200Tensor* lhs_data,
    const DGLB200Tensor* rhs_data, const DGLB200Tensor* out_data,
    const DGLB200Tensor* grad_out_data, DGLB200Tensor* grad_lhs_data, void* stream);

/* kernel._CAPI_DGLKernelBackwardRhsBinaryOpReduce(...)  (binary_reduce.cc:551-626). */
int DGLB200KernelBackwardRhsBinaryOpReduce(
    const char* reducer, const char* op, DGLB200GraphHandle graph, int lhs_target, int rhs_target,
    const DGLB200Tensor* lhs_mapping, const DGLB200Tensor* rhs_mapping,
    const DGLB200Tensor* out_mapping, const DGLB200Tensor* lhs_data,
    const DGLB200Tensor* rhs_data, const DGLB200Tensor* out_data,
    const DGLB200Tensor* grad_out_data, DGLB200Tensor* grad_rhs_data, void* stream);

/* kernel._CAPI_DGLKernelCopyReduce(reducer, graph, target, in, out, in_map, out_map)
 * (binary_reduce.cc:628-665). */
int DGLB200KernelCopyReduce(const char* reducer, DGLB200GraphHandle graph, int target,
                            const DGLB200Tensor* in_data, DGLB200Tensor* out_data,
                            const DGLB200Tensor* in_mapping, const DGLB200Tensor* out_mapping,
                            void* stream);

/* kernel._CAPI_DGLKernelBackwardCopyReduce(reducer, graph, target, in, out, grad_out, grad_in,
 *   in_map, out_map)  (binary_reduce.cc:667-716). */
int DGLB200KernelBackwardCopyReduce(const char* reducer, DGLB200GraphHandle graph, int target,
                                    const DGLB200Tensor* in_data, const DGLB200Tensor* out_data,
                                    const DGLB200Tensor* grad_out_data,
                                    DGLB200Tensor* grad_in_data,
                                    const DGLB200Tensor* in_mapping,
                                    const DGLB200Tensor* out_mapping, void* stream);

/* kernel._CAPI_DGLFusedGatKernel(graph, feat_src, el, er, sum, exp, ret, slope)
 * (src/kernel/binary_reduce.cc:380-396; kernels src/kernel/cuda/binary_reduce_impl.cu:51-113,
 * 359-422). feat_src [N,H,F], el/er [N,H,1] (any shape with H elements per row), outputs
 * sum [N,H,1], exp [E,H,1] (indexed by edge id), ret [N,H,F]:
 *   exp = v > 0 ? e^v : e^(slope v), v = el[src] + er[dst];  sum = sum over in-edges;
 *   ret = sum over in-edges of (exp / sum) * feat_src[src]   (no max-subtraction, as the reference). */
int DGLB200FusedGatKernel(DGLB200GraphHandle graph, const DGLB200Tensor* feat_src,
                          const DGLB200Tensor* el, const DGLB200Tensor* er, DGLB200Tensor* sum,
                          DGLB200Tensor* exp, DGLB200Tensor* ret, float slope, void* stream);

/* kernel._CAPI_DGLKernelBackwardFusedGat(graph, feat_src, el, er, sum, exp, ret, grad_out,
 *   grad_feat_src, grad_el, grad_er, slope)  (binary_reduce.cc:529-549; kernels
 * binary_reduce_impl.cu:123-213, 1248-1308). The three gradients are fully overwritten
 * (the reference accumulates with atomics into zero-filled buffers; same values). */
int DGLB200KernelBackwardFusedGat(DGLB200GraphHandle graph, const DGLB200Tensor* feat_src,
                                  const DGLB200Tensor* el, const DGLB200Tensor* er,
                                  const DGLB200Tensor* sum, const DGLB200Tensor* exp,
                                  const DGLB200Tensor* ret, const DGLB200Tensor* grad_out,
                                  DGLB200Tensor* grad_feat_src, DGLB200Tensor* grad_el,
                                  DGLB200Tensor* grad_er, float slope, void* stream);

/* graph.add_edges_with_type (Graph::AddEdgesWithType, src/graph/graph.cc:140-165) followed by
 * Graph::GetCsrSortedByEdgeType (graph.cc:690-741): attaches a relation id to every edge id
 * (etypes[e], int32 or int64, host or device, 0 <= type < num_types) and builds the typed in- and
 * out-adjacency on the device: indptr is the untyped CSR's, every row is re-ordered by
 * (type, edge id). The reference orders rows with std::sort on the type only, so the order
 * inside one type is unspecified there; this is its stable form. */
int DGLB200GraphSetEdgeTypes(DGLB200GraphHandle graph, const void* etypes, int idx_bits,
                             int64_t num_types, void* stream);

/* The four int32 arrays GetCsrSortedByEdgeType(transpose) returns -- ranges [N+1], ids [E],
 * eids [E], types [E] -- copied into caller buffers (host or device; NULL skips one).
 * transpose = 0: in-edges (ids = sources), 1: out-edges (ids = destinations). */
int DGLB200GraphCopyCsrSortedByEdgeType(DGLB200GraphHandle graph, int transpose, int32_t* indptr,
                                        int32_t* ids, int32_t* eids, int32_t* types,
                                        void* stream);

/* The same typed graph rebuilt from the four arrays the reference's
 * graph_index._CAPI_DGLGraphGetCsrSortedByEdgeType(g, transpose = 0) returns
 * (src/graph/graph_apis.cc:295-301): in-edge ranges [N+1], source ids [E], edge ids [E], edge
 * types [E]; int32 or int64, host or device. num_types <= 0: max type + 1. This is how the
 * PackedFunc shims of the RGCN ops obtain the graph from a reference graph object. */
int DGLB200GraphCreateFromTypedCSR(int64_t num_nodes, int64_t num_edges, int idx_bits,
                                   const void* indptr, const void* ids, const void* eids,
                                   const void* types, int64_t num_types, int device, void* stream,
                                   DGLB200GraphHandle* out);

/* Batched per-relation message passing (BASELINE config 4: HeteroGraphConv loops R
 * BinaryOpReduce / CopyReduce calls over R UnitGraphs, python/dgl/nn/pytorch/hetero.py:120-171,
 * src/kernel/binary_reduce.cc:246-269; R = 133 launches that each write every destination row).
 * From a typed graph (DGLB200GraphSetEdgeTypes) builds the RELATION-SEGMENT graph: one destination
 * row per (relation, destination node) pair that receives at least one edge, sources and edge
 * ids unchanged. Every kernel entry point of this header accepts the new handle, so ONE
 * CopyReduce / BinaryOpReduce call on it computes what the reference's loop over the relations'
 * unit graphs computes, for all relations at once, writing each receiving (relation, node) row
 * exactly once: row s of the output = the reduction over the in-edges of relation seg_type[s]
 * into node seg_dst[s]; (relation, node) pairs without edges have no row (the loop's outputs hold
 * the reducer identity there). Segments are ordered by (relation, node), so the rows of one
 * relation are a contiguous slice; inside a segment, edges are reduced in the unit graph's own
 * order (source ascending, edge id). Synchronises `stream` once
 * (the number of segments is needed on the host); free the handle with DGLB200GraphFree. */
int DGLB200GraphCreateRelationSegments(DGLB200GraphHandle typed_graph, void* stream,
                                       DGLB200GraphHandle* out);
/* seg_dst[s], seg_type[s] of a relation-segment graph (num_dst entries each, int32; host or
 * device buffers, either may be NULL). */
int DGLB200GraphCopyRelationSegments(DGLB200GraphHandle segment_graph, int32_t* seg_dst,
                                     int32_t* seg_type, void* stream);

/* kernel._CAPI_DGLRgcnLayer0(graph, weight, norm, ret)  (src/kernel/binary_reduce.cc:412-420;
 * kernel src/kernel/cuda/binary_reduce_impl.cu:913-975). weight [R, N, F] (the embedding of a
 * one-hot input), norm [E,1] by edge id, ret [N, F]:
 *   ret[v,:] = sum over in-edges e of weight[type(e), src(e), :] * norm[e]. */
int DGLB200RgcnLayer0(DGLB200GraphHandle graph, const DGLB200Tensor* weight,
                      const DGLB200Tensor* norm, DGLB200Tensor* ret, void* stream);

/* kernel._CAPI_DGLRgcnLayer0Backward(graph, grad_out, norm, grad_weight)  (binary_reduce.cc:
 * 422-430; kernel binary_reduce_impl.cu:982-1046). Writes grad_weight[type, u, :] for every
 * (type, u) that has an out-edge; other entries keep the caller's zeros, as in the reference.
 * accumulate != 0: the gradient, sum over the edges u -> v of that type of grad_out[v,:]*norm[e].
 * accumulate == 0: the reference's plain store (:1004) -- the LAST such edge in (type, edge id)
 * order wins; equal to the sum only when no (type, u) pair has two out-edges. */
int DGLB200RgcnLayer0Backward(DGLB200GraphHandle graph, const DGLB200Tensor* grad_out,
                              const DGLB200Tensor* norm, DGLB200Tensor* grad_weight,
                              int accumulate, void* stream);

/* kernel._CAPI_DGLRgcnLayer1(graph, hidden, weight, norm, ret)  (binary_reduce.cc:432-440;
 * kernel binary_reduce_impl.cu:1082-1155). hidden [N, in], weight [R, in, out], ret [N, out]
 * (fully overwritten; the reference atomically adds into a zero-filled ret):
 *   ret[v,x] = sum_e sum_y hidden[src(e),y] * weight[type(e),y,x] * norm[e];  in, out <= 128. */
int DGLB200RgcnLayer1(DGLB200GraphHandle graph, const DGLB200Tensor* hidden,
                      const DGLB200Tensor* weight, const DGLB200Tensor* norm, DGLB200Tensor* ret,
                      void* stream);

/* kernel._CAPI_DGLRgcnLayer1Backward(graph, hidden, weight, norm, grad_out, grad_hidden,
 *   grad_weight)  (binary_reduce.cc:442-450; kernel binary_reduce_impl.cu:1157-1246).
 * grad_hidden [N, in] and grad_weight [R, in, out] are fully overwritten (the reference
 * atomicAdds into zero-filled buffers; same values, here without atomics and deterministic).
 * in * out <= 1024 as in the reference (its block size). */
int DGLB200RgcnLayer1Backward(DGLB200GraphHandle graph, const DGLB200Tensor* hidden,
                              const DGLB200Tensor* weight, const DGLB200Tensor* norm,
                              const DGLB200Tensor* grad_out, DGLB200Tensor* grad_hidden,
                              DGLB200Tensor* grad_weight, void* stream);

/* ---- extensions (no reference PackedFunc; they replace multi-call sequences the reference's
 *      torch backend performs around the kernels) ---- */

/* copy_{src,edge} + mean in ONE pass: out[v] = (sum_e in[sel]) / max(deg(v), 1). Replaces the
 * sum + second copy_reduce over a ones-vector + clamp + divide of
 * python/dgl/backend/pytorch/tensor.py:529-540 (same arithmetic: fp32 sum in in-CSR order, then
 * one IEEE division by the exact degree). No id mappings. */
int DGLB200KernelCopyReduceMean(DGLB200GraphHandle graph, int target,
                                const DGLB200Tensor* in_data, DGLB200Tensor* out_data,
                                void* stream);

/* edge_softmax over the in-edges of every destination node, forward and backward, ONE kernel
 * each (python/dgl/nn/pytorch/softmax.py:16-110 EdgeSoftmax runs copy_e+max, e_sub_v, exp,
 * copy_e+sum, e_div_v forward and copy_e+sum, e_mul_v backward through the builtin kernels).
 * `score` / `out` / `grad_out` / `grad_score` are [num_edges, ...] float32 in edge-id order with
 * identical shapes; softmax runs independently per trailing feature column. */
int DGLB200EdgeSoftmaxForward(DGLB200GraphHandle graph, const DGLB200Tensor* score,
                              DGLB200Tensor* out, void* stream);
int DGLB200EdgeSoftmaxBackward(DGLB200GraphHandle graph, const DGLB200Tensor* out,
                               const DGLB200Tensor* grad_out, DGLB200Tensor* grad_score,
                               void* stream);

/* GAT attention projections (python/dgl/nn/pytorch/conv/gatconv.py:143-144):
 *   el[n,h] = sum_f feat[n,h,f] * attn_l[h,f],  er[n,h] = sum_f feat[n,h,f] * attn_r[h,f]
 * in one pass over feat [N, H, F]; attn_* are [1, H, F] (or [H, F]); el / er are [N, H, 1].
 * attn_r / er (or attn_l / el) may be NULL to compute one projection only (bipartite blocks
 * project source and destination features separately). Backward: grad_feat [N, H, F] =
 * grad_el * attn_l + grad_er * attn_r and grad_attn_k [H, F] = sum_n grad_k[n,h] * feat[n,h,f]
 * (two-stage, fixed fold order). H * F <= 1024 for the backward. */
int DGLB200GatAttnProj(const DGLB200Tensor* feat, const DGLB200Tensor* attn_l,
                       const DGLB200Tensor* attn_r, DGLB200Tensor* el, DGLB200Tensor* er,
                       void* stream);
int DGLB200GatAttnProjBackward(const DGLB200Tensor* feat, const DGLB200Tensor* attn_l,
                               const DGLB200Tensor* attn_r, const DGLB200Tensor* grad_el,
                               const DGLB200Tensor* grad_er, DGLB200Tensor* grad_feat,
                               DGLB200Tensor* grad_attn_l, DGLB200Tensor* grad_attn_r,
                               void* stream);

/* DGLB200GatAttnProjBackward for callers that cannot allocate device memory (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the
 * reference. Same arguments, checks and results
 * (bit-identical on a B200: same blocking, same sum orders) as DGLB200GatAttnProjBackward, but the
 * block partials of grad_attn live in a file-scope `static __device__` array of the library
 * (768 row blocks x 2 x 512 floats = 3 MiB, held by every GPU that loads the library until
 * exit), so the call performs no cudaMalloc / cudaMallocAsync / cudaFree and calls no allocating
 * library routine. Limit: H * F <= 512; larger widths are rejected with an error (use the
 * allocating entry point). The array is shared by ALL calls on a device and keeps the previous
 * call's contents (every partial that is read is overwritten first, nothing needs clearing):
 * calls that overlap on the GPU -- different streams, or concurrent graph launches -- race; run
 * them in stream order. */
int DGLB200GatAttnProjBackwardNoAlloc(const DGLB200Tensor* feat, const DGLB200Tensor* attn_l,
                                      const DGLB200Tensor* attn_r, const DGLB200Tensor* grad_el,
                                      const DGLB200Tensor* grad_er, DGLB200Tensor* grad_feat,
                                      DGLB200Tensor* grad_attn_l, DGLB200Tensor* grad_attn_r,
                                      void* stream);

/* copy_src + sum / mean with bf16 FEATURE STORAGE and fp32 accumulation: a second, explicit
 * precision tier (north_star: 1e-2 relative in bf16). The reference kernels are fp32-only
 * (DGL_DTYPE_SWITCH, src/kernel/common.h:49-55), so this has no reference PackedFunc; it halves
 * the gathered bytes per edge. in_data / out_data are bfloat16 (DLPack code 4, 16 bits) [rows, D]
 * with D % 8 == 0 (pad the storage; 16-byte loads of 8 features), float32 tensors are rejected,
 * never converted silently. transpose = 0: out[v] = row_scale[v] * (1 / max(deg v, 1) if mean) *
 * sum_{u -> v} col_scale[u] * in[u] over the (src, edge id)-ordered in-CSR; transpose = 1: the
 * same over the out-CSR (the gradient). Scales are float32 vectors or NULL. Sums are accumulated
 * in fp32 in a fixed order (deterministic) and rounded to bf16 once. */
int DGLB200KernelCopySumBF16(DGLB200GraphHandle graph, int transpose, const DGLB200Tensor* in_data,
                             const DGLB200Tensor* col_scale, const DGLB200Tensor* row_scale,
                             int mean, DGLB200Tensor* out_data, void* stream);

/* The dense projection either side of the aggregation -- GraphConv's `feat @ weight`
 * (python/dgl/nn/pytorch/conv/graphconv.py:144-176), SAGEConv's fc_self / fc_neigh
 * (sageconv.py:150-160), GATConv's fc (gatconv.py:125-144) -- which the reference leaves to
 * cuBLAS' fp32 SIMT GEMM. Hand-written sm_100a kernel (csrc/gemm.cu): TMA-staged SWIZZLE_128B
 * operand tiles, tcgen05.mma kind::tf32 into TMEM accumulators, error-compensated 3xTF32
 * (a_hi.b_hi + a_hi.b_lo + a_lo.b_hi) so the result stays within ~1e-6 of the fp32 product
 * relative to sum|a||b| (plain TF32: 3e-4), bias / ReLU fused into the TMEM read-out.
 *   c[M, N] = a[M, K] . B (+ bias[N]) (ReLU);  b_is_kn != 0: B = b[K, N];  b_is_kn == 0: B = b[N, K]^T
 * (the second form is the gradient w.r.t. the input, grad_a = grad_c . W^T with b = W).
 * fp32 compact row-major tensors on one CUDA device; N <= 256, K % 4 == 0, a 16-byte aligned;
 * other shapes are rejected (callers fall back to the library GEMM). bias may be NULL. */
int DGLB200DenseTF32x3(const DGLB200Tensor* a, const DGLB200Tensor* b, int b_is_kn,
                       const DGLB200Tensor* bias, int relu, DGLB200Tensor* c, void* stream);

/* Weight gradient of the dense projection, c[K, N] = a[M, K]^T . g[M, N] (a = the layer's
 * input, g = the gradient of its output), same tensor-core arithmetic (3xTF32, fp32 accumulate in
 * TMEM). The contraction runs over the long dimension, so both operands are MN-major; every CTA
 * accumulates a full [K, N] partial over its chunk of rows and the partials are added in a fixed
 * order (deterministic). K, N <= 256 and divisible by 4; 16-byte aligned. */
int DGLB200DenseTF32x3WeightGrad(const DGLB200Tensor* a, const DGLB200Tensor* g, DGLB200Tensor* c,
                                 void* stream);

/* The dense PER-RELATION weight contraction of RGCN on the tensor cores (north_star (c); the
 * reference's RgcnLayer1 kernels contract per edge with CUDA-core FMAs and atomics,
 * src/kernel/cuda/binary_reduce_impl.cu:1082-1117, 1157-1194): rows [group_ptr[g],
 * group_ptr[g+1]) of a[M, K] are multiplied by b[g] ([groups, K, N] if b_is_kn, else
 * [groups, N, K] read transposed) into c[M, N]; same kernel and 3xTF32 arithmetic as
 * DGLB200DenseTF32x3, with a tile -> (rows, group) table. With the relation-segment graph the
 * RGCN layer is: per-(relation, node) sums (one SpMM launch) -> this contraction -> segment ->
 * node reduction. group_ptr is a HOST array of groups + 1 ascending row offsets from 0 to M.
 * Synchronises `stream` once when groups > 1 (the tile table is built on the host). */
int DGLB200GroupedDenseTF32x3(const DGLB200Tensor* a, const DGLB200Tensor* b, int b_is_kn,
                              const int64_t* group_ptr, int groups, DGLB200Tensor* c,
                              void* stream);

/* Uniform neighbour sampling on the device: dgl.sampling.sample_neighbors(g, seeds, fanout,
 * replace) / dgl.in_subgraph(g, seeds) (python/dgl/sampling/neighbor.py, used by
 * examples/pytorch/graphsage/train_sampling_multi_gpu.py:27-38). seeds[n_seeds] and
 * offsets[n_seeds + 1] are DEVICE int64 arrays; offsets is the exclusive prefix sum of the
 * per-seed output counts the caller derives from the in-degrees: deg for fanout < 0 (all
 * in-edges, in-CSR order), (deg > 0 ? fanout : 0) with replacement, min(deg, fanout) without.
 * Outputs (device int64, offsets[n_seeds] entries): source node, destination (= the seed) and
 * edge id of every sampled edge, grouped by seed; out_slot (may be NULL) receives the index of
 * the seed in seeds[] -- the destination id inside a block over the same seeds. Reproducible for
 * a given rng_seed.
 * DGLB200SampleCounts computes those offsets on the device: offsets is a device int64
 * [n_seeds + 2]; [0 .. n_seeds] = the exclusive prefix sum (so offsets[n_seeds] = number of
 * sampled edges, which the caller reads back to size the outputs), [n_seeds + 1] = status bits:
 * 1 a seed is outside [0, num_dst), 2 more than 2^31 - 1 sampled edges (offsets invalid). */
int DGLB200SampleCounts(DGLB200GraphHandle graph, const int64_t* seeds, int64_t n_seeds,
                        int fanout, int replace, int64_t* offsets, void* stream);
int DGLB200SampleNeighbors(DGLB200GraphHandle graph, const int64_t* seeds, int64_t n_seeds,
                           int fanout, int replace, uint64_t rng_seed, const int64_t* offsets,
                           int64_t* out_src, int64_t* out_dst, int64_t* out_eid,
                           int64_t* out_slot, void* stream);

/* Id compaction of dgl.to_block (python/dgl/transform.py to_block, used by
 * examples/pytorch/graphsage/train_sampling_multi_gpu.py:31-36) on DEVICE int64 arrays: block
 * source nodes are dst_nodes[n_dst] first, then every other value of src[n_edges] in order of
 * first appearance. Writes src_nid (capacity n_dst + n_edges; counts[0] entries are valid),
 * src_local[n_edges], and -- when dst is not NULL -- dst_local[n_edges], the index of dst[e] in
 * dst_nodes (dst == NULL: dst_local is not touched). counts is a device int64[2]: [0] number of
 * block source nodes, [1] status bits the caller must check after the stream has run: 1 = a
 * destination node appears twice, 2 = an edge ends outside dst_nodes, 4 = negative id. */
int DGLB200CompactBlock(const int64_t* dst_nodes, int64_t n_dst, const int64_t* src,
                        const int64_t* dst, int64_t n_edges, int64_t* src_nid, int64_t* src_local,
                        int64_t* dst_local, int64_t* counts, void* stream);

/* Aggregation with GraphConv's degree scaling inside the kernel (python/dgl/nn/pytorch/conv/
 * graphconv.py:128-176 does `feat * norm`, update_all(copy_src, sum), `rst * norm` as three
 * passes): transpose = 0: out[v] = row_scale[v] * sum_{u -> v} col_scale[u] * in[u];
 * transpose = 1 (its gradient): out[u] = row_scale[u] * sum_{u -> v} col_scale[v] * in[v].
 * Either scale may be NULL. Same summation orders as CopyReduce / BackwardCopyReduce. */
int DGLB200KernelScaledCopySum(DGLB200GraphHandle graph, int transpose,
                               const DGLB200Tensor* in_data, const DGLB200Tensor* col_scale,
                               const DGLB200Tensor* row_scale, DGLB200Tensor* out_data,
                               void* stream);

/* The same with a second term: out = row_scale * (addend + sum ...), addend shaped like out_data
 * and allowed to BE out_data (each element is read, then written, by one thread). This is the
 * second pass of a two-pass aggregation -- partitioned graphs add the halo-source edges to the
 * inner-source sums with it (python/dgl/transform.py:551-586 semantics: inner and halo edges of a
 * part) -- without a separate add kernel or a third pass over the rows. */
int DGLB200KernelScaledCopySumAdd(DGLB200GraphHandle graph, int transpose,
                                  const DGLB200Tensor* in_data, const DGLB200Tensor* col_scale,
                                  const DGLB200Tensor* row_scale, const DGLB200Tensor* addend,
                                  DGLB200Tensor* out_data, void* stream);

/* copy_src + sum over a feature matrix that is spread over several GPUs of one node (partitioned
 * graphs, SURVEY 8e option iii: peer loads inside the SpMM, no pack / send / unpack). Gathered
 * node c of `graph` reads row (node_code[c] & (2^27 - 1)) of block slot_base[node_code[c] >> 27];
 * every block is a row-major fp32 matrix with the width of out_data, in this GPU's memory or in a
 * peer's (CUDA IPC mapping; the loads travel over NVLink / NVSwitch). slot_base is a HOST array of
 * n_slots <= 16 device pointers, node_code a device int32 array with one code per gathered node.
 * transpose = 0: out[v] = row_scale[v] * sum_{u -> v} X[code(u)] over the (src, edge id)-ordered
 * in-CSR; transpose = 1: out[u] = row_scale[u] * sum_{u -> v} X[code(v)] over the out-CSR.
 * row_scale may be NULL. Ordering against the writers of the blocks is the caller's (barrier). */
int DGLB200KernelPeerCopySum(DGLB200GraphHandle graph, int transpose, int n_slots,
                             const void* const* slot_base, const DGLB200Tensor* node_code,
                             const DGLB200Tensor* row_scale, DGLB200Tensor* out_data,
                             void* stream);

/* Dispatch introspection for tests / bench: name of the kernel family the LAST compute call in
 * this process used ("spmm", "generic_row", "generic_edge", ...), and how many kernel launches
 * the library has issued since load (all threads). */
const char* DGLB200LastKernel(void);
int64_t DGLB200LaunchCount(void);

/* ---- node partitioning (host code; GraphOp::MetisPartition, src/graph/metis_partition.cc:19-66,
 *      = METIS_PartGraphKway(nvtxs, ncon=1, xadj, adjncy, nparts=k) on the symmetrised graph) ---- */

/* Multilevel k-way partition of a symmetric graph in CSR form (host int64 arrays, as the
 * reference hands METIS its in-CSR): part[v] in [0, k) minimising the edge cut with every part's
 * weight <= 1.03 * total / k (METIS' default imbalance). vwgt = NULL means unit vertex weights
 * (what the reference passes); otherwise one non-negative weight per vertex. *objval receives the
 * edge cut. Deterministic for a given seed; the data-parallel host phases run on
 * min(16, hardware threads) std::threads (B200_KWAY_THREADS overrides) and the result does not
 * depend on that number. METIS itself is a submodule the reference links; this
 * is an independent implementation of the same published scheme, so assignments differ while the
 * invariants the reference tests (tests/compute/test_transform.py:223-272) hold. */
int DGLB200MetisPartition(int64_t nvtxs, const int64_t* xadj, const int64_t* adjncy,
                          const int64_t* vwgt, int k, uint64_t seed, int64_t* part,
                          int64_t* objval);

/* The graph the reference hands METIS: to_bidirected(g) as an in-CSR (python/dgl/transform.py:
 * 625-626) from a COO edge list (host int64): self loops dropped, parallel edges merged,
 * neighbours ascending. xadj[nvtxs + 1] and adjncy[2 * num_edges] are caller-allocated; *nnz
 * receives the number of adjncy entries written. */
int DGLB200SymmetrizeCOO(int64_t nvtxs, int64_t num_edges, const int64_t* src, const int64_t* dst,
                         int64_t* xadj, int64_t* adjncy, int64_t* nnz);

/* ---- multi-GPU halo exchange over peer memory (no reference PackedFunc; the partition semantics
 *      are partition_graph_with_halo's, python/dgl/transform.py:551-640) ---- */

/* cudaMalloc + cudaIpcGetMemHandle: a buffer other ranks of the same node can map. */
int DGLB200PeerAlloc(int64_t bytes, int device, void** ptr, unsigned char handle[64]);
/* cudaIpcOpenMemHandle of another rank's buffer (peer access enabled lazily). */
int DGLB200PeerOpen(const unsigned char handle[64], int device, void** ptr);
int DGLB200PeerClose(void* ptr);
int DGLB200PeerFree(void* ptr);
/* Fused pack + send: row i of the push list (local row rows[i] of x, or row i when rows is NULL),
 * beg[q] <= i < beg[q+1], is stored into row dst_row[q] + (i - beg[q]) of peer[q] (row pitch D
 * floats). peer / beg / dst_row are host arrays, n_peers <= 16; rows is a device int32 array.
 * The list is walked starting at peer first_peer (wrapping), so that ranks passing rank + 1 do
 * not all write to the same peer at once.
 * Ordering against the readers is the caller's (alternating buffers + one barrier per exchange). */
int DGLB200HaloPush(const float* x, int64_t D, int n_peers, void* const* peer, const int64_t* beg,
                    const int64_t* dst_row, const int32_t* rows, int first_peer, int device,
                    void* stream);

/* Bulk pull of halo rows over NVLink: out[dst_row[i]] = row (codes[i] & (2^27 - 1)) of block
 * slot_base[codes[i] >> 27], i < n_rows, rows of D floats (dst_row NULL: out[i]). slot_base: HOST
 * array of n_slots <= 16 device pointers (own / peer-mapped blocks), codes, dst_row: device
 * int32. One streaming kernel whose loads are all remote, so the de-duplicated halo arrives at
 * the link rate (list the rows round-robin over their owners: grouped by owner, every rank of
 * the node reads from the same peer at the same time); the aggregation then gathers from local
 * memory only. max_blocks > 0 caps the grid (a pull running beside a compute kernel on another
 * stream leaves it most of every SM; 2 CTAs per SM still saturate the link), <= 0 = default. */
int DGLB200HaloPull(int n_slots, const void* const* slot_base, const int32_t* codes,
                    const int32_t* dst_row, int64_t n_rows, int64_t D, float* out, int max_blocks,
                    int device, void* stream);

/* Push of halo rows over NVLink (the mirror of DGLB200HaloPull; peer STORES are posted, so the
 * link is not bounded by outstanding remote reads): row (codes[i] & (2^27 - 1)) of buffer
 * slot_base[codes[i] >> 27] = scale[rows[i]] * x[rows[i]], i < n_rows, rows of D floats; scale
 * NULL = plain copy. slot_base: HOST array of n_slots <= 16 device pointers (the ranks' receive
 * buffers, own / peer-mapped), codes, rows: device int32, scale: device float over x's rows.
 * The partition plan lists the rows round-robin over the destination ranks. The receivers read
 * their buffers after a barrier that follows this kernel in stream order. No counterpart in the
 * reference (python/dgl/transform.py:551-640 defines what a part's halo is, not how it moves). */
int DGLB200HaloScatter(int n_slots, void* const* slot_base, const int32_t* codes,
                       const int32_t* rows, const float* scale, int64_t n_rows, int64_t D,
                       const float* x, int max_blocks, int device, void* stream);

/* ---- the PackedFunc face: drop-in registration into a loaded libdgl ----
 * DGLValue (include/dgl/runtime/c_runtime_api.h:122-129; DGLType / DGLContext members are
 * 8 bytes as well) and DGLPackedCFunc (:301-306). */
typedef union {
  int64_t v_int64;
  double v_float64;
  void* v_handle;
  const char* v_str;
} DGLB200Value;
typedef int (*DGLB200PackedCFunc)(DGLB200Value* args, int* type_codes, int num_args, void* ret,
                                  void* resource_handle);

/* Registers DGLPackedCFunc shims for the reference's 12 kernel globals --
 * kernel._CAPI_DGLKernel{InferBinaryFeatureShape, BinaryOpReduce, BackwardLhsBinaryOpReduce,
 * BackwardRhsBinaryOpReduce, CopyReduce, BackwardCopyReduce}, kernel._CAPI_DGLFusedGatKernel,
 * kernel._CAPI_DGLKernelBackwardFusedGat, kernel._CAPI_DGLRgcnLayer{0,0Backward,1,1Backward}
 * (src/kernel/binary_reduce.cc:281, 357, 380, 412-450, 502, 529, 600, 649, 697) -- into the
 * global function table of the libdgl whose dlopen() handle is passed, with override = 1
 * (DGLFuncCreateFromCFunc / DGLFuncRegisterGlobal, c_runtime_api.h:319-346). Afterwards the
 * unmodified python/dgl/kernel.py (bound by _init_api("dgl.kernel"), kernel.py:438) runs on this
 * library whenever the tensors are on a CUDA device; calls with CPU tensors are forwarded to
 * the functions that were registered before. Positional arguments are decoded exactly as the
 * reference reads args[i]: kStr, kDLInt, kDLFloat, kArrayHandle / kNDArrayContainer (a DLTensor*
 * either way), kNull, kObjectHandle for the graph, whose CSR arrays are read through
 * graph_index._CAPI_DGLGraphGetAdj / heterograph_index._CAPI_DGLHeteroGetAdj and cached per
 * (graph, device). Kernels run on the stream set by DGLB200SetStream (thread-local; default =
 * the legacy default stream, as the reference's CUDAThreadEntry, cuda_common.h:51-66).
 * Registering twice into the same process is a no-op. */
int DGLB200RegisterInto(void* libdgl_handle);
int DGLB200SetStream(void* stream);
/* The shims themselves, for hosts that register them on their own. */
int DGLB200PackedFuncCount(void);
const char* DGLB200PackedFuncName(int i);
DGLB200PackedCFunc DGLB200PackedFuncGet(int i);
/* out[0] = shim invocations, out[1] = of those forwarded to the previously registered (CPU)
 * function, out[2] = device graph indices built from reference graph objects. */
int DGLB200PackedStats(int64_t out[3]);
/* Drops every cached (reference graph -> device index) entry. */
int DGLB200PackedCacheClear(void);

#ifdef __cplusplus
}
#endif
#endif /* DGL_B200_H_ */

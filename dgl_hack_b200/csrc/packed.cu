// The PackedFunc face of the library: DGLPackedCFunc-signature entry points for the reference's
// `kernel._CAPI_*` globals and their registration into a loaded libdgl
// (include/dgl/runtime/c_runtime_api.h:255-362: DGLFuncCall, DGLPackedCFunc,
// DGLFuncCreateFromCFunc, DGLFuncRegisterGlobal; names and positional arguments as
// src/kernel/binary_reduce.cc:281-293, 357-378, 380-396, 412-450, 502-527, 529-549, 600-626,
// 649-665, 697-716 read `args[i]`). With the shims registered (override = 1) the UNMODIFIED
// python/dgl/kernel.py drives the sm_100a kernels: tensors arrive as NDArray containers /
// DLTensor handles (both start with a DLTensor, packed_func.h:408-418), the graph as an object
// handle that only libdgl can read -- its CSR arrays are fetched once per graph through the public
// PackedFunc `graph_index._CAPI_DGLGraphGetAdj(g, transpose, "csr")` (src/graph/graph_apis.cc:
// 285-292; `heterograph_index._CAPI_DGLHeteroGetAdj` for unit graphs) and wrapped with
// DGLB200GraphCreateFromCSR. Calls whose tensors are NOT on a CUDA device are forwarded to the
// function that was registered before (the reference's own CPU kernels), so the override only
// replaces the GPU path.
//
// This file has no link-time dependency on libdgl: every runtime function is looked up with
// dlsym on the handle the caller passes to DGLB200RegisterInto.
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace b200 {
namespace {

// DGLTypeCode (c_runtime_api.h:66-90) + DLPack's kDLInt / kDLFloat
enum { tInt = 0, tFloat = 2, tHandle = 3, tNull = 4, tArrayHandle = 7, tObjectHandle = 8,
       tFuncHandle = 10, tStr = 11, tNDArrayContainer = 13 };

typedef DGLB200Value Value;
typedef int (*PackedCFunc)(Value*, int*, int, void*, void*);

struct RefApi {
  int (*FuncGetGlobal)(const char*, void**) = nullptr;
  int (*FuncCall)(void*, Value*, int*, int, Value*, int*) = nullptr;
  int (*FuncFree)(void*) = nullptr;
  int (*FuncCreateFromCFunc)(PackedCFunc, void*, void (*)(void*), void**) = nullptr;
  int (*FuncRegisterGlobal)(const char*, void*, int) = nullptr;
  int (*CFuncSetReturn)(void*, Value*, int*, int) = nullptr;
  int (*ArrayAlloc)(const int64_t*, int, int, int, int, int, int, void**) = nullptr;
  int (*ArrayFree)(void*) = nullptr;
  void (*SetLastError)(const char*) = nullptr;
  const char* (*GetLastError)() = nullptr;
  bool ok() const {
    return FuncGetGlobal && FuncCall && FuncFree && FuncCreateFromCFunc && FuncRegisterGlobal &&
           CFuncSetReturn && ArrayAlloc && ArrayFree && SetLastError && GetLastError;
  }
};

struct Shim {           // resource handle of one registered function
  RefApi* api;
  const char* name;
  void* previous;       // the function registered under this name before (CPU path), may be NULL
};

RefApi g_api;
thread_local void* t_stream = nullptr;
std::atomic<int64_t> g_calls{0}, g_forwarded{0}, g_graphs_built{0};

// ---- argument decoding
struct Args {
  Value* v;
  int* c;
  int n;
  const char* fn;
  void need(int count) const {
    B200_CHECK(n == count, std::string(fn) + ": expected " + std::to_string(count) +
                               " arguments, got " + std::to_string(n));
  }
  const char* str(int i) const {
    B200_CHECK(c[i] == tStr, std::string(fn) + ": argument " + std::to_string(i) + " must be a string");
    return v[i].v_str;
  }
  int64_t integer(int i) const {
    B200_CHECK(c[i] == tInt, std::string(fn) + ": argument " + std::to_string(i) + " must be an int");
    return v[i].v_int64;
  }
  double real(int i) const {
    if (c[i] == tInt) return (double)v[i].v_int64;
    B200_CHECK(c[i] == tFloat, std::string(fn) + ": argument " + std::to_string(i) + " must be a float");
    return v[i].v_float64;
  }
  DGLB200Tensor* tensor(int i) const {
    if (c[i] == tNull) return nullptr;
    B200_CHECK(c[i] == tArrayHandle || c[i] == tNDArrayContainer,
               std::string(fn) + ": argument " + std::to_string(i) + " must be an NDArray");
    return static_cast<DGLB200Tensor*>(v[i].v_handle);
  }
  bool is_object(int i) const { return c[i] == tObjectHandle; }
};

static bool on_gpu(const DGLB200Tensor* t) {
  return t && t->ndim >= 1 && t->shape && t->shape[0] != 0 && t->ctx.device_type == 2;
}

// ---- calling back into libdgl
struct Returned {       // one value returned by DGLFuncCall, released on scope exit
  RefApi* api;
  Value v{};
  int code = tNull;
  explicit Returned(RefApi* a) : api(a) {}
  ~Returned() {
    if (code == tFuncHandle && v.v_handle) api->FuncFree(v.v_handle);
    if (code == tNDArrayContainer && v.v_handle) api->ArrayFree(v.v_handle);
  }
  void* release() {
    void* h = v.v_handle;
    v.v_handle = nullptr;
    return h;
  }
  Returned(const Returned&) = delete;
  Returned& operator=(const Returned&) = delete;
};

static void* global_func(RefApi* api, const char* name) {
  void* f = nullptr;
  if (api->FuncGetGlobal(name, &f) != 0) return nullptr;
  return f;
}

struct GlobalFunc {     // a looked-up global (DGLFuncGetGlobal hands out a copy the caller frees)
  RefApi* api;
  void* f;
  GlobalFunc(RefApi* a, const char* name) : api(a), f(global_func(a, name)) {}
  ~GlobalFunc() {
    if (f) api->FuncFree(f);
  }
};

// adjacency triple of `graph` through the reference's own PackedFuncs; false if the object is
// not of that kind (the reference throws on the handle type check)
static bool fetch_adj(RefApi* api, const Value& graph, bool hetero, bool transpose,
                      void* arrays[3]) {
  GlobalFunc get(api, hetero ? "heterograph_index._CAPI_DGLHeteroGetAdj"
                             : "graph_index._CAPI_DGLGraphGetAdj");
  if (!get.f) return false;
  Value a[4];
  int c[4];
  int n = 0;
  a[n] = graph; c[n++] = tObjectHandle;
  if (hetero) { a[n].v_int64 = 0; c[n++] = tInt; }       // edge type 0 of the unit graph
  a[n].v_int64 = transpose ? 1 : 0; c[n++] = tInt;
  a[n].v_str = "csr"; c[n++] = tStr;
  Returned list(api);
  if (api->FuncCall(get.f, a, c, n, &list.v, &list.code) != 0) return false;
  if (list.code != tFuncHandle) return false;
  for (int i = 0; i < 3; ++i) {
    Value idx;
    idx.v_int64 = i;
    int ic = tInt;
    Returned arr(api);
    if (api->FuncCall(list.v.v_handle, &idx, &ic, 1, &arr.v, &arr.code) != 0 ||
        arr.code != tNDArrayContainer) {
      for (int k = 0; k < i; ++k) api->ArrayFree(arrays[k]);
      return false;
    }
    arrays[i] = arr.release();
  }
  return true;
}

// ---- graph cache: one device index per (reference graph, device). The key is the address of
// the graph's own in-CSR indptr storage; the cache keeps a reference on that NDArray, so the
// address cannot be recycled by another graph while the entry lives.
struct CachedGraph {
  const void* key;
  int device;
  void* keep[3];
  DGLB200GraphHandle g;
  uint64_t used;
};
std::mutex g_cache_mu;
std::vector<CachedGraph> g_cache;
uint64_t g_tick = 0;
constexpr size_t kCacheCap = 32;

static void drop_entry(RefApi* api, CachedGraph& e) {
  for (void* h : e.keep)
    if (h) api->ArrayFree(h);
  DGLB200GraphFree(e.g);
}

static DGLB200GraphHandle resolve_graph(RefApi* api, const Value& graph, int device) {
  void* in[3] = {nullptr, nullptr, nullptr};
  bool hetero = false;
  if (!fetch_adj(api, graph, false, false, in)) {
    hetero = true;
    B200_CHECK(fetch_adj(api, graph, true, false, in),
               std::string("cannot read the graph's CSR through _CAPI_DGLGraphGetAdj / "
                           "_CAPI_DGLHeteroGetAdj: ") + api->GetLastError());
  }
  const DGLB200Tensor* ip = static_cast<const DGLB200Tensor*>(in[0]);
  const void* key = static_cast<const char*>(ip->data) + ip->byte_offset;
  std::lock_guard<std::mutex> lock(g_cache_mu);
  for (CachedGraph& e : g_cache) {
    if (e.key == key && e.device == device) {
      for (void* h : in) api->ArrayFree(h);
      e.used = ++g_tick;
      return e.g;
    }
  }
  void* out[3] = {nullptr, nullptr, nullptr};
  if (!fetch_adj(api, graph, hetero, true, out)) {
    for (void* h : in) api->ArrayFree(h);
    throw ApiError{std::string("cannot read the graph's out-CSR: ") + api->GetLastError()};
  }
  auto T = [](void* h) { return static_cast<const DGLB200Tensor*>(h); };
  auto P = [](const DGLB200Tensor* t) -> const void* {
    return static_cast<const char*>(t->data) + t->byte_offset;
  };
  const int bits = T(in[0])->dtype.bits;
  const int64_t num_dst = T(in[0])->shape[0] - 1, num_src = T(out[0])->shape[0] - 1;
  const int64_t num_edges = T(in[1])->shape[0];
  DGLB200GraphHandle g = nullptr;
  const int rc = DGLB200GraphCreateFromCSR(num_src, num_dst, num_edges, bits, P(T(in[0])),
                                           P(T(in[1])), P(T(in[2])), P(T(out[0])), P(T(out[1])),
                                           P(T(out[2])), device, t_stream, &g);
  for (void* h : out) api->ArrayFree(h);
  if (rc != 0) {
    for (void* h : in) api->ArrayFree(h);
    throw ApiError{DGLB200GetLastError()};
  }
  if (g_cache.size() >= kCacheCap) {
    auto lru = std::min_element(g_cache.begin(), g_cache.end(),
                                [](const CachedGraph& a, const CachedGraph& b) { return a.used < b.used; });
    drop_entry(api, *lru);
    g_cache.erase(lru);
  }
  g_cache.push_back(CachedGraph{key, device, {in[0], in[1], in[2]}, g, ++g_tick});
  g_graphs_built.fetch_add(1);
  return g;
}

// typed graphs of the RGCN hack ops: the reference hands the kernels
// Graph::GetCsrSortedByEdgeType(transpose) (src/graph/graph.cc:690-741), reachable from outside
// through graph_index._CAPI_DGLGraphGetCsrSortedByEdgeType (src/graph/graph_apis.cc:295-301)
static DGLB200GraphHandle resolve_typed_graph(RefApi* api, const Value& graph, int device) {
  GlobalFunc get(api, "graph_index._CAPI_DGLGraphGetCsrSortedByEdgeType");
  B200_CHECK(get.f, "graph_index._CAPI_DGLGraphGetCsrSortedByEdgeType is not registered");
  Value a[2];
  int c[2] = {tObjectHandle, tInt};
  a[0] = graph;
  a[1].v_int64 = 0;
  Returned list(api);
  B200_CHECK(api->FuncCall(get.f, a, c, 2, &list.v, &list.code) == 0 && list.code == tFuncHandle,
             std::string("_CAPI_DGLGraphGetCsrSortedByEdgeType failed: ") + api->GetLastError());
  void* arr[4] = {nullptr, nullptr, nullptr, nullptr};
  struct Free4 {
    RefApi* api; void** a;
    ~Free4() { for (int i = 0; i < 4; ++i) if (a[i]) api->ArrayFree(a[i]); }
  } guard{api, arr};
  for (int i = 0; i < 4; ++i) {
    Value idx;
    idx.v_int64 = i;
    int ic = tInt;
    Returned r(api);
    B200_CHECK(api->FuncCall(list.v.v_handle, &idx, &ic, 1, &r.v, &r.code) == 0 &&
                   r.code == tNDArrayContainer, "typed CSR: bad array");
    arr[i] = r.release();
  }
  const DGLB200Tensor* ip = static_cast<const DGLB200Tensor*>(arr[0]);
  const void* key = static_cast<const char*>(ip->data) + ip->byte_offset;
  std::lock_guard<std::mutex> lock(g_cache_mu);
  for (CachedGraph& e : g_cache)
    if (e.key == key && e.device == device) {
      e.used = ++g_tick;
      return e.g;
    }
  auto T = [](void* h) { return static_cast<const DGLB200Tensor*>(h); };
  auto P = [](const DGLB200Tensor* t) -> const void* {
    return static_cast<const char*>(t->data) + t->byte_offset;
  };
  DGLB200GraphHandle g = nullptr;
  const int rc = DGLB200GraphCreateFromTypedCSR(
      T(arr[0])->shape[0] - 1, T(arr[1])->shape[0], T(arr[0])->dtype.bits, P(T(arr[0])),
      P(T(arr[1])), P(T(arr[2])), P(T(arr[3])), -1, device, t_stream, &g);
  if (rc != 0) throw ApiError{DGLB200GetLastError()};
  if (g_cache.size() >= kCacheCap) {
    auto lru = std::min_element(g_cache.begin(), g_cache.end(),
                                [](const CachedGraph& a, const CachedGraph& b) { return a.used < b.used; });
    drop_entry(api, *lru);
    g_cache.erase(lru);
  }
  g_cache.push_back(CachedGraph{key, device, {arr[0], nullptr, nullptr}, g, ++g_tick});
  g_graphs_built.fetch_add(1);
  arr[0] = nullptr;  // kept by the cache
  return g;
}

// ---- shim bodies: decode, dispatch (GPU -> this library, otherwise -> the previous function)
static void forward_to_previous(Shim* sh, Value* v, int* c, int n, void* ret) {
  B200_CHECK(sh->previous != nullptr,
             std::string(sh->name) + ": tensors are not on a CUDA device and no previous "
                                     "(CPU) implementation was registered");
  g_forwarded.fetch_add(1);
  Returned r(sh->api);
  B200_CHECK(sh->api->FuncCall(sh->previous, v, c, n, &r.v, &r.code) == 0,
             std::string(sh->api->GetLastError()));
  if (r.code != tNull) sh->api->CFuncSetReturn(ret, &r.v, &r.code, 1);
}

static void own_call(int rc) {
  if (rc != 0) throw ApiError{DGLB200GetLastError()};
}

static int device_of(std::initializer_list<const DGLB200Tensor*> ts) {
  for (const DGLB200Tensor* t : ts)
    if (on_gpu(t)) return t->ctx.device_id;
  return -1;
}

typedef void (*Body)(Shim*, const Args&, void* ret);

template <Body BODY>
int shim_entry(Value* v, int* c, int n, void* ret, void* resource) {
  Shim* sh = static_cast<Shim*>(resource);
  g_calls.fetch_add(1);
  try {
    Args a{v, c, n, sh->name};
    BODY(sh, a, ret);
    return 0;
  } catch (const ApiError& e) {
    sh->api->SetLastError(e.msg.c_str());
  } catch (const std::exception& e) {
    sh->api->SetLastError(e.what());
  }
  return -1;
}

// kernel._CAPI_DGLKernelInferBinaryFeatureShape(op, lhs, rhs) -> int64 NDArray on the CPU
static void body_infer_shape(Shim* sh, const Args& a, void* ret) {
  a.need(3);
  int32_t nd = 0;
  int64_t shape[16];
  own_call(DGLB200KernelInferBinaryFeatureShape(a.str(0), a.tensor(1), a.tensor(2), &nd, shape));
  const int64_t len = nd;
  void* arr = nullptr;
  B200_CHECK(sh->api->ArrayAlloc(&len, 1, tInt, 64, 1, /*kDLCPU*/ 1, 0, &arr) == 0,
             std::string(sh->api->GetLastError()));
  DGLB200Tensor* t = static_cast<DGLB200Tensor*>(arr);
  std::memcpy(static_cast<char*>(t->data) + t->byte_offset, shape, sizeof(int64_t) * nd);
  Value rv;
  rv.v_handle = arr;
  int code = tNDArrayContainer;
  sh->api->CFuncSetReturn(ret, &rv, &code, 1);  // takes its own reference
  sh->api->ArrayFree(arr);
}

// (reducer, op, graph, lhs_tgt, rhs_tgt, lhs, rhs, out, lhs_map, rhs_map, out_map)
static void body_binary_op_reduce(Shim* sh, const Args& a, void* ret) {
  a.need(11);
  const int dev = device_of({a.tensor(5), a.tensor(6), a.tensor(7)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_graph(sh->api, a.v[2], dev);
  own_call(DGLB200KernelBinaryOpReduce(a.str(0), a.str(1), g, (int)a.integer(3), (int)a.integer(4),
                                       a.tensor(5), a.tensor(6), a.tensor(7), a.tensor(8),
                                       a.tensor(9), a.tensor(10), t_stream));
}

// (reducer, op, graph, lhs_tgt, rhs_tgt, lhs_map, rhs_map, out_map, lhs, rhs, out, grad_out, grad)
template <bool LHS>
static void body_backward_binary(Shim* sh, const Args& a, void* ret) {
  a.need(13);
  const int dev = device_of({a.tensor(8), a.tensor(9), a.tensor(10), a.tensor(11), a.tensor(12)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_graph(sh->api, a.v[2], dev);
  auto fn = LHS ? DGLB200KernelBackwardLhsBinaryOpReduce : DGLB200KernelBackwardRhsBinaryOpReduce;
  own_call(fn(a.str(0), a.str(1), g, (int)a.integer(3), (int)a.integer(4), a.tensor(5),
              a.tensor(6), a.tensor(7), a.tensor(8), a.tensor(9), a.tensor(10), a.tensor(11),
              a.tensor(12), t_stream));
}

// (reducer, graph, target, in, out, in_map, out_map)
static void body_copy_reduce(Shim* sh, const Args& a, void* ret) {
  a.need(7);
  const int dev = device_of({a.tensor(3), a.tensor(4)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_graph(sh->api, a.v[1], dev);
  own_call(DGLB200KernelCopyReduce(a.str(0), g, (int)a.integer(2), a.tensor(3), a.tensor(4),
                                   a.tensor(5), a.tensor(6), t_stream));
}

// (reducer, graph, target, in, out, grad_out, grad_in, in_map, out_map)
static void body_backward_copy_reduce(Shim* sh, const Args& a, void* ret) {
  a.need(9);
  const int dev = device_of({a.tensor(3), a.tensor(4), a.tensor(5), a.tensor(6)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_graph(sh->api, a.v[1], dev);
  own_call(DGLB200KernelBackwardCopyReduce(a.str(0), g, (int)a.integer(2), a.tensor(3),
                                           a.tensor(4), a.tensor(5), a.tensor(6), a.tensor(7),
                                           a.tensor(8), t_stream));
}

// (graph, feat_src, el, er, sum, exp, ret, slope)
static void body_fused_gat(Shim* sh, const Args& a, void* ret) {
  a.need(8);
  const int dev = device_of({a.tensor(1), a.tensor(6)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_graph(sh->api, a.v[0], dev);
  own_call(DGLB200FusedGatKernel(g, a.tensor(1), a.tensor(2), a.tensor(3), a.tensor(4),
                                 a.tensor(5), a.tensor(6), (float)a.real(7), t_stream));
}

// (graph, feat_src, el, er, sum, exp, ret, grad_out, grad_feat_src, grad_el, grad_er, slope)
static void body_backward_fused_gat(Shim* sh, const Args& a, void* ret) {
  a.need(12);
  const int dev = device_of({a.tensor(1), a.tensor(7)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_graph(sh->api, a.v[0], dev);
  own_call(DGLB200KernelBackwardFusedGat(g, a.tensor(1), a.tensor(2), a.tensor(3), a.tensor(4),
                                         a.tensor(5), a.tensor(6), a.tensor(7), a.tensor(8),
                                         a.tensor(9), a.tensor(10), (float)a.real(11), t_stream));
}

// (graph, weight, norm, ret)
static void body_rgcn_l0(Shim* sh, const Args& a, void* ret) {
  a.need(4);
  const int dev = device_of({a.tensor(1), a.tensor(3)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_typed_graph(sh->api, a.v[0], dev);
  own_call(DGLB200RgcnLayer0(g, a.tensor(1), a.tensor(2), a.tensor(3), t_stream));
}

// (graph, grad_out, norm, grad_weight): the reference's plain-store semantics (accumulate = 0)
static void body_rgcn_l0_bwd(Shim* sh, const Args& a, void* ret) {
  a.need(4);
  const int dev = device_of({a.tensor(1), a.tensor(3)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_typed_graph(sh->api, a.v[0], dev);
  own_call(DGLB200RgcnLayer0Backward(g, a.tensor(1), a.tensor(2), a.tensor(3), 0, t_stream));
}

// (graph, hidden, weight, norm, ret)
static void body_rgcn_l1(Shim* sh, const Args& a, void* ret) {
  a.need(5);
  const int dev = device_of({a.tensor(1), a.tensor(4)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_typed_graph(sh->api, a.v[0], dev);
  own_call(DGLB200RgcnLayer1(g, a.tensor(1), a.tensor(2), a.tensor(3), a.tensor(4), t_stream));
}

// (graph, hidden, weight, norm, grad_out, grad_hidden, grad_weight)
static void body_rgcn_l1_bwd(Shim* sh, const Args& a, void* ret) {
  a.need(7);
  const int dev = device_of({a.tensor(1), a.tensor(4)});
  if (dev < 0) return forward_to_previous(sh, a.v, a.c, a.n, ret);
  DGLB200GraphHandle g = resolve_typed_graph(sh->api, a.v[0], dev);
  own_call(DGLB200RgcnLayer1Backward(g, a.tensor(1), a.tensor(2), a.tensor(3), a.tensor(4),
                                     a.tensor(5), a.tensor(6), t_stream));
}

struct Entry {
  const char* name;
  PackedCFunc fn;
};

const Entry kEntries[] = {
    {"kernel._CAPI_DGLKernelInferBinaryFeatureShape", shim_entry<body_infer_shape>},
    {"kernel._CAPI_DGLKernelBinaryOpReduce", shim_entry<body_binary_op_reduce>},
    {"kernel._CAPI_DGLKernelBackwardLhsBinaryOpReduce", shim_entry<body_backward_binary<true>>},
    {"kernel._CAPI_DGLKernelBackwardRhsBinaryOpReduce", shim_entry<body_backward_binary<false>>},
    {"kernel._CAPI_DGLKernelCopyReduce", shim_entry<body_copy_reduce>},
    {"kernel._CAPI_DGLKernelBackwardCopyReduce", shim_entry<body_backward_copy_reduce>},
    {"kernel._CAPI_DGLFusedGatKernel", shim_entry<body_fused_gat>},
    {"kernel._CAPI_DGLKernelBackwardFusedGat", shim_entry<body_backward_fused_gat>},
    {"kernel._CAPI_DGLRgcnLayer0", shim_entry<body_rgcn_l0>},
    {"kernel._CAPI_DGLRgcnLayer0Backward", shim_entry<body_rgcn_l0_bwd>},
    {"kernel._CAPI_DGLRgcnLayer1", shim_entry<body_rgcn_l1>},
    {"kernel._CAPI_DGLRgcnLayer1Backward", shim_entry<body_rgcn_l1_bwd>},
};
constexpr int kNumEntries = sizeof(kEntries) / sizeof(kEntries[0]);

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int DGLB200PackedFuncCount(void) { return kNumEntries; }

const char* DGLB200PackedFuncName(int i) {
  return (i >= 0 && i < kNumEntries) ? kEntries[i].name : nullptr;
}

DGLB200PackedCFunc DGLB200PackedFuncGet(int i) {
  return (i >= 0 && i < kNumEntries) ? reinterpret_cast<DGLB200PackedCFunc>(kEntries[i].fn) : nullptr;
}

int DGLB200SetStream(void* stream) {
  t_stream = stream;
  return 0;
}

int DGLB200RegisterInto(void* libdgl_handle) {
  try {
    B200_CHECK(libdgl_handle != nullptr, "DGLB200RegisterInto: NULL library handle");
    RefApi api;
#define B200_SYM(field, sym)                                                                \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(libdgl_handle, sym));             \
  B200_CHECK(api.field != nullptr, std::string("DGLB200RegisterInto: the library does not export ") + sym)
    B200_SYM(FuncGetGlobal, "DGLFuncGetGlobal");
    B200_SYM(FuncCall, "DGLFuncCall");
    B200_SYM(FuncFree, "DGLFuncFree");
    B200_SYM(FuncCreateFromCFunc, "DGLFuncCreateFromCFunc");
    B200_SYM(FuncRegisterGlobal, "DGLFuncRegisterGlobal");
    B200_SYM(CFuncSetReturn, "DGLCFuncSetReturn");
    B200_SYM(ArrayAlloc, "DGLArrayAlloc");
    B200_SYM(ArrayFree, "DGLArrayFree");
    B200_SYM(SetLastError, "DGLAPISetLastError");
    B200_SYM(GetLastError, "DGLGetLastError");
#undef B200_SYM
    static bool registered = false;
    if (registered) return 0;
    g_api = api;
    registered = true;
    for (int i = 0; i < kNumEntries; ++i) {
      Shim* sh = new Shim{&g_api, kEntries[i].name, nullptr};  // lives as long as the registration
      sh->previous = global_func(&g_api, kEntries[i].name);
      void* f = nullptr;
      B200_CHECK(g_api.FuncCreateFromCFunc(kEntries[i].fn, sh, nullptr, &f) == 0,
                 std::string("DGLFuncCreateFromCFunc failed: ") + g_api.GetLastError());
      const int rc = g_api.FuncRegisterGlobal(kEntries[i].name, f, /*override=*/1);
      g_api.FuncFree(f);  // the registry holds its own copy
      B200_CHECK(rc == 0, std::string("DGLFuncRegisterGlobal failed: ") + g_api.GetLastError());
    }
    return 0;
  } catch (const ApiError& e) {
    set_last_error(e.msg);
  } catch (const std::exception& e) {
    set_last_error(e.what());
  }
  return -1;
}

int DGLB200PackedStats(int64_t out[3]) {
  out[0] = g_calls.load();
  out[1] = g_forwarded.load();
  out[2] = g_graphs_built.load();
  return 0;
}

int DGLB200PackedCacheClear(void) {
  std::lock_guard<std::mutex> lock(g_cache_mu);
  if (g_api.ok())
    for (CachedGraph& e : g_cache) drop_entry(&g_api, e);
  g_cache.clear();
  return 0;
}

}  // extern "C"

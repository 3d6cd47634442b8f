// Device-side index preparation and the graph handle.
//
// Replaces the reference's single-threaded CPU path for the arrays the kernels consume:
//   aten::COOToCSR      src/array/cpu/spmat_op_impl_coo.cc:230-283  (stable counting sort by row)
//   aten::CSRTranspose  src/array/cpu/spmat_op_impl.cc:323-369      (stable)
//   aten::CSRToCOO      src/array/cpu/spmat_op_impl.cc:371-385
//   AsNumBits / CopyTo  src/graph/immutable_graph.cc:548-590
// Both conversions are "stable sort of positions by a small integer key", done here with a
// hand-written LSD radix sort (8 bits per pass, only as many passes as the key range needs).
// Integer work: results are bit-identical to the reference.
#include <algorithm>
#include <vector>

#include "common.cuh"

namespace b200 {

// Graph arrays live in the device's stream-ordered memory pool (cudaMallocAsync) with the release
// threshold raised, so building and dropping a graph -- which the mini-batch path does for every
// sampled block -- costs no cudaMalloc / cudaFree round trips (each one a device-wide
// synchronisation plus ~100 us of driver work; 15 arrays per graph). An array is allocated on the
// stream that first writes it and freed in stream order (or after a device synchronise when the
// whole graph goes away).
static void pool_init(int device) {
  static bool done[64] = {};
  if (device < 0 || device >= 64 || done[device]) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
    uint64_t keep = UINT64_MAX;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  done[device] = true;
}

template <typename T>
static cudaError_t pool_alloc(T** p, size_t bytes, cudaStream_t s) {
  return cudaMallocAsync(reinterpret_cast<void**>(p), bytes ? bytes : 4, s);
}

static void pool_free(void* p, cudaStream_t s) {
  if (p) cudaFreeAsync(p, s);
}

// ---------------------------------------------------------------- small utilities
__global__ void fill_f32_kernel(float* __restrict__ p, int64_t n, float v) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // 128-bit stores on the aligned body
  if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
    const int64_t n4 = n >> 2;
    float4 v4 = make_float4(v, v, v, v);
    for (int64_t k = i; k < n4; k += stride) reinterpret_cast<float4*>(p)[k] = v4;
    for (int64_t k = (n4 << 2) + i; k < n; k += stride) p[k] = v;
  } else {
    for (int64_t k = i; k < n; k += stride) p[k] = v;
  }
}

void fill_f32(float* p, int64_t n, float v, cudaStream_t s) {
  if (n <= 0) return;
  int blocks = (int)std::min<int64_t>((n / 4 + 255) / 256 + 1, 148 * 16);
  fill_f32_kernel<<<blocks, 256, 0, s>>>(p, n, v);
  note_kernel(nullptr, 1);
}

__global__ void fill_i32_kernel(int32_t* __restrict__ p, int64_t n, int32_t v) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride) p[k] = v;
}
static void fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t s) {
  if (n <= 0) return;
  int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 16);
  fill_i32_kernel<<<blocks, 256, 0, s>>>(p, n, v);
  note_kernel(nullptr, 1);
}

__global__ void narrow_kernel(const int64_t* __restrict__ in, int32_t* __restrict__ out,
                              int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride)
    out[k] = (int32_t)in[k];
}
// int64 -> int32 with a range check against [0, hi): an id outside (negative, >= hi, or with
// high bits set) is replaced by 0 and reported through *bad, so that nothing downstream
// (count_keys atomics, radix scatter, gathers) can leave its arrays.
template <typename T>
__global__ void narrow_checked_kernel(const T* __restrict__ in, int32_t* __restrict__ out,
                                      int64_t n, int64_t hi, int32_t* __restrict__ bad) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  bool any = false;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride) {
    const int64_t v = (int64_t)in[k];
    const bool ok = v >= 0 && v < hi;
    out[k] = ok ? (int32_t)v : 0;
    any |= !ok;
  }
  if (any) atomicOr(bad, 1);
}

// indptr of a caller-provided CSR: starts at 0, ends at nnz, never decreases
__global__ void check_indptr_kernel(const int32_t* __restrict__ indptr, int64_t n_rows,
                                    int64_t nnz, int32_t* __restrict__ bad) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  bool any = false;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k <= n_rows; k += stride) {
    const int32_t v = indptr[k];
    if (k == 0 && v != 0) any = true;
    if (k == n_rows && v != nnz) any = true;
    if (k < n_rows && indptr[k + 1] < v) any = true;
    if (v < 0 || v > nnz) any = true;
  }
  if (any) atomicOr(bad, 1);
}

static void narrow(const int64_t* in, int32_t* out, int64_t n, cudaStream_t s) {
  if (n <= 0) return;
  int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 16);
  narrow_kernel<<<blocks, 256, 0, s>>>(in, out, n);
  note_kernel(nullptr, 1);
}

__global__ void widen_diff_kernel(const int32_t* __restrict__ indptr, int64_t* __restrict__ out,
                                  int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride)
    out[k] = (int64_t)(indptr[k + 1] - indptr[k]);
}

__global__ void gather_i32_kernel(const int32_t* __restrict__ table,
                                  const int32_t* __restrict__ idx, int32_t* __restrict__ out,
                                  int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride)
    out[k] = table[idx[k]];
}
static void gather_i32(const int32_t* table, const int32_t* idx, int32_t* out, int64_t n,
                       cudaStream_t s) {
  if (n <= 0) return;
  int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 16);
  gather_i32_kernel<<<blocks, 256, 0, s>>>(table, idx, out, n);
  note_kernel(nullptr, 1);
}

// out[idx[k]] = k  (inverse permutation)
__global__ void invert_perm_kernel(const int32_t* __restrict__ idx, int32_t* __restrict__ out,
                                   int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride)
    out[idx[k]] = (int32_t)k;
}

__global__ void count_keys_kernel(const int32_t* __restrict__ keys, int64_t n,
                                  int32_t* __restrict__ counts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride)
    atomicAdd(&counts[keys[k]], 1);
}

// row[p] = r for p in [indptr[r], indptr[r+1]); one warp per row, lanes stride the row.
__global__ void expand_rows_kernel(const int32_t* __restrict__ indptr, int64_t n_rows,
                                   int32_t* __restrict__ row) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n_rows; r += n_warps) {
    const int beg = indptr[r], end = indptr[r + 1];
    for (int p = beg + lane; p < end; p += 32) row[p] = (int32_t)r;
  }
}
static void expand_rows(const int32_t* indptr, int64_t n_rows, int32_t* row, cudaStream_t s) {
  if (n_rows <= 0) return;
  int blocks = (int)std::min<int64_t>((n_rows * 32 + 255) / 256, 148 * 32);
  expand_rows_kernel<<<blocks, 256, 0, s>>>(indptr, n_rows, row);
  note_kernel(nullptr, 1);
}

// ---------------------------------------------------------------- exclusive scan (int32)
// Three-phase: per-block reduce -> scan of block sums (recursive) -> per-block scan + offset.
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int block_exclusive_scan_256(int v, int* total) {
  // exclusive scan of one int per thread over 256 threads
  __shared__ int warp_sums[8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_sums[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    int w = lane < 8 ? warp_sums[lane] : 0;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    if (lane < 8) warp_sums[lane] = w;  // inclusive over warps
  }
  __syncthreads();
  const int warp_off = wid == 0 ? 0 : warp_sums[wid - 1];
  if (total) *total = warp_sums[7];
  const int res = warp_off + inc - v;
  __syncthreads();  // warp_sums reusable by the caller's next invocation
  return res;
}

__global__ void scan_reduce_kernel(const int32_t* __restrict__ data, int64_t n,
                                   int32_t* __restrict__ block_sums) {
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  int sum = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    const int64_t k = base + i * kScanThreads + threadIdx.x;
    if (k < n) sum += data[k];
  }
  int total;
  block_exclusive_scan_256(sum, &total);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

__global__ void scan_apply_kernel(int32_t* __restrict__ data, int64_t n,
                                  const int32_t* __restrict__ block_offsets) {
  // thread t owns kScanItems CONSECUTIVE elements so the scan order is the array order
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int v[kScanItems];
  int sum = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = (base + i < n) ? data[base + i] : 0;
    sum += v[i];
  }
  int run = block_exclusive_scan_256(sum, nullptr) + (block_offsets ? block_offsets[blockIdx.x] : 0);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) data[base + i] = run;
    run += v[i];
  }
}

void exclusive_scan_i32(int32_t* data, int64_t n, cudaStream_t s) {
  if (n <= 0) return;
  const int blocks = ceil_div(n, kScanTile);
  if (blocks == 1) {
    scan_apply_kernel<<<1, kScanThreads, 0, s>>>(data, n, nullptr);
    note_kernel(nullptr, 1);
    return;
  }
  Scratch sums((size_t)blocks * sizeof(int32_t), s);
  scan_reduce_kernel<<<blocks, kScanThreads, 0, s>>>(data, n, sums.as<int32_t>());
  note_kernel(nullptr, 1);
  exclusive_scan_i32(sums.as<int32_t>(), blocks, s);
  scan_apply_kernel<<<blocks, kScanThreads, 0, s>>>(data, n, sums.as<int32_t>());
  note_kernel(nullptr, 1);
}

// ---------------------------------------------------------------- stable LSD radix sort
// Sorts (key, value) pairs by 32-bit non-negative keys, 8 bits per pass; values default to the
// element index on the first pass. Stability is what makes the result equal to the reference's
// counting sort: ties keep their input order.
constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;  // per thread
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096
constexpr int kRadix = 256;

__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const int32_t* __restrict__ keys, int64_t n, int shift, int n_blocks,
                  int32_t* __restrict__ block_hist /* [kRadix][n_blocks] */) {
  __shared__ int hist[kRadix];
  hist[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kSortTile;
#pragma unroll
  for (int i = 0; i < kSortItems; ++i) {
    const int64_t k = base + i * kSortThreads + threadIdx.x;
    if (k < n) atomicAdd(&hist[(keys[k] >> shift) & (kRadix - 1)], 1);
  }
  __syncthreads();
  block_hist[(int64_t)threadIdx.x * n_blocks + blockIdx.x] = hist[threadIdx.x];
}

__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const int32_t* __restrict__ keys_in, const int32_t* __restrict__ vals_in,
                     int64_t n, int shift, int n_blocks,
                     const int32_t* __restrict__ block_offs /* scanned [kRadix][n_blocks] */,
                     int32_t* __restrict__ keys_out, int32_t* __restrict__ vals_out) {
  __shared__ int warp_cnt[kSortThreads / 32][kRadix];
  __shared__ int digit_base[kRadix];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (kSortThreads / 32) * kRadix; i += kSortThreads)
    (&warp_cnt[0][0])[i] = 0;
  digit_base[threadIdx.x] = block_offs[(int64_t)threadIdx.x * n_blocks + blockIdx.x];
  __syncthreads();

  // warp `wid` owns the contiguous slice [wbase, wbase + 32*kSortItems) of this block's tile
  const int64_t wbase = (int64_t)blockIdx.x * kSortTile + (int64_t)wid * (32 * kSortItems);
  int key[kSortItems], rank[kSortItems];
  const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const int64_t k = wbase + r * 32 + lane;
    const bool valid = k < n;
    key[r] = valid ? keys_in[k] : -1;
    const unsigned digit = valid ? (unsigned)((key[r] >> shift) & (kRadix - 1)) : 0xffffffffu;
    const unsigned peers = __match_any_sync(0xffffffffu, digit);
    const int before = __popc(peers & lt_mask);
    int old = 0;
    if (valid) old = warp_cnt[wid][digit];
    __syncwarp();
    if (valid && before == 0) warp_cnt[wid][digit] = old + __popc(peers);
    __syncwarp();
    rank[r] = old + before;
  }
  __syncthreads();
  {
    // exclusive scan of each digit's count over the warps (warp order = element order)
    const int d = threadIdx.x;
    int run = 0;
#pragma unroll
    for (int w = 0; w < kSortThreads / 32; ++w) {
      const int t = warp_cnt[w][d];
      warp_cnt[w][d] = run;
      run += t;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const int64_t k = wbase + r * 32 + lane;
    if (k < n) {
      const int digit = (key[r] >> shift) & (kRadix - 1);
      const int64_t pos = (int64_t)digit_base[digit] + warp_cnt[wid][digit] + rank[r];
      keys_out[pos] = key[r];
      vals_out[pos] = vals_in ? vals_in[k] : (int32_t)k;
    }
  }
}

// Sorts n pairs stably by keys in [0, key_range). On return *sorted_keys / *sorted_vals point at
// whichever of the ping-pong buffers holds the result. vals_in == nullptr means iota.
// keys_a is the input (may be overwritten if more than one pass is needed).
static void radix_sort_pairs(int32_t* keys_a, int32_t* keys_b, const int32_t* vals_in,
                             int32_t* vals_a, int32_t* vals_b, int64_t n, int64_t key_range,
                             int32_t** sorted_keys, int32_t** sorted_vals, cudaStream_t s) {
  int bits = 1;
  while (bits < 31 && ((int64_t)1 << bits) < key_range) ++bits;
  const int passes = std::max(1, (bits + 7) / 8);
  const int n_blocks = std::max(1, ceil_div(n, kSortTile));
  Scratch hist((size_t)kRadix * n_blocks * sizeof(int32_t), s);
  int32_t* kin = keys_a;
  int32_t* kout = keys_b;
  const int32_t* vin = vals_in;
  int32_t* vout = vals_a;
  for (int p = 0; p < passes; ++p) {
    const int shift = p * 8;
    radix_hist_kernel<<<n_blocks, kSortThreads, 0, s>>>(kin, n, shift, n_blocks,
                                                        hist.as<int32_t>());
    note_kernel(nullptr, 1);
    exclusive_scan_i32(hist.as<int32_t>(), (int64_t)kRadix * n_blocks, s);
    radix_scatter_kernel<<<n_blocks, kSortThreads, 0, s>>>(kin, vin, n, shift, n_blocks,
                                                           hist.as<int32_t>(), kout, vout);
    note_kernel(nullptr, 1);
    std::swap(kin, kout);
    vin = vout;
    vout = (vout == vals_a) ? vals_b : vals_a;
  }
  *sorted_keys = kin;
  *sorted_vals = const_cast<int32_t*>(vin);
}

// ---------------------------------------------------------------- CSR construction
// rows/cols: device int32 [nnz]. Produces indptr [n_rows+1], indices [nnz], data [nnz] with
// data[p] = original position, stable in input order (== aten::COOToCSR unsorted branch).
// If row_of_pos != nullptr it also receives the sorted row keys (the COO row expansion).
static void coo_to_csr_device(int64_t n_rows, int64_t nnz, const int32_t* rows,
                              const int32_t* cols, const int32_t* data_in, int32_t* indptr,
                              int32_t* indices, int32_t* data, int32_t* row_of_pos,
                              int32_t* src_pos /* optional: sorted position permutation */,
                              cudaStream_t s) {
  fill_i32(indptr, n_rows + 1, 0, s);
  if (nnz > 0) {
    int blocks = (int)std::min<int64_t>((nnz + 255) / 256, 148 * 16);
    count_keys_kernel<<<blocks, 256, 0, s>>>(rows, nnz, indptr);
    note_kernel(nullptr, 1);
  }
  exclusive_scan_i32(indptr, n_rows + 1, s);
  if (nnz == 0) return;
  Scratch ka((size_t)nnz * 4, s), kb((size_t)nnz * 4, s), va((size_t)nnz * 4, s),
      vb((size_t)nnz * 4, s);
  B200_CUDA(cudaMemcpyAsync(ka.ptr, rows, (size_t)nnz * 4, cudaMemcpyDeviceToDevice, s));
  int32_t *sk, *sv;
  radix_sort_pairs(ka.as<int32_t>(), kb.as<int32_t>(), nullptr, va.as<int32_t>(),
                   vb.as<int32_t>(), nnz, n_rows, &sk, &sv, s);
  // sv[p] = input position of the p-th sorted element
  gather_i32(cols, sv, indices, nnz, s);
  if (data_in) {
    gather_i32(data_in, sv, data, nnz, s);
  } else {
    B200_CUDA(cudaMemcpyAsync(data, sv, (size_t)nnz * 4, cudaMemcpyDeviceToDevice, s));
  }
  if (row_of_pos)
    B200_CUDA(cudaMemcpyAsync(row_of_pos, sk, (size_t)nnz * 4, cudaMemcpyDeviceToDevice, s));
  if (src_pos)
    B200_CUDA(cudaMemcpyAsync(src_pos, sv, (size_t)nnz * 4, cudaMemcpyDeviceToDevice, s));
}

// ---------------------------------------------------------------- lazily built per-CSR arrays
static void ensure_device(int device) {
  B200_CUDA(cudaSetDevice(device));
  pool_init(device);
}

const int32_t* csr_rows(DGLB200Graph* g, int walk, cudaStream_t s) {
  Csr& c = (walk == kWalkOut) ? g->out : g->in;  // in_sorted shares the in-CSR's indptr
  if (!c.row) {
    B200_CUDA(pool_alloc(&c.row, std::max<int64_t>(g->num_edges, 1) * 4, s));
    expand_rows(c.indptr, c.n_rows, c.row, s);
  }
  return c.row;
}

const int32_t* csr_to_other(DGLB200Graph* g, int walk, cudaStream_t s) {
  const bool out_csr = (walk == kWalkOut);
  if (walk != kWalkInSorted) in_plain(g, s);  // both directions read the in-CSR's edge ids
  Csr& c = out_csr ? g->out : (walk == kWalkInSorted ? g->in_s : g->in);
  Csr& o = out_csr ? g->in : g->out;
  if (!c.to_other) {
    const int64_t e = g->num_edges;
    B200_CUDA(pool_alloc(&c.to_other, std::max<int64_t>(e, 1) * 4, s));
    if (e > 0) {
      // pos_of_eid_in_other[o.eids[q]] = q ; to_other[p] = pos_of_eid_in_other[c.eids[p]]
      Scratch inv((size_t)e * 4, s);
      int blocks = (int)std::min<int64_t>((e + 255) / 256, 148 * 16);
      invert_perm_kernel<<<blocks, 256, 0, s>>>(o.eids, inv.as<int32_t>(), e);
      note_kernel(nullptr, 1);
      gather_i32(inv.as<int32_t>(), c.eids, c.to_other, e, s);
    }
  }
  return c.to_other;
}

// ---------------------------------------------------------------- segment plan
// Also emits the scheduling key of every row (degree class, longest first: rows that take the
// longest sequential walk start first, so the kernel's tail is made of the shortest rows; rows of
// one class keep their natural order) and the largest degree.
__global__ void seg_count_kernel(const int32_t* __restrict__ indptr, int64_t n_rows, int seg_len,
                                 int32_t* __restrict__ nseg, int32_t* __restrict__ heavy,
                                 int32_t* __restrict__ key, int32_t* __restrict__ max_deg) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int m = 0;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n_rows; r += stride) {
    const int deg = indptr[r + 1] - indptr[r];
    const bool h = deg > seg_len;
    nseg[r] = h ? (deg + seg_len - 1) / seg_len : 0;
    heavy[r] = h ? 1 : 0;
    // 4 classes per octave of degree; heavy rows (they only exit) go last
    int cls = 0;
    if (deg > 0) {
      const int lg = 31 - __clz(deg);
      cls = 4 * lg + (lg >= 2 ? ((deg >> (lg - 2)) & 3) : 0);
    }
    key[r] = h ? 255 : (127 - cls);
    m = max(m, deg);
    if (!h && deg >= kLongRow) atomicAdd(max_deg + 1, 1);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(max_deg, m);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    nseg[n_rows] = 0;
    heavy[n_rows] = 0;
  }
}

__global__ void seg_fill_kernel(const int32_t* __restrict__ indptr, int64_t n_rows, int seg_len,
                                const int32_t* __restrict__ seg_off,
                                const int32_t* __restrict__ heavy_idx,
                                int32_t* __restrict__ seg_row, int32_t* __restrict__ seg_beg,
                                int32_t* __restrict__ seg_end, int32_t* __restrict__ heavy_row,
                                int32_t* __restrict__ heavy_ptr) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n_rows; r += stride) {
    const int beg = indptr[r], end = indptr[r + 1];
    if (end - beg <= seg_len) continue;
    const int h = heavy_idx[r], o = seg_off[r];
    heavy_row[h] = (int32_t)r;
    heavy_ptr[h] = o;
    const int n = (end - beg + seg_len - 1) / seg_len;
    for (int k = 0; k < n; ++k) {
      seg_row[o + k] = (int32_t)r;
      seg_beg[o + k] = beg + k * seg_len;
      seg_end[o + k] = min(end, beg + (k + 1) * seg_len);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    heavy_ptr[heavy_idx[n_rows]] = seg_off[n_rows];  // sentinel: totals after the scans
}

// Keys of the "natural" schedule: heavy rows last, rows of >= kLongRow edges by degree class
// (longest first), every shorter row in ONE class -- the stable sort keeps their natural order.
// Also probes the numbering: of each row's first (up to) four neighbours, how many lie within
// 1/32 of the id range of the row. Community-ordered graphs (METIS-style numbering, the
// products-shaped SBM graph: 85 %) keep a community's feature rows in L2 when the rows are
// walked in natural order; on randomly numbered graphs (a few %) the order buys nothing and the
// degree classes pack the CTAs better.
__global__ void natural_keys_kernel(const int32_t* __restrict__ indptr,
                                    const int32_t* __restrict__ indices, int64_t n_rows,
                                    int seg_len, int32_t* __restrict__ key,
                                    unsigned long long* __restrict__ counts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t window = max((int64_t)1, n_rows / 32);
  unsigned near = 0, seen = 0;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n_rows; r += stride) {
    const int beg = indptr[r], deg = indptr[r + 1] - beg;
    int cls = 0;
    if (deg > 0) {
      const int lg = 31 - __clz(deg);
      cls = 4 * lg + (lg >= 2 ? ((deg >> (lg - 2)) & 3) : 0);
    }
    key[r] = deg > seg_len ? 255 : (deg >= kLongRow ? 127 - cls : 104);
    for (int k = 0; k < min(deg, 4); ++k) {
      const int64_t d = (int64_t)indices[beg + k] - r;
      near += (d < window && d > -window) ? 1u : 0u;
      ++seen;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    near += __shfl_xor_sync(0xffffffffu, near, o);
    seen += __shfl_xor_sync(0xffffffffu, seen, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(counts, (unsigned long long)near);
    atomicAdd(counts + 1, (unsigned long long)seen);
  }
}

static void free_plan(Csr::Plan& p, cudaStream_t s) {
  pool_free(p.sched_nat, s);
  pool_free(p.seg_row, s);
  pool_free(p.seg_beg, s);
  pool_free(p.seg_end, s);
  pool_free(p.heavy_row, s);
  pool_free(p.heavy_ptr, s);
  pool_free(p.sched, s);
  p = Csr::Plan();
}

SegPlan csr_segments(DGLB200Graph* g, int walk, int seg_len, cudaStream_t s, bool want_natural) {
  Csr& csr = (walk == kWalkOut) ? g->out : g->in;  // in_sorted has the in-CSR's degrees
  int slot = -1;
  for (int k = 0; k < Csr::kPlanSlots; ++k)
    if (csr.plan[k].seg_len == seg_len) slot = k;
  const bool build = slot < 0;
  if (build) {
    for (int k = 0; k < Csr::kPlanSlots && slot < 0; ++k)
      if (csr.plan[k].seg_len == 0) slot = k;
    if (slot < 0) {
      slot = csr.plan_next;
      csr.plan_next = (csr.plan_next + 1) % Csr::kPlanSlots;
      // a kernel enqueued earlier may still read the evicted plan
      B200_CUDA(cudaStreamSynchronize(s));
      free_plan(csr.plan[slot], s);
    }
  }
  struct View {  // the CSR's shape + the chosen slot's fields under the names used below
    Csr::Plan& p;
    int32_t*& seg_row; int32_t*& seg_beg; int32_t*& seg_end;
    int32_t*& heavy_row; int32_t*& heavy_ptr; int32_t*& sched;
    int32_t& n_seg; int32_t& n_heavy; int32_t& n_long; int32_t& seg_len;
    const int32_t* indptr; int64_t n_rows;
  };
  Csr::Plan& pl = csr.plan[slot];
  View c{pl, pl.seg_row, pl.seg_beg, pl.seg_end, pl.heavy_row, pl.heavy_ptr, pl.sched,
         pl.n_seg, pl.n_heavy, pl.n_long, pl.seg_len, csr.indptr, csr.n_rows};
  if (build) {
    c.n_seg = c.n_heavy = c.n_long = 0;
    if (c.n_rows > 0) {
      Scratch nseg((size_t)(c.n_rows + 1) * 4, s), heavy((size_t)(c.n_rows + 1) * 4, s);
      Scratch key((size_t)c.n_rows * 4, s), mdeg(8, s);  // [max degree, rows >= kLongRow]
      B200_CUDA(cudaMemsetAsync(mdeg.ptr, 0, 8, s));
      int blocks = (int)std::min<int64_t>((c.n_rows + 255) / 256, 148 * 8);
      seg_count_kernel<<<blocks, 256, 0, s>>>(c.indptr, c.n_rows, seg_len, nseg.as<int32_t>(),
                                              heavy.as<int32_t>(), key.as<int32_t>(),
                                              mdeg.as<int32_t>());
      note_kernel(nullptr, 1);
      exclusive_scan_i32(nseg.as<int32_t>(), c.n_rows + 1, s);
      exclusive_scan_i32(heavy.as<int32_t>(), c.n_rows + 1, s);
      int32_t tot[4] = {0, 0, 0, 0};
      B200_CUDA(cudaMemcpyAsync(&tot[2], mdeg.ptr, 8, cudaMemcpyDeviceToHost, s));
      B200_CUDA(cudaMemcpyAsync(&tot[0], nseg.as<int32_t>() + c.n_rows, 4, cudaMemcpyDeviceToHost, s));
      B200_CUDA(cudaMemcpyAsync(&tot[1], heavy.as<int32_t>() + c.n_rows, 4, cudaMemcpyDeviceToHost, s));
      B200_CUDA(cudaStreamSynchronize(s));
      ensure_valid_ids(g, s);
      c.n_seg = tot[0];
      c.n_heavy = tot[1];
      c.n_long = tot[3];
      // Skewed degrees: order the row work items by degree class. (Uniform-ish graphs keep the
      // natural order and with it whatever locality the node numbering has.)
      const int64_t mean = (g->num_edges + c.n_rows - 1) / c.n_rows;
      if ((int64_t)tot[2] > 4 * std::max<int64_t>(mean, 1) && c.n_rows > 1) {
        Scratch kb((size_t)c.n_rows * 4, s), va((size_t)c.n_rows * 4, s), vb((size_t)c.n_rows * 4, s);
        int32_t *sk, *sv;
        radix_sort_pairs(key.as<int32_t>(), kb.as<int32_t>(), nullptr, va.as<int32_t>(),
                         vb.as<int32_t>(), c.n_rows, 256, &sk, &sv, s);
        B200_CUDA(pool_alloc(&c.sched, (size_t)c.n_rows * 4, s));
        B200_CUDA(cudaMemcpyAsync(c.sched, sv, (size_t)c.n_rows * 4, cudaMemcpyDeviceToDevice, s));
      }
      if (c.n_seg > 0) {
        B200_CUDA(pool_alloc(&c.seg_row, (size_t)c.n_seg * 4, s));
        B200_CUDA(pool_alloc(&c.seg_beg, (size_t)c.n_seg * 4, s));
        B200_CUDA(pool_alloc(&c.seg_end, (size_t)c.n_seg * 4, s));
        B200_CUDA(pool_alloc(&c.heavy_row, (size_t)c.n_heavy * 4, s));
        B200_CUDA(pool_alloc(&c.heavy_ptr, (size_t)(c.n_heavy + 1) * 4, s));
        seg_fill_kernel<<<blocks, 256, 0, s>>>(c.indptr, c.n_rows, seg_len, nseg.as<int32_t>(),
                                               heavy.as<int32_t>(), c.seg_row, c.seg_beg,
                                               c.seg_end, c.heavy_row, c.heavy_ptr);
        note_kernel(nullptr, 1);
      }
    }
    c.seg_len = seg_len;
  }
  // ---- natural schedule (lazy): only for callers whose rows do not share a warp, only when a
  // degree schedule exists at all, and only if the numbering turns out to be local
  const int32_t* use_sched = c.sched;
  {
    const char* env = std::getenv("B200_SPMM_SCHED");  // "fine" / "natural" force one of them
    const bool force_fine = env && env[0] == 'f', force_nat = env && env[0] == 'n';
    if (want_natural && c.sched && !force_fine && csr.n_rows == csr.n_cols) {
      if (pl.nat_state == 0) {
        const int32_t* idx = (walk == kWalkOut) ? g->out.indices : g->in_s.indices;
        Scratch key((size_t)c.n_rows * 4, s), cnt(16, s);
        B200_CUDA(cudaMemsetAsync(cnt.ptr, 0, 16, s));
        const int blocks = (int)std::min<int64_t>((c.n_rows + 255) / 256, 148 * 8);
        natural_keys_kernel<<<blocks, 256, 0, s>>>(c.indptr, idx, c.n_rows, seg_len,
                                                   key.as<int32_t>(),
                                                   cnt.as<unsigned long long>());
        note_kernel(nullptr, 1);
        unsigned long long h[2] = {0, 0};
        B200_CUDA(cudaMemcpyAsync(h, cnt.ptr, 16, cudaMemcpyDeviceToHost, s));
        B200_CUDA(cudaStreamSynchronize(s));
        const bool local = h[1] > 0 && (double)h[0] >= 0.5 * (double)h[1];
        if (local || force_nat) {
          Scratch kb((size_t)c.n_rows * 4, s), va((size_t)c.n_rows * 4, s),
              vb((size_t)c.n_rows * 4, s);
          int32_t *sk, *sv;
          radix_sort_pairs(key.as<int32_t>(), kb.as<int32_t>(), nullptr, va.as<int32_t>(),
                           vb.as<int32_t>(), c.n_rows, 256, &sk, &sv, s);
          B200_CUDA(pool_alloc(&pl.sched_nat, (size_t)c.n_rows * 4, s));
          B200_CUDA(cudaMemcpyAsync(pl.sched_nat, sv, (size_t)c.n_rows * 4,
                                    cudaMemcpyDeviceToDevice, s));
          B200_CUDA(cudaStreamSynchronize(s));  // the scratch buffers go out of scope
          pl.nat_state = 1;
        } else {
          pl.nat_state = -1;
        }
      }
      if (pl.nat_state == 1) use_sched = pl.sched_nat;
    }
  }
  SegPlan p;
  p.seg_len = seg_len;
  p.n_seg = c.n_seg;
  p.n_heavy = c.n_heavy;
  p.n_long = c.sched ? c.n_long : 0;
  p.seg_row = c.seg_row;
  p.seg_beg = c.seg_beg;
  p.seg_end = c.seg_end;
  p.heavy_row = c.heavy_row;
  p.heavy_ptr = c.heavy_ptr;
  p.sched = use_sched;
  return p;
}

static void free_csr(Csr& c, cudaStream_t s = nullptr) {
  pool_free(c.indptr, s);
  pool_free(c.indices, s);
  pool_free(c.eids, s);
  pool_free(c.row, s);
  pool_free(c.to_other, s);
  for (int k = 0; k < Csr::kPlanSlots; ++k) free_plan(c.plan[k], s);
  c = Csr();
}

static void alloc_csr(Csr& c, int64_t n_rows, int64_t n_cols, int64_t nnz, cudaStream_t s) {
  c.n_rows = n_rows;
  c.n_cols = n_cols;
  B200_CUDA(pool_alloc(&c.indptr, (size_t)(n_rows + 1) * 4, s));
  B200_CUDA(pool_alloc(&c.indices, (size_t)std::max<int64_t>(nnz, 1) * 4, s));
  B200_CUDA(pool_alloc(&c.eids, (size_t)std::max<int64_t>(nnz, 1) * 4, s));
}

// in_s = Transpose(out-CSR): stable sort of the out-CSR positions by destination.
static void build_in_sorted(DGLB200Graph* g, cudaStream_t s) {
  const int64_t e = g->num_edges;
  Csr& c = g->in_s;
  c.n_rows = g->num_dst;
  c.n_cols = g->num_src;
  c.indptr = g->in.indptr;  // alias (never freed through in_s)
  B200_CUDA(pool_alloc(&c.indices, (size_t)std::max<int64_t>(e, 1) * 4, s));
  B200_CUDA(pool_alloc(&c.eids, (size_t)std::max<int64_t>(e, 1) * 4, s));
  B200_CUDA(pool_alloc(&c.to_other, (size_t)std::max<int64_t>(e, 1) * 4, s));
  if (e == 0) return;
  const int32_t* out_row = csr_rows(g, kWalkOut, s);
  Scratch ip((size_t)(g->num_dst + 1) * 4, s);
  coo_to_csr_device(g->num_dst, e, g->out.indices, out_row, g->out.eids, ip.as<int32_t>(),
                    c.indices, c.eids, nullptr, c.to_other, s);
}

// The plain in-CSR's indices / eids, rebuilt on demand: COOToCSR keyed by dst of the edges in
// edge-id order (immutable_graph.cc:407-421), the endpoints by edge id coming from the out-CSR.
const Csr& in_plain(DGLB200Graph* g, cudaStream_t s) {
  Csr& c = g->in;
  if (c.indices) return c;
  const int64_t e = g->num_edges;
  B200_CUDA(pool_alloc(&c.indices, (size_t)std::max<int64_t>(e, 1) * 4, s));
  B200_CUDA(pool_alloc(&c.eids, (size_t)std::max<int64_t>(e, 1) * 4, s));
  if (e == 0) return c;
  const bool had_coo = g->coo_src != nullptr;
  const int32_t* src_of = graph_coo(g, 0, s);
  const int32_t* dst_of = graph_coo(g, 1, s);
  Scratch ip((size_t)(g->num_dst + 1) * 4, s);
  coo_to_csr_device(g->num_dst, e, dst_of, src_of, nullptr, ip.as<int32_t>(), c.indices, c.eids,
                    nullptr, nullptr, s);
  if (!had_coo) {  // built for this only: not worth 2 x E resident ids
    pool_free(g->coo_src, s);
    pool_free(g->coo_dst, s);
    g->coo_src = g->coo_dst = nullptr;
  }
  return c;
}

// B200_INDEX_LEAN=0 keeps every by-product of the construction resident (A/B runs).
static bool index_lean() {
  const char* e = std::getenv("B200_INDEX_LEAN");
  return !(e && e[0] == '0');
}

// After construction from COO only the arrays the hot paths walk stay resident: in.indptr,
// out.(indptr, indices, eids), in_s.(indices, eids) = 4 x E ids, the reference's two CSRs
// (1.83 GB at the Reddit shape; 11 x E with every by-product kept). Row expansions, position
// maps and the plain in-CSR come back on demand (csr_rows, csr_to_other, in_plain).
static void drop_construction_byproducts(DGLB200Graph* g, bool keep_in_plain, cudaStream_t s) {
  if (!index_lean()) return;
  auto drop = [&](int32_t*& p) {
    pool_free(p, s);
    p = nullptr;
  };
  drop(g->in.row);
  drop(g->out.row);
  drop(g->out.to_other);
  drop(g->in_s.to_other);
  if (keep_in_plain) return;  // a caller-supplied in-CSR may order a row's edges its own way
  drop(g->in.indices);
  drop(g->in.eids);
}

// ---------------------------------------------------------------- typed adjacency
__global__ void scatter_coo_kernel(const int32_t* __restrict__ row, const int32_t* __restrict__ col,
                                   const int32_t* __restrict__ eid, int64_t n,
                                   int32_t* __restrict__ row_of, int32_t* __restrict__ col_of) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t e = eid[i];
    row_of[e] = row[i];
    col_of[e] = col[i];
  }
}

__global__ void check_range_kernel(const int32_t* __restrict__ v, int64_t n, int32_t hi,
                                   int32_t* __restrict__ bad) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    if (v[i] < 0 || v[i] >= hi) atomicOr(bad, 1);
}

const int32_t* graph_coo(DGLB200Graph* g, int which, cudaStream_t s) {
  if (!g->coo_src) {
    const int64_t e = g->num_edges;
    const size_t eb = (size_t)std::max<int64_t>(e, 1) * 4;
    int32_t *cs = nullptr, *cd = nullptr;
    B200_CUDA(pool_alloc(&cs, eb, s));
    B200_CUDA(pool_alloc(&cd, eb, s));
    if (e > 0) {
      // from the out-CSR (always resident): row = src, column = dst
      const int32_t* out_row = csr_rows(g, kWalkOut, s);
      const int blocks = (int)std::min<int64_t>((e + 255) / 256, 148 * 16);
      scatter_coo_kernel<<<blocks, 256, 0, s>>>(out_row, g->out.indices, g->out.eids, e, cs, cd);
      note_kernel(nullptr, 1);
    }
    g->coo_dst = cd;
    g->coo_src = cs;
  }
  return which ? g->coo_dst : g->coo_src;
}

static void free_typed(TypeIndex& t, cudaStream_t s = nullptr) {
  pool_free(t.in.ids, s);
  pool_free(t.in.eids, s);
  pool_free(t.in.types, s);
  pool_free(t.out.ids, s);
  pool_free(t.out.eids, s);
  pool_free(t.out.types, s);
  pool_free(t.bt_src, s);
  pool_free(t.bt_dst, s);
  pool_free(t.bt_eid, s);
  pool_free(t.type_ptr, s);
  t = TypeIndex();
}

// et: device int32 [E], type of every edge id. Orders: the reference sorts each adjacency row by
// type with std::sort (graph.cc:727-733), which leaves the order inside one type unspecified;
// here it is (type, edge id), i.e. the stable version of the same sort.
static void build_typed(DGLB200Graph* g, const int32_t* et, int64_t num_types, cudaStream_t s) {
  const int64_t e = g->num_edges;
  const size_t eb = (size_t)std::max<int64_t>(e, 1) * 4;
  TypeIndex& t = g->ty;
  free_typed(t, s);  // a previous typing: released in stream order
  int32_t** arrays[] = {&t.in.ids, &t.in.eids, &t.in.types, &t.out.ids, &t.out.eids,
                        &t.out.types, &t.bt_src, &t.bt_dst, &t.bt_eid};
  for (auto a : arrays) B200_CUDA(pool_alloc(a, eb, s));
  B200_CUDA(pool_alloc(&t.type_ptr, (size_t)(num_types + 1) * 4, s));
  t.num_types = num_types;
  fill_i32(t.type_ptr, num_types + 1, 0, s);
  if (e == 0) return;
  const int blocks = (int)std::min<int64_t>((e + 255) / 256, 148 * 16);
  {
    Scratch bad(4, s);
    fill_i32(bad.as<int32_t>(), 1, 0, s);
    check_range_kernel<<<blocks, 256, 0, s>>>(et, e, (int32_t)num_types, bad.as<int32_t>());
    note_kernel(nullptr, 1);
    int32_t h = 0;
    B200_CUDA(cudaMemcpyAsync(&h, bad.ptr, 4, cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));
    B200_CHECK(h == 0, "edge type id out of range [0, num_types)");
  }
  // endpoints by edge id
  const int32_t* src_of_p = graph_coo(g, 0, s);
  const int32_t* dst_of_p = graph_coo(g, 1, s);
  // type-major list: stable sort of edge ids by type
  count_keys_kernel<<<blocks, 256, 0, s>>>(et, e, t.type_ptr);
  note_kernel(nullptr, 1);
  exclusive_scan_i32(t.type_ptr, num_types + 1, s);
  Scratch ka(eb, s), kb(eb, s), va(eb, s), vb(eb, s), bt_type(eb, s), perm(eb, s),
      ip((size_t)(std::max(g->num_src, g->num_dst) + 1) * 4, s);
  B200_CUDA(cudaMemcpyAsync(ka.ptr, et, eb, cudaMemcpyDeviceToDevice, s));
  int32_t *sk, *sv;
  radix_sort_pairs(ka.as<int32_t>(), kb.as<int32_t>(), nullptr, va.as<int32_t>(),
                   vb.as<int32_t>(), e, num_types, &sk, &sv, s);
  B200_CUDA(cudaMemcpyAsync(t.bt_eid, sv, eb, cudaMemcpyDeviceToDevice, s));
  B200_CUDA(cudaMemcpyAsync(bt_type.ptr, sk, eb, cudaMemcpyDeviceToDevice, s));
  gather_i32(src_of_p, t.bt_eid, t.bt_src, e, s);
  gather_i32(dst_of_p, t.bt_eid, t.bt_dst, e, s);
  // rows: stable sort of the type-major list by destination (in) / source (out)
  coo_to_csr_device(g->num_dst, e, t.bt_dst, t.bt_src, t.bt_eid, ip.as<int32_t>(), t.in.ids,
                    t.in.eids, nullptr, perm.as<int32_t>(), s);
  gather_i32(bt_type.as<int32_t>(), perm.as<int32_t>(), t.in.types, e, s);
  coo_to_csr_device(g->num_src, e, t.bt_src, t.bt_dst, t.bt_eid, ip.as<int32_t>(), t.out.ids,
                    t.out.eids, nullptr, perm.as<int32_t>(), s);
  gather_i32(bt_type.as<int32_t>(), perm.as<int32_t>(), t.out.types, e, s);
}

static bool is_device_ptr(const void* p) {
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// copies `n` index values (int32 or int64, host or device) into a device int32 array
// upload_idx with the range check of narrow_checked_kernel (ids must lie in [0, hi))
static void upload_ids_checked(const void* src, int bits, int64_t n, int64_t hi, int32_t* dst,
                               int32_t* bad, cudaStream_t s) {
  if (n <= 0) return;
  const bool dev = is_device_ptr(src);
  const size_t bytes = (size_t)n * (bits / 8);
  Scratch tmp;
  const void* in = src;
  if (!dev) {
    tmp.alloc(bytes, s);
    B200_CUDA(cudaMemcpyAsync(tmp.ptr, src, bytes, cudaMemcpyHostToDevice, s));
    in = tmp.ptr;
  }
  const int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 16);
  if (bits == 64)
    narrow_checked_kernel<int64_t><<<blocks, 256, 0, s>>>(static_cast<const int64_t*>(in), dst, n, hi, bad);
  else
    narrow_checked_kernel<int32_t><<<blocks, 256, 0, s>>>(static_cast<const int32_t*>(in), dst, n, hi, bad);
  note_kernel(nullptr, 1);
  // the host buffer must stay valid until the copy has been consumed
  if (!dev) B200_CUDA(cudaStreamSynchronize(s));
}

static void read_bad_flag(DGLB200Graph* g, cudaStream_t s, const char* what) {
  int32_t h = 0;
  B200_CUDA(cudaMemcpyAsync(&h, g->bad_ids, 4, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  g->ids_checked = true;
  B200_CHECK(h == 0, what);
}

void ensure_valid_ids(DGLB200Graph* g, cudaStream_t s) {
  if (g->ids_checked || !g->bad_ids) return;
  read_bad_flag(g, s, "graph was built from node ids outside [0, num_src) x [0, num_dst)");
}

static void upload_idx(const void* src, int bits, int64_t n, int32_t* dst, cudaStream_t s) {
  if (n <= 0) return;
  const bool dev = is_device_ptr(src);
  if (bits == 32) {
    B200_CUDA(cudaMemcpyAsync(dst, src, (size_t)n * 4,
                              dev ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s));
  } else {
    if (dev) {
      narrow(static_cast<const int64_t*>(src), dst, n, s);
    } else {
      Scratch tmp((size_t)n * 8, s);
      B200_CUDA(cudaMemcpyAsync(tmp.ptr, src, (size_t)n * 8, cudaMemcpyHostToDevice, s));
      narrow(tmp.as<int64_t>(), dst, n, s);
      // the host buffer must stay valid until the copy has been consumed
      B200_CUDA(cudaStreamSynchronize(s));
    }
  }
}

}  // namespace b200

using namespace b200;

#define API_BEGIN() try {
#define API_END()                                     \
  }                                                   \
  catch (const ::b200::ApiError& e) {                 \
    ::b200::set_last_error(e.msg);                    \
    return -1;                                        \
  }                                                   \
  catch (const std::exception& e) {                   \
    ::b200::set_last_error(e.what());                 \
    return -1;                                        \
  }                                                   \
  return 0;

extern "C" {

int DGLB200GraphCreateFromCOO(int64_t num_src, int64_t num_dst, int64_t num_edges,
                              const int64_t* src, const int64_t* dst, int device, void* stream,
                              DGLB200GraphHandle* out) {
  API_BEGIN();
  B200_CHECK(out != nullptr, "DGLB200GraphCreateFromCOO: out handle is NULL");
  B200_CHECK(num_src >= 0 && num_dst >= 0 && num_edges >= 0, "negative graph size");
  // graph_index.py:941-952 picks 64-bit ids only if N or E >= 2^31; the GPU kernels of the
  // reference are int32-only (src/kernel/common.h:62-81) and so are these.
  B200_CHECK(num_src < (1ll << 31) - 1 && num_dst < (1ll << 31) - 1 && num_edges < (1ll << 31) - 1,
             "graph needs 64-bit ids; the GPU kernels are int32-only (as in the reference)");
  B200_CHECK(num_edges == 0 || (src && dst), "src/dst arrays are NULL");
  ensure_device(device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DGLB200Graph* g = new DGLB200Graph();
  g->num_src = num_src;
  g->num_dst = num_dst;
  g->num_edges = num_edges;
  g->device = device;
  try {
    alloc_csr(g->in, num_dst, num_src, num_edges, s);
    alloc_csr(g->out, num_src, num_dst, num_edges, s);
    const int64_t e = num_edges;
    Scratch s32((size_t)std::max<int64_t>(e, 1) * 4, s), d32((size_t)std::max<int64_t>(e, 1) * 4, s);
    B200_CUDA(pool_alloc(&g->bad_ids, 4, s));
    fill_i32(g->bad_ids, 1, 0, s);
    upload_ids_checked(src, 64, e, num_src, s32.as<int32_t>(), g->bad_ids, s);
    upload_ids_checked(dst, 64, e, num_dst, d32.as<int32_t>(), g->bad_ids, s);
    g->ids_checked = false;
    // host arrays: the uploads synchronised already, so the verdict costs nothing extra;
    // device arrays: deferred to the handle's first synchronisation point (ensure_valid_ids)
    if (e > 0 && !is_device_ptr(src))
      read_bad_flag(g, s, "DGLB200GraphCreateFromCOO: node id outside [0, num_src) x [0, num_dst)");
    // in-CSR = COOToCSR keyed by dst (immutable_graph.cc:407-421); keep the row expansion,
    // the transpose needs it.
    B200_CUDA(pool_alloc(&g->in.row, (size_t)std::max<int64_t>(e, 1) * 4, s));
    coo_to_csr_device(num_dst, e, d32.as<int32_t>(), s32.as<int32_t>(), nullptr, g->in.indptr,
                      g->in.indices, g->in.eids, g->in.row, nullptr, s);
    // out-CSR = in_csr.Transpose() (immutable_graph.cc:424-437): stable sort of in-CSR positions
    // by src; column = dst row of the position, data = its edge id. The sorted position
    // permutation is out->in ("to_other" of the out-CSR), kept because it is free here.
    B200_CUDA(pool_alloc(&g->out.to_other, (size_t)std::max<int64_t>(e, 1) * 4, s));
    coo_to_csr_device(num_src, e, g->in.indices, g->in.row, g->in.eids, g->out.indptr,
                      g->out.indices, g->out.eids, nullptr, g->out.to_other, s);
    build_in_sorted(g, s);
    drop_construction_byproducts(g, false, s);
  } catch (...) {
    g->in_s.indptr = nullptr;
    free_csr(g->in_s, s);
    free_csr(g->in, s);
    free_csr(g->out, s);
    pool_free(g->bad_ids, s);
    delete g;
    throw;
  }
  *out = g;
  API_END();
}

int DGLB200GraphCreateFromCSR(int64_t num_src, int64_t num_dst, int64_t num_edges, int idx_bits,
                              const void* in_indptr, const void* in_indices, const void* in_eids,
                              const void* out_indptr, const void* out_indices,
                              const void* out_eids, int device, void* stream,
                              DGLB200GraphHandle* out) {
  API_BEGIN();
  B200_CHECK(out != nullptr, "DGLB200GraphCreateFromCSR: out handle is NULL");
  B200_CHECK(idx_bits == 32 || idx_bits == 64, "idx_bits must be 32 or 64");
  B200_CHECK(num_src < (1ll << 31) - 1 && num_dst < (1ll << 31) - 1 && num_edges < (1ll << 31) - 1,
             "graph needs 64-bit ids; the GPU kernels are int32-only (as in the reference)");
  ensure_device(device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DGLB200Graph* g = new DGLB200Graph();
  g->num_src = num_src;
  g->num_dst = num_dst;
  g->num_edges = num_edges;
  g->device = device;
  try {
    alloc_csr(g->in, num_dst, num_src, num_edges, s);
    alloc_csr(g->out, num_src, num_dst, num_edges, s);
    B200_CUDA(pool_alloc(&g->bad_ids, 4, s));
    fill_i32(g->bad_ids, 1, 0, s);
    // every array is validated while it is narrowed: indptr values in [0, E] (and monotone,
    // 0 .. E), column ids in [0, n_cols), edge ids in [0, E)
    upload_ids_checked(in_indptr, idx_bits, num_dst + 1, num_edges + 1, g->in.indptr, g->bad_ids, s);
    upload_ids_checked(in_indices, idx_bits, num_edges, num_src, g->in.indices, g->bad_ids, s);
    upload_ids_checked(in_eids, idx_bits, num_edges, num_edges, g->in.eids, g->bad_ids, s);
    upload_ids_checked(out_indptr, idx_bits, num_src + 1, num_edges + 1, g->out.indptr, g->bad_ids, s);
    upload_ids_checked(out_indices, idx_bits, num_edges, num_dst, g->out.indices, g->bad_ids, s);
    upload_ids_checked(out_eids, idx_bits, num_edges, num_edges, g->out.eids, g->bad_ids, s);
    check_indptr_kernel<<<ceil_div(num_dst + 1, 256), 256, 0, s>>>(g->in.indptr, num_dst, num_edges, g->bad_ids);
    check_indptr_kernel<<<ceil_div(num_src + 1, 256), 256, 0, s>>>(g->out.indptr, num_src, num_edges, g->bad_ids);
    note_kernel(nullptr, 2);
    g->ids_checked = false;
    read_bad_flag(g, s, "DGLB200GraphCreateFromCSR: indptr / indices / edge ids out of range");
    build_in_sorted(g, s);
    drop_construction_byproducts(g, true, s);
    B200_CUDA(cudaStreamSynchronize(s));  // host buffers may be released by the caller now
  } catch (...) {
    g->in_s.indptr = nullptr;
    free_csr(g->in_s, s);
    free_csr(g->in, s);
    free_csr(g->out, s);
    pool_free(g->bad_ids, s);
    delete g;
    throw;
  }
  *out = g;
  API_END();
}

int DGLB200GraphFree(DGLB200GraphHandle g) {
  API_BEGIN();
  if (g) {
    cudaSetDevice(g->device);
    cudaDeviceSynchronize();  // nothing on any stream still reads the arrays
    cudaStream_t s = nullptr;
    free_typed(g->ty);
    pool_free(g->coo_src, s);
    pool_free(g->coo_dst, s);
    pool_free(g->bad_ids, s);
    pool_free(g->seg_dst, s);
    pool_free(g->seg_type, s);
    g->in_s.indptr = nullptr;
    free_csr(g->in_s);
    free_csr(g->in);
    free_csr(g->out);
    delete g;
  }
  API_END();
}

// Bytes of index arrays the handle holds right now (CSR arrays, lazily built by-products,
// COO by edge id, typed adjacency, segment plans excluded).
int DGLB200GraphIndexBytes(DGLB200GraphHandle g, int64_t* bytes) {
  API_BEGIN();
  B200_CHECK(g && bytes, "NULL argument");
  const int64_t e4 = std::max<int64_t>(g->num_edges, 1) * 4;
  int64_t n = 0;
  auto csr = [&](const Csr& c, bool own_indptr) {
    if (c.indptr && own_indptr) n += (c.n_rows + 1) * 4;
    for (const int32_t* p : {c.indices, c.eids, c.row, c.to_other})
      if (p) n += e4;
  };
  csr(g->in, true);
  csr(g->out, true);
  csr(g->in_s, false);
  for (const int32_t* p : {g->coo_src, g->coo_dst, g->ty.in.ids, g->ty.in.eids, g->ty.in.types,
                           g->ty.out.ids, g->ty.out.eids, g->ty.out.types, g->ty.bt_src,
                           g->ty.bt_dst, g->ty.bt_eid})
    if (p) n += e4;
  *bytes = n;
  API_END();
}

int DGLB200GraphInfo(DGLB200GraphHandle g, int64_t* info) {
  API_BEGIN();
  B200_CHECK(g && info, "NULL argument");
  info[0] = g->num_src;
  info[1] = g->num_dst;
  info[2] = g->num_edges;
  info[3] = g->device;
  API_END();
}

int DGLB200GraphGetCSR(DGLB200GraphHandle g, int transpose, const int32_t** indptr,
                       const int32_t** indices, const int32_t** eids) {
  API_BEGIN();
  B200_CHECK(g, "NULL graph");
  if (!transpose && !g->in.indices) {  // rebuilt on the default stream; usable from any stream
    ensure_device(g->device);
    in_plain(g, nullptr);
    B200_CUDA(cudaStreamSynchronize(nullptr));
  }
  const Csr& c = transpose ? g->out : g->in;
  if (indptr) *indptr = c.indptr;
  if (indices) *indices = c.indices;
  if (eids) *eids = c.eids;
  API_END();
}

int DGLB200GraphCopyCSR(DGLB200GraphHandle g, int transpose, int32_t* indptr, int32_t* indices,
                        int32_t* eids, void* stream) {
  API_BEGIN();
  B200_CHECK(g, "NULL graph");
  ensure_device(g->device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const Csr& c = transpose ? g->out : ((indices || eids) ? in_plain(g, s) : g->in);
  bool host = false;
  auto copy = [&](int32_t* dst, const int32_t* src, int64_t n) {
    if (!dst || n <= 0) return;
    const bool dev = is_device_ptr(dst);
    host = host || !dev;
    B200_CUDA(cudaMemcpyAsync(dst, src, (size_t)n * 4,
                              dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
  };
  copy(indptr, c.indptr, c.n_rows + 1);
  copy(indices, c.indices, g->num_edges);
  copy(eids, c.eids, g->num_edges);
  if (host) {
    B200_CUDA(cudaStreamSynchronize(s));
    ensure_valid_ids(g, s);
  }
  API_END();
}

// graph.add_edges_with_type (python/dgl/graph.py, Graph::AddEdgesWithType graph.cc:140-165) +
// Graph::GetCsrSortedByEdgeType (graph.cc:690-741): attaches a type to every edge id and builds
// both typed adjacencies on the device.
int DGLB200GraphSetEdgeTypes(DGLB200GraphHandle g, const void* etypes, int idx_bits,
                             int64_t num_types, void* stream) {
  API_BEGIN();
  B200_CHECK(g, "NULL graph");
  B200_CHECK(idx_bits == 32 || idx_bits == 64, "idx_bits must be 32 or 64");
  B200_CHECK(num_types > 0 && num_types < (1ll << 31) - 1, "num_types out of range");
  B200_CHECK(g->num_edges == 0 || etypes, "etypes is NULL");
  ensure_device(g->device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  Scratch et((size_t)std::max<int64_t>(g->num_edges, 1) * 4, s);
  upload_idx(etypes, idx_bits, g->num_edges, et.as<int32_t>(), s);
  try {
    build_typed(g, et.as<int32_t>(), num_types, s);
  } catch (...) {
    free_typed(g->ty, s);
    throw;
  }
  B200_CUDA(cudaStreamSynchronize(s));
  API_END();
}

// (position p of the typed in-CSR) -> COO by edge id: src[eid] = ids[p], dst[eid] = row(p),
// type[eid] = types[p]; an edge id outside [0, E) is reported and skipped
__global__ void typed_csr_to_coo_kernel(const int32_t* __restrict__ row,
                                        const int32_t* __restrict__ ids,
                                        const int32_t* __restrict__ eids,
                                        const int32_t* __restrict__ types, int64_t e,
                                        int64_t* __restrict__ src, int64_t* __restrict__ dst,
                                        int32_t* __restrict__ type_by_eid,
                                        int32_t* __restrict__ max_type, int32_t* __restrict__ bad) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int32_t mx = 0;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < e; p += stride) {
    const int32_t id = eids[p];
    if (id < 0 || id >= e) {
      atomicOr(bad, 1);
      continue;
    }
    src[id] = ids[p];
    dst[id] = row[p];
    const int32_t t = types[p];
    type_by_eid[id] = t;
    mx = max(mx, t);
  }
  atomicMax(max_type, mx);
}

int DGLB200GraphCreateFromTypedCSR(int64_t num_nodes, int64_t num_edges, int idx_bits,
                                   const void* indptr, const void* ids, const void* eids,
                                   const void* types, int64_t num_types, int device, void* stream,
                                   DGLB200GraphHandle* out) {
  API_BEGIN();
  B200_CHECK(out != nullptr, "DGLB200GraphCreateFromTypedCSR: out handle is NULL");
  B200_CHECK(idx_bits == 32 || idx_bits == 64, "idx_bits must be 32 or 64");
  B200_CHECK(num_nodes >= 0 && num_edges >= 0 && num_nodes < (1ll << 31) - 1 &&
                 num_edges < (1ll << 31) - 1, "graph size out of the int32 range");
  B200_CHECK(num_edges == 0 || (indptr && ids && eids && types), "NULL typed CSR array");
  ensure_device(device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t e = num_edges;
  const size_t eb = (size_t)std::max<int64_t>(e, 1);
  Scratch ip((size_t)(num_nodes + 1) * 4, s), row(eb * 4, s), id32(eb * 4, s), eid32(eb * 4, s),
      ty32(eb * 4, s), src(eb * 8, s), dst(eb * 8, s), tye(eb * 4, s), flags(8, s);
  fill_i32(flags.as<int32_t>(), 2, 0, s);
  int32_t* bad = flags.as<int32_t>();
  upload_ids_checked(indptr, idx_bits, num_nodes + 1, e + 1, ip.as<int32_t>(), bad, s);
  check_indptr_kernel<<<ceil_div(num_nodes + 1, 256), 256, 0, s>>>(ip.as<int32_t>(), num_nodes, e, bad);
  note_kernel(nullptr, 1);
  upload_ids_checked(ids, idx_bits, e, num_nodes, id32.as<int32_t>(), bad, s);
  upload_ids_checked(eids, idx_bits, e, e, eid32.as<int32_t>(), bad, s);
  upload_ids_checked(types, idx_bits, e, (1ll << 31) - 1, ty32.as<int32_t>(), bad, s);
  int32_t h[2] = {0, 0};
  B200_CUDA(cudaMemcpyAsync(h, flags.ptr, 4, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  B200_CHECK(h[0] == 0, "DGLB200GraphCreateFromTypedCSR: indptr / ids / edge ids / types out of range");
  if (e > 0) {
    expand_rows(ip.as<int32_t>(), num_nodes, row.as<int32_t>(), s);
    const int blocks = (int)std::min<int64_t>((e + 255) / 256, 148 * 16);
    typed_csr_to_coo_kernel<<<blocks, 256, 0, s>>>(
        row.as<int32_t>(), id32.as<int32_t>(), eid32.as<int32_t>(), ty32.as<int32_t>(), e,
        src.as<int64_t>(), dst.as<int64_t>(), tye.as<int32_t>(), flags.as<int32_t>() + 1, bad);
    note_kernel(nullptr, 1);
  }
  B200_CUDA(cudaMemcpyAsync(h, flags.ptr, 8, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  B200_CHECK(h[0] == 0, "DGLB200GraphCreateFromTypedCSR: edge ids are not a permutation of [0, E)");
  if (num_types <= 0) num_types = (int64_t)h[1] + 1;
  B200_CHECK(h[1] < num_types, "DGLB200GraphCreateFromTypedCSR: edge type >= num_types");
  DGLB200GraphHandle g = nullptr;
  B200_CHECK(DGLB200GraphCreateFromCOO(num_nodes, num_nodes, e, src.as<int64_t>(),
                                       dst.as<int64_t>(), device, stream, &g) == 0,
             std::string(DGLB200GetLastError()));
  if (DGLB200GraphSetEdgeTypes(g, tye.ptr, 32, num_types, stream) != 0) {
    const std::string msg = DGLB200GetLastError();
    DGLB200GraphFree(g);
    throw ApiError{msg};
  }
  *out = g;
  API_END();
}

// ---- relation segments: every maximal run of one edge type inside a destination row of the
// typed in-CSR is one (relation, destination) pair that receives messages
__global__ void seg_runs_count_kernel(const int32_t* __restrict__ indptr,
                                      const int32_t* __restrict__ types, int64_t n_rows,
                                      int32_t* __restrict__ counts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n_rows; r += stride) {
    const int32_t b = indptr[r], e = indptr[r + 1];
    int32_t c = 0, prev = -1;
    for (int32_t p = b; p < e; ++p) {
      const int32_t t = types[p];
      c += (t != prev);
      prev = t;
    }
    counts[r] = c;
  }
}

__global__ void seg_runs_fill_kernel(const int32_t* __restrict__ indptr,
                                     const int32_t* __restrict__ types,
                                     const int32_t* __restrict__ eids,
                                     const int32_t* __restrict__ src_of_eid, int64_t n_rows,
                                     const int32_t* __restrict__ seg_ptr,
                                     int32_t* __restrict__ seg_dst, int32_t* __restrict__ seg_type,
                                     int64_t* __restrict__ coo_src, int64_t* __restrict__ coo_seg) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n_rows; r += stride) {
    const int32_t b = indptr[r], e = indptr[r + 1];
    int32_t seg = seg_ptr[r] - 1, prev = -1;
    for (int32_t p = b; p < e; ++p) {
      const int32_t t = types[p];
      if (t != prev) {
        ++seg;
        seg_dst[seg] = (int32_t)r;
        seg_type[seg] = t;
        prev = t;
      }
      const int32_t id = eids[p];
      coo_src[id] = src_of_eid[id];
      coo_seg[id] = seg;
    }
  }
}

__global__ void remap_i64_kernel(int64_t* __restrict__ v, const int32_t* __restrict__ table,
                                 int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += stride)
    v[k] = table[v[k]];
}

int DGLB200GraphCreateRelationSegments(DGLB200GraphHandle g, void* stream,
                                       DGLB200GraphHandle* out) {
  API_BEGIN();
  B200_CHECK(g && out, "NULL argument");
  B200_CHECK(g->ty.num_types > 0, "graph has no edge types (call DGLB200GraphSetEdgeTypes)");
  ensure_device(g->device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t e = g->num_edges, n = g->num_dst;
  const size_t eb = (size_t)std::max<int64_t>(e, 1);
  Scratch seg_ptr((size_t)(n + 1) * 4, s), src64(eb * 8, s), seg64(eb * 8, s);
  fill_i32(seg_ptr.as<int32_t>(), n + 1, 0, s);
  const int rb = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, 148 * 16));
  if (n > 0) {
    seg_runs_count_kernel<<<rb, 256, 0, s>>>(g->in.indptr, g->ty.in.types, n, seg_ptr.as<int32_t>());
    note_kernel(nullptr, 1);
  }
  exclusive_scan_i32(seg_ptr.as<int32_t>(), n + 1, s);
  int32_t n_seg = 0;
  B200_CUDA(cudaMemcpyAsync(&n_seg, seg_ptr.as<int32_t>() + n, 4, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  ensure_valid_ids(g, s);
  int32_t *seg_dst = nullptr, *seg_type = nullptr;
  B200_CUDA(pool_alloc(&seg_dst, (size_t)std::max(n_seg, 1) * 4, s));
  B200_CUDA(pool_alloc(&seg_type, (size_t)std::max(n_seg, 1) * 4, s));
  DGLB200GraphHandle sg = nullptr;
  try {
    if (n > 0 && e > 0) {
      seg_runs_fill_kernel<<<rb, 256, 0, s>>>(g->in.indptr, g->ty.in.types, g->ty.in.eids,
                                              graph_coo(g, 0, s), n, seg_ptr.as<int32_t>(), seg_dst,
                                              seg_type, src64.as<int64_t>(), seg64.as<int64_t>());
      note_kernel(nullptr, 1);
    }
    // renumber the segments relation-major (stable: nodes ascending inside a relation), so that
    // the rows of one relation are contiguous -- "the output of relation r" is a slice, and a
    // grouped GEMM can contract each slice with that relation's weight
    if (n_seg > 1) {
      const size_t sb = (size_t)n_seg * 4;
      Scratch ka(sb, s), kb(sb, s), va(sb, s), vb(sb, s), inv(sb, s), dst_old(sb, s);
      B200_CUDA(cudaMemcpyAsync(ka.ptr, seg_type, sb, cudaMemcpyDeviceToDevice, s));
      B200_CUDA(cudaMemcpyAsync(dst_old.ptr, seg_dst, sb, cudaMemcpyDeviceToDevice, s));
      int32_t *sk, *sv;
      radix_sort_pairs(ka.as<int32_t>(), kb.as<int32_t>(), nullptr, va.as<int32_t>(),
                       vb.as<int32_t>(), n_seg, g->ty.num_types, &sk, &sv, s);
      B200_CUDA(cudaMemcpyAsync(seg_type, sk, sb, cudaMemcpyDeviceToDevice, s));
      gather_i32(dst_old.as<int32_t>(), sv, seg_dst, n_seg, s);
      const int sblocks = (int)std::min<int64_t>((n_seg + 255) / 256, 148 * 16);
      invert_perm_kernel<<<sblocks, 256, 0, s>>>(sv, inv.as<int32_t>(), n_seg);
      remap_i64_kernel<<<(int)std::min<int64_t>((e + 255) / 256, 148 * 16), 256, 0, s>>>(
          seg64.as<int64_t>(), inv.as<int32_t>(), e);
      note_kernel(nullptr, 2);
    }
    // sources -> segments, edge ids preserved: the standard constructor gives this graph the same
    // index invariants (and kernels, reducers, gradients) as any other
    B200_CHECK(DGLB200GraphCreateFromCOO(g->num_src, n_seg, e, src64.as<int64_t>(),
                                         seg64.as<int64_t>(), g->device, stream, &sg) == 0,
               std::string(DGLB200GetLastError()));
  } catch (...) {
    pool_free(seg_dst, s);
    pool_free(seg_type, s);
    throw;
  }
  sg->seg_dst = seg_dst;
  sg->seg_type = seg_type;
  *out = sg;
  API_END();
}

int DGLB200GraphCopyRelationSegments(DGLB200GraphHandle sg, int32_t* seg_dst, int32_t* seg_type,
                                     void* stream) {
  API_BEGIN();
  B200_CHECK(sg && sg->seg_dst, "not a relation-segment graph");
  ensure_device(sg->device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  bool host = false;
  auto copy = [&](int32_t* dst, const int32_t* src) {
    if (!dst || sg->num_dst <= 0) return;
    const bool dev = is_device_ptr(dst);
    host = host || !dev;
    B200_CUDA(cudaMemcpyAsync(dst, src, (size_t)sg->num_dst * 4,
                              dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
  };
  copy(seg_dst, sg->seg_dst);
  copy(seg_type, sg->seg_type);
  if (host) B200_CUDA(cudaStreamSynchronize(s));
  API_END();
}

int DGLB200GraphCopyCsrSortedByEdgeType(DGLB200GraphHandle g, int transpose, int32_t* indptr,
                                        int32_t* ids, int32_t* eids, int32_t* types,
                                        void* stream) {
  API_BEGIN();
  B200_CHECK(g, "NULL graph");
  B200_CHECK(g->ty.num_types > 0, "graph has no edge types (call DGLB200GraphSetEdgeTypes)");
  ensure_device(g->device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const Csr& c = transpose ? g->out : g->in;
  const TypedCsr& t = transpose ? g->ty.out : g->ty.in;
  bool host = false;
  auto copy = [&](int32_t* dst, const int32_t* src, int64_t n) {
    if (!dst || n <= 0) return;
    const bool dev = is_device_ptr(dst);
    host = host || !dev;
    B200_CUDA(cudaMemcpyAsync(dst, src, (size_t)n * 4,
                              dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
  };
  copy(indptr, c.indptr, c.n_rows + 1);
  copy(ids, t.ids, g->num_edges);
  copy(eids, t.eids, g->num_edges);
  copy(types, t.types, g->num_edges);
  if (host) {
    B200_CUDA(cudaStreamSynchronize(s));
    ensure_valid_ids(g, s);
  }
  API_END();
}

int DGLB200GraphDegrees(DGLB200GraphHandle g, int which, int64_t* out, void* stream) {
  API_BEGIN();
  B200_CHECK(g && out, "NULL argument");
  ensure_device(g->device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const Csr& c = which ? g->out : g->in;
  if (c.n_rows > 0) {
    const bool dev = is_device_ptr(out);
    Scratch tmp;
    int64_t* d = out;
    if (!dev) {
      tmp.alloc((size_t)c.n_rows * 8, s);
      d = tmp.as<int64_t>();
    }
    int blocks = (int)std::min<int64_t>((c.n_rows + 255) / 256, 148 * 16);
    widen_diff_kernel<<<blocks, 256, 0, s>>>(c.indptr, d, c.n_rows);
    note_kernel(nullptr, 1);
    if (!dev)
      B200_CUDA(cudaMemcpyAsync(out, d, (size_t)c.n_rows * 8, cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));
    ensure_valid_ids(g, s);
  }
  API_END();
}

int DGLB200COOToCSR(int64_t num_rows, int64_t nnz, const int32_t* row, const int32_t* col,
                    int32_t* indptr, int32_t* indices, int32_t* data, void* stream) {
  API_BEGIN();
  B200_CHECK(num_rows >= 0 && nnz >= 0, "negative size");
  coo_to_csr_device(num_rows, nnz, row, col, nullptr, indptr, indices, data, nullptr, nullptr,
                    static_cast<cudaStream_t>(stream));
  API_END();
}

int DGLB200CSRTranspose(int64_t num_rows, int64_t num_cols, int64_t nnz, const int32_t* indptr,
                        const int32_t* indices, const int32_t* data, int32_t* t_indptr,
                        int32_t* t_indices, int32_t* t_data, void* stream) {
  API_BEGIN();
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  Scratch row((size_t)std::max<int64_t>(nnz, 1) * 4, s);
  expand_rows(indptr, num_rows, row.as<int32_t>(), s);
  // transpose: key = column, new column = old row, data follows (or position if absent)
  coo_to_csr_device(num_cols, nnz, indices, row.as<int32_t>(), data, t_indptr, t_indices, t_data,
                    nullptr, nullptr, s);
  API_END();
}

int DGLB200CSRToCOORows(int64_t num_rows, int64_t nnz, const int32_t* indptr, int32_t* row,
                        void* stream) {
  API_BEGIN();
  (void)nnz;
  expand_rows(indptr, num_rows, row, static_cast<cudaStream_t>(stream));
  API_END();
}

int DGLB200AsInt32(int64_t n, const int64_t* in, int32_t* out, void* stream) {
  API_BEGIN();
  narrow(in, out, n, static_cast<cudaStream_t>(stream));
  API_END();
}

}  // extern "C"

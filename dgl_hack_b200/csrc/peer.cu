// Peer-memory halo exchange over NVLink: instead of pack -> NCCL send/recv -> unpack,
// ONE kernel reads the rows a peer needs from this rank's feature matrix and stores them straight
// into that peer's halo buffer (CUDA IPC mapping of a cudaMalloc'ed buffer; P2P stores travel over
// NVLink / NVSwitch at full rate and need no staging copy). The reference has no counterpart (its
// multi-GPU story is DDP over sampled mini-batches); the partition semantics follow
// python/dgl/transform.py:551-640 (partition_graph_with_halo).
//
// Synchronisation is the caller's (dgl_hack_b200/partition.py): pushes go to alternating buffers
// and every exchange ends with one stream-ordered NCCL barrier, so a buffer is rewritten only
// after every rank has passed the barrier that follows its readers.
#include <algorithm>

#include "common.cuh"

namespace b200 {

constexpr int kMaxPeers = 16;

struct PushArgs {
  const float* x;          // [*, D] local rows
  const int32_t* rows;     // [total] local row of every pushed row (NULL = identity)
  int64_t D;
  int n_peers;
  float* peer[kMaxPeers];      // base of each destination buffer (peer-mapped or local)
  int64_t beg[kMaxPeers + 1];  // rows [beg[q], beg[q+1]) of the push list go to peer q
  int64_t dst_row[kMaxPeers];  // first destination row inside peer q's buffer
  int64_t rot;                 // the push list is walked from here (wrapping): ranks start at
                               // different peers, so no peer's NVLink ingress sees all senders at once
};

// one warp per pushed row; 128-bit moves when the row pitch allows it
template <int VEC>
__global__ void __launch_bounds__(256) halo_push_kernel(const PushArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t total = a.beg[a.n_peers];
  const int nvec = (int)(a.D / VEC);
  for (int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; j < total;
       j += n_warps) {
    int64_t i = j + a.rot;
    if (i >= total) i -= total;
    int q = 0;
    while (i >= a.beg[q + 1]) ++q;
    const int64_t r = a.rows ? (int64_t)__ldg(a.rows + i) : i;
    const float* src = a.x + r * a.D;
    float* dst = a.peer[q] + (a.dst_row[q] + (i - a.beg[q])) * a.D;
    for (int c = lane; c < nvec; c += 32) {
      if (VEC == 4) {
        reinterpret_cast<float4*>(dst)[c] = __ldg(reinterpret_cast<const float4*>(src) + c);
      } else {
        dst[c] = __ldg(src + c);
      }
    }
  }
}

// Bulk pull of halo rows: out[dst_row[i]] = row (code & kPullRowMask) of block tab[code >> 27],
// code = codes[i] (dst_row == NULL: i). Every load is a peer load (NVLink). Work item = one
// 16-byte chunk of one row, consecutive threads = consecutive chunks (no idle lanes at D = 100),
// four items in flight per thread (64 B x 2048 threads = 128 KB per SM against the several-us
// NVLink round trip). The caller lists the rows ROUND-ROBIN over the owners (partition.py
// PeerGather): with the rows grouped by owner every rank of the node would read from peer 0
// first, then from peer 1, ... and one GPU's egress would serve all of them at a time.
constexpr int kPullSlotShift = 27;
constexpr int kPullRowMask = (1 << kPullSlotShift) - 1;
constexpr int kPullUnroll = 4;
struct PullArgs {
  const float* tab[kMaxPeers];
  const int32_t* codes;
  const int32_t* dst_row;
  int64_t n_rows, D;
  float* out;
};

template <int VEC>
__global__ void __launch_bounds__(256) halo_pull_kernel(const PullArgs a) {
  const int nvec = (int)(a.D / VEC);
  const int64_t total = a.n_rows * nvec;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; base < total;
       base += stride * kPullUnroll) {
    float4 v[kPullUnroll];
    int64_t o[kPullUnroll];
#pragma unroll
    for (int k = 0; k < kPullUnroll; ++k) {
      const int64_t idx = base + k * stride;
      o[k] = -1;
      if (idx < total) {
        const int64_t i = idx / nvec;
        const int c = (int)(idx - i * nvec);
        const int32_t code = __ldg(a.codes + i);
        const float* src = a.tab[(code >> kPullSlotShift) & (kMaxPeers - 1)] +
                           (int64_t)(code & kPullRowMask) * a.D + c * VEC;
        if (VEC == 4) v[k] = __ldg(reinterpret_cast<const float4*>(src));
        else v[k].x = __ldg(src);
        o[k] = (int64_t)(a.dst_row ? __ldg(a.dst_row + i) : i) * a.D + c * VEC;
      }
    }
#pragma unroll
    for (int k = 0; k < kPullUnroll; ++k) {
      if (o[k] >= 0) {
        if (VEC == 4) *reinterpret_cast<float4*>(a.out + o[k]) = v[k];
        else a.out[o[k]] = v[k].x;
      }
    }
  }
}

// Push of halo rows (the mirror of the pull): row rows[i] of x, times scale[rows[i]] when a scale
// is given, is stored as row (code & kPullRowMask) of buffer tab[code >> 27], code = codes[i].
// Every store is a peer store (posted: no round trip, so the link is not limited by the number
// of outstanding remote reads as the pull is). Same work split as the pull: one 16-byte chunk
// per item, four items per thread, the list interleaved over the destination ranks by the
// caller.
struct ScatterArgs {
  float* tab[kMaxPeers];
  const int32_t* codes;
  const int32_t* rows;
  const float* scale;
  const float* x;
  int64_t n_rows, D;
};

template <int VEC>
__global__ void __launch_bounds__(256) halo_scatter_kernel(const ScatterArgs a) {
  const int nvec = (int)(a.D / VEC);
  const int64_t total = a.n_rows * nvec;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; base < total;
       base += stride * kPullUnroll) {
    float4 v[kPullUnroll];
    float* o[kPullUnroll];
#pragma unroll
    for (int k = 0; k < kPullUnroll; ++k) {
      const int64_t idx = base + k * stride;
      o[k] = nullptr;
      if (idx < total) {
        const int64_t i = idx / nvec;
        const int c = (int)(idx - i * nvec);
        const int32_t code = __ldg(a.codes + i);
        const int64_t r = __ldg(a.rows + i);
        const float* src = a.x + r * a.D + c * VEC;
        if (VEC == 4) v[k] = __ldg(reinterpret_cast<const float4*>(src));
        else v[k].x = __ldg(src);
        if (a.scale) {
          const float sc = __ldg(a.scale + r);
          v[k].x = __fmul_rn(v[k].x, sc);
          if (VEC == 4) {
            v[k].y = __fmul_rn(v[k].y, sc);
            v[k].z = __fmul_rn(v[k].z, sc);
            v[k].w = __fmul_rn(v[k].w, sc);
          }
        }
        o[k] = a.tab[(code >> kPullSlotShift) & (kMaxPeers - 1)] +
               (int64_t)(code & kPullRowMask) * a.D + c * VEC;
      }
    }
#pragma unroll
    for (int k = 0; k < kPullUnroll; ++k) {
      if (o[k]) {
        if (VEC == 4) *reinterpret_cast<float4*>(o[k]) = v[k];
        else *o[k] = v[k].x;
      }
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

int DGLB200PeerAlloc(int64_t bytes, int device, void** ptr, unsigned char handle[64]) {
  API_BEGIN();
  B200_CHECK(ptr && handle && bytes > 0, "DGLB200PeerAlloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  B200_CUDA(cudaSetDevice(device));
  B200_CUDA(cudaMalloc(ptr, (size_t)bytes));
  cudaIpcMemHandle_t h;
  B200_CUDA(cudaIpcGetMemHandle(&h, *ptr));
  std::copy(reinterpret_cast<unsigned char*>(&h), reinterpret_cast<unsigned char*>(&h) + 64, handle);
  API_END();
}

int DGLB200PeerOpen(const unsigned char handle[64], int device, void** ptr) {
  API_BEGIN();
  B200_CHECK(ptr && handle, "DGLB200PeerOpen: bad argument");
  B200_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  std::copy(handle, handle + 64, reinterpret_cast<unsigned char*>(&h));
  B200_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  API_END();
}

int DGLB200PeerClose(void* ptr) {
  API_BEGIN();
  if (ptr) B200_CUDA(cudaIpcCloseMemHandle(ptr));
  API_END();
}

int DGLB200PeerFree(void* ptr) {
  API_BEGIN();
  if (ptr) B200_CUDA(cudaFree(ptr));
  API_END();
}

// Pushes rows of x into the peers' buffers: row i of the push list (local row rows[i], or i when
// rows is NULL) with beg[q] <= i < beg[q+1] lands in row dst_row[q] + (i - beg[q]) of peer[q].
// peer / beg / dst_row are HOST arrays (n_peers <= 16), rows is a device array.
int DGLB200HaloPush(const float* x, int64_t D, int n_peers, void* const* peer, const int64_t* beg,
                    const int64_t* dst_row, const int32_t* rows, int first_peer, int device,
                    void* stream) {
  API_BEGIN();
  B200_CHECK(n_peers >= 0 && n_peers <= kMaxPeers, "DGLB200HaloPush: at most 16 peers");
  B200_CHECK(D > 0 && peer && beg && dst_row, "DGLB200HaloPush: bad argument");
  PushArgs a{};
  a.x = x;
  a.rows = rows;
  a.D = D;
  a.n_peers = n_peers;
  bool al = (reinterpret_cast<uintptr_t>(x) % 16 == 0) && (D % 4 == 0);
  for (int q = 0; q < n_peers; ++q) {
    a.peer[q] = static_cast<float*>(peer[q]);
    a.beg[q] = beg[q];
    a.dst_row[q] = dst_row[q];
    al = al && (reinterpret_cast<uintptr_t>(peer[q]) % 16 == 0);
  }
  a.beg[n_peers] = beg[n_peers];
  const int64_t total = beg[n_peers];
  a.rot = (first_peer >= 0 && first_peer < n_peers) ? beg[first_peer] : 0;
  if (a.rot >= total) a.rot = 0;
  if (total == 0) return 0;
  B200_CUDA(cudaSetDevice(device));
  const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((total + 7) / 8, 148 * 16));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (al) halo_push_kernel<4><<<blocks, 256, 0, s>>>(a);
  else halo_push_kernel<1><<<blocks, 256, 0, s>>>(a);
  B200_CUDA(cudaGetLastError());
  note_kernel("halo_push", 1);
  API_END();
}

// Pulls n_rows rows of D floats: out[dst_row[i]] = block[codes[i] >> 27][codes[i] & (2^27 - 1)]
// (dst_row NULL: out[i]; max_blocks <= 0: default grid). slot_base is a HOST array of n_slots <= 16 device pointers (own or
// peer-mapped blocks of row pitch D), codes / dst_row DEVICE int32 arrays. Ordering against the
// writers of the blocks is the caller's.
int DGLB200HaloPull(int n_slots, const void* const* slot_base, const int32_t* codes,
                    const int32_t* dst_row, int64_t n_rows, int64_t D, float* out, int max_blocks,
                    int device, void* stream) {
  API_BEGIN();
  B200_CHECK(n_slots >= 1 && n_slots <= kMaxPeers && slot_base, "DGLB200HaloPull: 1 to 16 slots");
  B200_CHECK(D > 0 && n_rows >= 0, "DGLB200HaloPull: bad size");
  if (n_rows == 0) return 0;
  B200_CHECK(codes && out, "DGLB200HaloPull: NULL argument");
  PullArgs a{};
  bool al = (D % 4 == 0) && (reinterpret_cast<uintptr_t>(out) % 16 == 0);
  for (int q = 0; q < n_slots; ++q) {
    B200_CHECK(slot_base[q] != nullptr, "DGLB200HaloPull: NULL slot base");
    a.tab[q] = static_cast<const float*>(slot_base[q]);
    al = al && (reinterpret_cast<uintptr_t>(slot_base[q]) % 16 == 0);
  }
  a.codes = codes;
  a.dst_row = dst_row;
  a.n_rows = n_rows;
  a.D = D;
  a.out = out;
  B200_CUDA(cudaSetDevice(device));
  const int64_t items = n_rows * (al ? D / 4 : D);
  // max_blocks > 0 caps the grid: a pull that shares the GPU with a compute kernel (the
  // inner-edge SpMM of the split mode) leaves most registers / warp slots of every SM to it;
  // two CTAs per SM keep 32 KB of loads in flight per SM, enough for the link
  const int cap = max_blocks > 0 ? max_blocks : 148 * 8;
  const int blocks = (int)std::max<int64_t>(
      1, std::min<int64_t>((items + 256 * kPullUnroll - 1) / (256 * kPullUnroll), cap));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (al) halo_pull_kernel<4><<<blocks, 256, 0, s>>>(a);
  else halo_pull_kernel<1><<<blocks, 256, 0, s>>>(a);
  B200_CUDA(cudaGetLastError());
  note_kernel("halo_pull", 1);
  API_END();
}

// Pushes n_rows rows of D floats: buffer[codes[i] >> 27][codes[i] & (2^27 - 1)] =
// scale[rows[i]] * x[rows[i]] (scale NULL: plain copy). slot_base is a HOST array of n_slots <= 16
// device pointers (the ranks' receive buffers of row pitch D, own or peer-mapped), codes / rows
// DEVICE int32 arrays, scale a device float vector over x's rows. max_blocks as in
// DGLB200HaloPull. Ordering against the readers of the buffers is the caller's.
int DGLB200HaloScatter(int n_slots, void* const* slot_base, const int32_t* codes,
                       const int32_t* rows, const float* scale, int64_t n_rows, int64_t D,
                       const float* x, int max_blocks, int device, void* stream) {
  API_BEGIN();
  B200_CHECK(n_slots >= 1 && n_slots <= kMaxPeers && slot_base, "DGLB200HaloScatter: 1 to 16 slots");
  B200_CHECK(D > 0 && n_rows >= 0, "DGLB200HaloScatter: bad size");
  if (n_rows == 0) return 0;
  B200_CHECK(codes && rows && x, "DGLB200HaloScatter: NULL argument");
  ScatterArgs a{};
  bool al = (D % 4 == 0) && (reinterpret_cast<uintptr_t>(x) % 16 == 0);
  for (int q = 0; q < n_slots; ++q) {
    // a rank that receives nothing may have no buffer: its slot is never addressed
    a.tab[q] = static_cast<float*>(slot_base[q]);
    al = al && (reinterpret_cast<uintptr_t>(slot_base[q]) % 16 == 0);
  }
  a.codes = codes;
  a.rows = rows;
  a.scale = scale;
  a.x = x;
  a.n_rows = n_rows;
  a.D = D;
  B200_CUDA(cudaSetDevice(device));
  const int64_t items = n_rows * (al ? D / 4 : D);
  const int cap = max_blocks > 0 ? max_blocks : 148 * 8;
  const int blocks = (int)std::max<int64_t>(
      1, std::min<int64_t>((items + 256 * kPullUnroll - 1) / (256 * kPullUnroll), cap));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (al) halo_scatter_kernel<4><<<blocks, 256, 0, s>>>(a);
  else halo_scatter_kernel<1><<<blocks, 256, 0, s>>>(a);
  B200_CUDA(cudaGetLastError());
  note_kernel("halo_scatter", 1);
  API_END();
}

}  // extern "C"

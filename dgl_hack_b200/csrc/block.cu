// Block relabelling: the id compaction of dgl.to_block
// (python/dgl/transform.py to_block -> src/graph/transform/to_bipartite.cc in later DGL; used by
// examples/pytorch/graphsage/train_sampling_multi_gpu.py:31-36) as three linear passes over an
// open-addressing hash table instead of sort-based unique + argsort:
//
//   sequence  v[0 .. n) = dst_nodes ++ src                       (n = n_dst + n_edges)
//   insert    key table <- v[i]; first[slot] = min position i holding that value
//   flag      is_first[i] = (first[slot(i)] == i); exclusive scan -> new id of every first occurrence
//   emit      src_nid[id] = v[i] at first occurrences; src_local[e] = id(first[slot(n_dst + e)]);
//             dst_local[e] = id of dst[e] looked up in the table (must be one of dst_nodes)
//
// so block sources are the destination nodes first (ids 0 .. n_dst-1 when they are unique, which is
// checked), then every other source in order of first appearance -- the reference's order. The
// result does not depend on the launch shape: the only atomics are CAS on keys and min on
// positions, both order-independent.
#include <algorithm>

#include "common.cuh"

namespace b200 {
namespace {

constexpr unsigned long long kEmpty = ~0ull;

__device__ __forceinline__ uint32_t hash_slot(int64_t v, uint32_t mask) {
  uint64_t z = (uint64_t)v * 0x9E3779B97F4A7C15ull;
  z ^= z >> 29;
  return (uint32_t)z & mask;
}

struct BlockArgs {
  const int64_t *dst_nodes, *src, *dst;
  int64_t n_dst, n_edges, n;
  unsigned long long* keys;  // [cap]
  int32_t* first;            // [cap] smallest position of the key in v
  int32_t* slot_of;          // [n]
  int32_t* rank;             // [n + 1] flags, then their exclusive scan
  uint32_t mask;
  int64_t *src_nid, *src_local, *dst_local, *counts;  // counts = {n_src, status}
};

__device__ __forceinline__ int64_t value_at(const BlockArgs& a, int64_t i) {
  return i < a.n_dst ? a.dst_nodes[i] : a.src[i - a.n_dst];
}

__global__ void __launch_bounds__(256) block_insert_kernel(BlockArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    const int64_t v = value_at(a, i);
    if (v < 0) {  // not a node id (and -1 is the empty key)
      atomicOr((unsigned long long*)(a.counts + 1), 4ull);
      a.slot_of[i] = -1;
      continue;
    }
    uint32_t slot = hash_slot(v, a.mask);
    while (true) {
      const unsigned long long prev = atomicCAS(a.keys + slot, kEmpty, (unsigned long long)v);
      if (prev == kEmpty || prev == (unsigned long long)v) break;
      slot = (slot + 1) & a.mask;
    }
    atomicMin(a.first + slot, (int32_t)i);
    a.slot_of[i] = (int32_t)slot;
  }
}

__global__ void __launch_bounds__(256) block_flag_kernel(BlockArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= a.n; i += stride) {
    int f = 0;
    if (i < a.n && a.slot_of[i] >= 0) {
      f = a.first[a.slot_of[i]] == (int32_t)i;
      if (!f && i < a.n_dst) atomicOr((unsigned long long*)(a.counts + 1), 1ull);  // duplicate dst
    }
    a.rank[i] = f;
  }
}

__global__ void __launch_bounds__(256) block_emit_kernel(BlockArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid == 0) a.counts[0] = a.rank[a.n];
  for (int64_t i = tid; i < a.n; i += stride) {
    const int32_t slot = a.slot_of[i];
    const int32_t f = slot < 0 ? 0 : a.first[slot];
    const int32_t id = a.rank[f];
    if (f == (int32_t)i) a.src_nid[id] = value_at(a, i);
    if (i >= a.n_dst) a.src_local[i - a.n_dst] = id;
  }
  if (a.dst == nullptr) return;
  for (int64_t e = tid; e < a.n_edges; e += stride) {
    const int64_t v = a.dst[e];
    int64_t id = -1;
    if (v >= 0) {
      uint32_t slot = hash_slot(v, a.mask);
      while (true) {
        const unsigned long long k = a.keys[slot];
        if (k == kEmpty) break;
        if (k == (unsigned long long)v) {
          const int32_t f = a.first[slot];
          if (f < a.n_dst) id = a.rank[f];
          break;
        }
        slot = (slot + 1) & a.mask;
      }
    }
    if (id < 0) atomicOr((unsigned long long*)(a.counts + 1), 2ull);  // not a destination node
    a.dst_local[e] = id < 0 ? 0 : id;
  }
}

}  // namespace

// counts (device, 2 x int64): [0] = number of block source nodes, [1] = status bits
// (1 duplicate destination node, 2 edge destination outside dst_nodes, 4 negative id).
void compact_block(const int64_t* dst_nodes, int64_t n_dst, const int64_t* src, const int64_t* dst,
                   int64_t n_edges, int64_t* src_nid, int64_t* src_local, int64_t* dst_local,
                   int64_t* counts, cudaStream_t s) {
  const int64_t n = n_dst + n_edges;
  B200_CHECK(n < (int64_t)1 << 30, "CompactBlock: more than 2^30 ids");
  B200_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(int64_t), s));
  if (n == 0) return;
  uint64_t cap = 1024;
  while (cap < (uint64_t)n * 2) cap <<= 1;
  Scratch keys(cap * 8, s), first(cap * 4, s), slot_of((size_t)n * 4, s), rank((size_t)(n + 1) * 4, s);
  B200_CUDA(cudaMemsetAsync(keys.ptr, 0xFF, cap * 8, s));
  B200_CUDA(cudaMemsetAsync(first.ptr, 0x7F, cap * 4, s));
  BlockArgs a{dst_nodes, src, dst, n_dst, n_edges, n, keys.as<unsigned long long>(),
              first.as<int32_t>(), slot_of.as<int32_t>(), rank.as<int32_t>(), (uint32_t)(cap - 1),
              src_nid, src_local, dst_local, counts};
  const int blocks = (int)std::min<int64_t>((n + 256) / 256, 148 * 8);
  block_insert_kernel<<<blocks, 256, 0, s>>>(a);
  block_flag_kernel<<<blocks, 256, 0, s>>>(a);
  exclusive_scan_i32(a.rank, n + 1, s);
  block_emit_kernel<<<blocks, 256, 0, s>>>(a);
  note_kernel("compact_block", 3);
  B200_CUDA(cudaGetLastError());
}

}  // namespace b200

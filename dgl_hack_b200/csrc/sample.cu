// Uniform neighbour sampling over the in-CSR: the device form of
// dgl.sampling.sample_neighbors(g, seeds, fanout, replace) and dgl.in_subgraph(g, seeds) that the
// fork's mini-batch trainers call on the host (examples/pytorch/graphsage/
// train_sampling_multi_gpu.py:22-38, python/dgl/sampling/neighbor.py, src/graph/sampling/).
//
// The caller sizes the output from the in-degrees (offsets[i] = first output slot of seed i):
//   fanout < 0           : every in-edge of the seed, in in-CSR (edge id) order   -> in_subgraph
//   replace              : `fanout` independent uniform picks (none for in-degree 0)
//   !replace, deg<=fanout: every in-edge once
//   !replace, deg> fanout: `fanout` distinct in-edges (Floyd's algorithm; selection sampling for
//                          fanouts above the per-thread table)
// Randomness is a counter-based hash of (rng_seed, seed index, draw), so a call is reproducible
// and independent of the launch shape. Outputs are global ids (int64): src, dst = seed, edge id.
#include <algorithm>

#include "common.cuh"

namespace b200 {
namespace {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ uint32_t draw(uint64_t key, uint64_t i, uint64_t k, uint32_t n) {
  const uint64_t r = mix64(mix64(key ^ (i * 0xD1B54A32D192ED03ull)) + k);
  return (uint32_t)(((r >> 32) * (uint64_t)n) >> 32);  // unbiased enough for n << 2^32
}

struct SampleArgs {
  const int32_t *indptr, *indices, *eids;
  const int64_t *seeds, *offsets;
  int64_t *out_src, *out_dst, *out_eid, *out_slot;  // out_slot (optional): index of the seed
  int64_t n_seeds;
  int fanout, replace;
  uint64_t key;
};

constexpr int kTable = 64;  // distinct picks kept per thread (Floyd)

// with replacement / copy-all: one thread per output slot of a seed, a warp per seed
__global__ void __launch_bounds__(256) sample_simple_kernel(SampleArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < a.n_seeds; i += warps) {
    const int64_t v = a.seeds[i];
    const int beg = __ldg(a.indptr + v), deg = __ldg(a.indptr + v + 1) - beg;
    const int64_t o = a.offsets[i];
    const int cnt = (int)(a.offsets[i + 1] - o);
    const bool all = cnt == deg && !(a.replace && a.fanout >= 0);
    for (int k = lane; k < cnt; k += 32) {
      const int p = beg + (all ? k : (int)draw(a.key, (uint64_t)i, (uint64_t)k, (uint32_t)deg));
      a.out_src[o + k] = __ldg(a.indices + p);
      a.out_dst[o + k] = v;
      a.out_eid[o + k] = __ldg(a.eids + p);
      if (a.out_slot) a.out_slot[o + k] = i;
    }
  }
}

// per-seed output counts (int32, scanned in place afterwards) + range check of the seeds; the
// 64-bit total guards the int32 scan (repeated seeds with fanout < 0 can exceed 2^31 edges)
__global__ void __launch_bounds__(256) sample_count_kernel(const int32_t* __restrict__ indptr,
                                                           const int64_t* __restrict__ seeds,
                                                           int64_t n_seeds, int64_t n_rows,
                                                           int fanout, int replace,
                                                           int32_t* __restrict__ cnt,
                                                           unsigned long long* __restrict__ flags) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  unsigned long long mine = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n_seeds; i += stride) {
    int c = 0;
    if (i < n_seeds) {
      const int64_t v = seeds[i];
      if (v < 0 || v >= n_rows) {
        atomicOr(flags, 1ull);
      } else {
        const int deg = __ldg(indptr + v + 1) - __ldg(indptr + v);
        c = fanout < 0 ? deg : (replace ? (deg > 0 ? fanout : 0) : min(deg, fanout));
      }
    }
    cnt[i] = c;
    mine += (unsigned long long)c;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(flags + 1, mine);
}

// flags[0] = 1 if a seed is out of range, flags[1] = total count in 64 bits
__global__ void __launch_bounds__(256) sample_offsets_kernel(const int32_t* __restrict__ scanned,
                                                             const unsigned long long* __restrict__ flags,
                                                             int64_t n_seeds,
                                                             int64_t* __restrict__ offsets) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n_seeds + 1; i += stride)
    offsets[i] = i <= n_seeds ? (int64_t)scanned[i]
                              : (int64_t)(flags[0] | (flags[1] > 0x7fffffffull ? 2ull : 0ull));
}

// without replacement, in-degree above the fanout: one thread per seed
__global__ void __launch_bounds__(128) sample_distinct_kernel(SampleArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n_seeds; i += stride) {
    const int64_t v = a.seeds[i];
    const int beg = __ldg(a.indptr + v), deg = __ldg(a.indptr + v + 1) - beg;
    if (deg <= a.fanout) continue;  // copied by sample_simple_kernel
    const int64_t o = a.offsets[i];
    const int f = a.fanout;
    if (f <= kTable) {
      int pick[kTable];
      for (int k = 0; k < f; ++k) {  // Floyd: j runs over deg - f .. deg - 1
        const int j = deg - f + k;
        int t = (int)draw(a.key, (uint64_t)i, (uint64_t)k, (uint32_t)(j + 1));
        for (int q = 0; q < k; ++q)
          if (pick[q] == t) { t = j; break; }
        pick[k] = t;
      }
      for (int k = 0; k < f; ++k) {
        const int p = beg + pick[k];
        a.out_src[o + k] = __ldg(a.indices + p);
        a.out_dst[o + k] = v;
        a.out_eid[o + k] = __ldg(a.eids + p);
      }
    } else {  // selection sampling: position q is taken with probability needed / remaining
      int need = f, k = 0;
      for (int q = 0; q < deg && need > 0; ++q) {
        if (draw(a.key, (uint64_t)i, (uint64_t)q, (uint32_t)(deg - q)) < (uint32_t)need) {
          const int p = beg + q;
          a.out_src[o + k] = __ldg(a.indices + p);
          a.out_dst[o + k] = v;
          a.out_eid[o + k] = __ldg(a.eids + p);
          ++k, --need;
        }
      }
    }
  }
}

}  // namespace

// offsets[0 .. n_seeds] = exclusive prefix sum of the per-seed output counts (so offsets[n_seeds]
// is the number of sampled edges), offsets[n_seeds + 1] = status bits: 1 a seed is out of range,
// 2 more than 2^31 - 1 sampled edges.
void sample_counts(DGLB200Graph* g, const int64_t* seeds, int64_t n_seeds, int fanout, bool replace,
                   int64_t* offsets, cudaStream_t s) {
  Scratch cnt((size_t)(n_seeds + 1) * 4, s), flags(16, s);
  B200_CUDA(cudaMemsetAsync(flags.ptr, 0, 16, s));
  const int blocks = (int)std::min<int64_t>((n_seeds + 2 + 255) / 256, 148 * 8);
  sample_count_kernel<<<blocks, 256, 0, s>>>(g->in.indptr, seeds, n_seeds, g->num_dst, fanout,
                                             replace ? 1 : 0, cnt.as<int32_t>(),
                                             flags.as<unsigned long long>());
  exclusive_scan_i32(cnt.as<int32_t>(), n_seeds + 1, s);
  sample_offsets_kernel<<<blocks, 256, 0, s>>>(cnt.as<int32_t>(), flags.as<unsigned long long>(),
                                               n_seeds, offsets);
  note_kernel("sample_neighbors", 2);
  B200_CUDA(cudaGetLastError());
}

void sample_neighbors(DGLB200Graph* g, const int64_t* seeds, int64_t n_seeds, int fanout,
                      bool replace, uint64_t rng_seed, const int64_t* offsets, int64_t* out_src,
                      int64_t* out_dst, int64_t* out_eid, int64_t* out_slot, cudaStream_t s) {
  if (n_seeds == 0) return;
  const Csr& in = in_plain(g, s);
  SampleArgs a{in.indptr, in.indices, in.eids, seeds, offsets, out_src, out_dst, out_eid,
               out_slot, n_seeds, fanout, replace ? 1 : 0, mix64(rng_seed)};
  const int blocks = (int)std::min<int64_t>((n_seeds + 7) / 8, 148 * 16);
  if (replace || fanout < 0) {
    sample_simple_kernel<<<blocks, 256, 0, s>>>(a);
    note_kernel("sample_neighbors", 1);
  } else {
    // rows with deg <= fanout are copied whole; the others are overwritten by distinct picks
    sample_simple_kernel<<<blocks, 256, 0, s>>>(a);
    sample_distinct_kernel<<<(int)std::min<int64_t>((n_seeds + 127) / 128, 148 * 16), 128, 0, s>>>(a);
    note_kernel("sample_neighbors", 2);
  }
  B200_CUDA(cudaGetLastError());
}

}  // namespace b200

// SpMM gather-reduce with TMA bulk-copy staging (the wide-row / high-degree path).
//
// Same contract as spmm_kernel (spmm_kernels.cuh): one warp reduces one CSR row sequentially in
// CSR order with its accumulators in registers. The difference is how neighbour rows reach the
// SM: instead of per-lane LDGs whose in-flight data lives in registers, each warp runs its own
// asynchronous pipeline -- lane 0 issues one `cp.async.bulk` (TMA, SASS UBLKCP) per neighbour row
// into a per-warp ring of shared-memory slots, completion is tracked by one mbarrier per slot
// (expect_tx / complete_tx), and the warp consumes slot i while slots i+1 .. i+S-1 are in flight.
// In-flight bytes per SM are bounded by shared memory (up to ~200 KB) rather than by registers,
// which is what a latency-bound random-row gather needs on B200 (HBM3e latency x bandwidth ~ 6 MB
// chip-wide). No inter-warp synchronisation exists in the kernel.
//
// Bulk copies need 16-byte aligned global/shared addresses and sizes: a row that starts at
// `off` = address mod 16 is fetched from address - off with its size rounded up to 16, and read
// back from the slot at byte offset off. The only row for which that could touch bytes past the
// tensor (the last row, when its end is not 16-byte aligned) is staged with ordinary loads.
#pragma once
#include "common.cuh"
#include "spmm_kernels.cuh"

namespace b200 {

struct TmaArgs {
  const int32_t* indptr;
  const int32_t* a_pos_ids;   // per CSR position: column (kFromCol) or edge id (kFromEid)
  const int32_t* a_node_map;  // optional map applied to a_pos_ids values (node maps)
  int n_seg;  // work items [0, n_seg) are heavy-row segments, see SpmmArgs
  int seg_len;
  const int32_t *seg_row, *seg_beg, *seg_end;
  const int32_t* sched;  // order of the row work items (NULL = natural)
  float* partial;
  int n_rows;
  int D;       // floats per row
  int nvec;    // D / VEC
  const float* A;
  int64_t a_rows;  // rows of A (to recognise the last row)
  const float* W;             // optional scalar-per-edge operand ([.,1])
  const int32_t* w_pos_ids;   // per position row id of W
  int w_is_lhs;
  int op, red, mean;
  float* out;
  const int32_t* out_map;
  int stages;      // ring depth S (2..16)
  int slot_bytes;  // bytes per slot (multiple of 16)
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!ok);
}
// global -> shared bulk copy, completion signalled on an mbarrier (TMA 1-D, UBLKCP)
__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst, const void* src, uint32_t bytes,
                                              uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}

template <int VEC>
__device__ __forceinline__ Vec<VEC> lds_vec(const unsigned char* p);
template <>
__device__ __forceinline__ Vec<1> lds_vec<1>(const unsigned char* p) {
  Vec<1> r;
  r.v[0] = *reinterpret_cast<const float*>(p);
  return r;
}
template <>
__device__ __forceinline__ Vec<2> lds_vec<2>(const unsigned char* p) {
  const float2 t = *reinterpret_cast<const float2*>(p);
  Vec<2> r;
  r.v[0] = t.x;
  r.v[1] = t.y;
  return r;
}
template <>
__device__ __forceinline__ Vec<4> lds_vec<4>(const unsigned char* p) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  Vec<4> r;
  r.v[0] = t.x;
  r.v[1] = t.y;
  r.v[2] = t.z;
  r.v[3] = t.w;
  return r;
}

constexpr int kTmaMaxWarps = 16;  // warps per CTA (runtime: blockDim.x / 32), each with a private ring

template <int VEC, int CH, bool SUM, bool HAS_W>
__global__ void __launch_bounds__(kTmaMaxWarps * 32) spmm_tma_kernel(const TmaArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int S = a.stages;
  const int kTmaWarps = blockDim.x >> 5;
  unsigned char* ring = smem + (size_t)wid * S * a.slot_bytes;
  const uint32_t ring_u32 = smem_u32(ring);
  const uint32_t bars_u32 =
      smem_u32(smem + (size_t)kTmaWarps * S * a.slot_bytes) + (uint32_t)(wid * S * 8);
  if (lane == 0) {
    for (int s = 0; s < S; ++s) mbar_init(bars_u32 + s * 8, 1);
    mbar_fence_init();
  }
  __syncwarp();

  const int64_t item = (int64_t)blockIdx.x * kTmaWarps + wid;
  if (item >= (int64_t)a.n_seg + a.n_rows) return;  // no CTA-wide synchronisation below
  const bool is_seg = item < a.n_seg;
  int r, beg, deg;
  if (is_seg) {
    r = __ldg(a.seg_row + item);
    beg = __ldg(a.seg_beg + item);
    deg = __ldg(a.seg_end + item) - beg;
  } else {
    r = (int)(item - a.n_seg);
    if (a.sched) r = __ldg(a.sched + r);
    beg = __ldg(a.indptr + r);
    deg = __ldg(a.indptr + r + 1) - beg;
    if (a.n_seg > 0 && deg > a.seg_len) return;  // heavy row: served by its segments
  }
  const uint32_t row_bytes = (uint32_t)a.D * 4u;

  int col[CH];
  bool cok[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int vc = c * 32 + lane;
    cok[c] = vc < a.nvec;
    col[c] = vc * VEC * 4;  // byte offset inside the row
  }
  const float ident = SUM ? 0.f : red_identity(a.red);
  Vec<VEC> acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[c].v[e] = ident;

  // ---- neighbour ids (and scalar weights) of the current and the next block of 32 positions
  auto load_ids = [&](int block) -> int {
    const int j = block * 32 + lane;
    if (j >= deg) return 0;
    int v = __ldg(a.a_pos_ids + beg + j);
    if (a.a_node_map) v = __ldg(a.a_node_map + v);
    return v;
  };
  auto load_w = [&](int block) -> float {
    const int j = block * 32 + lane;
    if (!HAS_W || j >= deg) return 0.f;
    return __ldg(a.W + __ldg(a.w_pos_ids + beg + j));
  };
  int blk = 0;
  int id_cur = load_ids(0), id_next = load_ids(1);
  float w_cur = load_w(0), w_next = load_w(1);

  // ---- stage neighbour j into slot j % S
  auto issue = [&](int j) {
    const int sel = ((j >> 5) == blk) ? id_cur : id_next;  // j's block is blk or blk + 1
    const int ia = __shfl_sync(0xffffffffu, sel, j & 31);
    const int slot = j % S;
    const float* grow = a.A + (int64_t)ia * a.D;
    const uint32_t off = (uint32_t)(reinterpret_cast<uintptr_t>(grow) & 15u);
    const uint32_t bytes = (off + row_bytes + 15u) & ~15u;
    const uint32_t bar = bars_u32 + slot * 8;
    const bool past_end = (ia == a.a_rows - 1) && (((off + row_bytes) & 15u) != 0u);
    if (!past_end) {
      if (lane == 0) {
        mbar_arrive_expect_tx(bar, bytes);
        bulk_copy_g2s(ring_u32 + slot * a.slot_bytes,
                      reinterpret_cast<const unsigned char*>(grow) - off, bytes, bar);
      }
    } else {
      float* d = reinterpret_cast<float*>(ring + (size_t)slot * a.slot_bytes + off);
      for (int k = lane; k < a.D; k += 32) d[k] = __ldg(grow + k);
      __syncwarp();
      if (lane == 0) mbar_arrive(bar);
    }
  };

  const int pre = min(S - 1, deg);
  for (int j = 0; j < pre; ++j) issue(j);

  for (int i = 0; i < deg; ++i) {
    if ((i & 31) == 0 && i > 0) {
      ++blk;
      id_cur = id_next;
      w_cur = w_next;
      id_next = load_ids(blk + 1);
      w_next = load_w(blk + 1);
    }
    __syncwarp();  // every lane has finished reading the slot that is about to be refilled
    if (i + S - 1 < deg) issue(i + S - 1);
    const int slot = i % S;
    const int ia = __shfl_sync(0xffffffffu, id_cur, i & 31);
    const uint32_t off =
        (uint32_t)(reinterpret_cast<uintptr_t>(a.A + (int64_t)ia * a.D) & 15u);
    float wv = 0.f;
    if (HAS_W) wv = __shfl_sync(0xffffffffu, w_cur, i & 31);
    mbar_wait(bars_u32 + slot * 8, (uint32_t)((i / S) & 1));
    const unsigned char* sp = ring + (size_t)slot * a.slot_bytes + off;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      if (cok[c]) {
        const Vec<VEC> x = lds_vec<VEC>(sp + col[c]);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
          float val = x.v[e];
          if (HAS_W) val = spmm_value(a.op, a.w_is_lhs, val, wv);
          else if (a.op == kOpNegA) val = __fmul_rn(val, -1.f);
          acc[c].v[e] = SUM ? __fadd_rn(acc[c].v[e], val) : red_call(a.red, acc[c].v[e], val);
        }
      }
    }
  }

  float* op;
  if (is_seg) {
    op = a.partial + item * a.D;
  } else {
    const int64_t orow = a.out_map ? __ldg(a.out_map + r) : r;
    op = a.out + orow * a.D;
  }
  const float dn = (float)(deg > 1 ? deg : 1);
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    if (cok[c]) {
      if (a.mean && !is_seg) {
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[c].v[e] = __fdiv_rn(acc[c].v[e], dn);
      }
      st_vec<VEC>(op + col[c] / 4, acc[c]);
    }
  }
}

template <int VEC, int CH>
void launch_tma_shape(const TmaArgs& a, bool sum, bool has_w, int blocks, int warps, size_t smem,
                      cudaStream_t s) {
#define B200_TMA_LAUNCH(SUMV, WV)                                                          \
  do {                                                                                     \
    auto kfn = spmm_tma_kernel<VEC, CH, SUMV, WV>;                                         \
    cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);     \
    kfn<<<blocks, warps * 32, smem, s>>>(a);                                           \
  } while (0)
  if (sum) {
    if (has_w) B200_TMA_LAUNCH(true, true);
    else B200_TMA_LAUNCH(true, false);
  } else {
    if (has_w) B200_TMA_LAUNCH(false, true);
    else B200_TMA_LAUNCH(false, false);
  }
#undef B200_TMA_LAUNCH
}

// ch must be one of {1,2,3,4,5,6,8,10,12,16}
template <int VEC>
void launch_tma_vec(const TmaArgs& a, int ch, bool sum, bool has_w, int blocks, int warps,
                    size_t smem, cudaStream_t s) {
  switch (ch) {
    case 1: launch_tma_shape<VEC, 1>(a, sum, has_w, blocks, warps, smem, s); break;
    case 2: launch_tma_shape<VEC, 2>(a, sum, has_w, blocks, warps, smem, s); break;
    case 3: launch_tma_shape<VEC, 3>(a, sum, has_w, blocks, warps, smem, s); break;
    case 4: launch_tma_shape<VEC, 4>(a, sum, has_w, blocks, warps, smem, s); break;
    case 5: launch_tma_shape<VEC, 5>(a, sum, has_w, blocks, warps, smem, s); break;
    case 6: launch_tma_shape<VEC, 6>(a, sum, has_w, blocks, warps, smem, s); break;
    case 8: launch_tma_shape<VEC, 8>(a, sum, has_w, blocks, warps, smem, s); break;
    case 10: launch_tma_shape<VEC, 10>(a, sum, has_w, blocks, warps, smem, s); break;
    case 12: launch_tma_shape<VEC, 12>(a, sum, has_w, blocks, warps, smem, s); break;
    default: launch_tma_shape<VEC, 16>(a, sum, has_w, blocks, warps, smem, s); break;
  }
}

}  // namespace b200

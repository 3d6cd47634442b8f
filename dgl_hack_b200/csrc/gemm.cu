// Dense projection on the 5th-generation tensor cores at fp32 accuracy:
//
//     C[M, N] = A[M, K] . Bt[N, K]^T  (+ bias[N]) (ReLU)          fp32 in, fp32 out
//
// the X.W products either side of the aggregation in GraphConv / SAGEConv / GATConv
// (python/dgl/nn/pytorch/conv/graphconv.py:144-176, sageconv.py:150-160, gatconv.py:125-144), which
// the reference leaves to cuBLAS' fp32 SIMT GEMM (46 % of the products-shaped GCN epoch).
//
// Arithmetic: error-compensated 3xTF32. Every operand is split into hi = its TF32 truncation
// (low 13 mantissa bits cleared) and lo = x - hi (exact in fp32); the kernel accumulates
// a_hi.b_hi + a_hi.b_lo + a_lo.b_hi in fp32 (TMEM). The dropped a_lo.b_lo term and the TF32
// rounding of the lo parts are each <= 2^-20 of |a||b|, i.e. ~1e-6 relative to sum|a||b| -- inside
// the 1e-5 tolerance plain TF32 (2^-11 per operand) breaks.
//
// Structure (one CTA per SM, persistent over 128-row tiles of A, whole N per tile):
//   warp 0     TMA producer: A tile [128 x 32] fp32 and the pre-split Bt_hi / Bt_lo tiles
//              [Npad x 32] into a ring of shared-memory stages (SWIZZLE_128B, K-major)
//   warps 4-7  splitter: a_hi in place, a_lo into the stage's second A buffer (element-wise, so
//              the swizzled layout is preserved), fence.proxy.async, arrive
//   warp 1     MMA issuer: one elected thread, 3 x tcgen05.mma.kind::tf32 per 8-wide K step into
//              one of two TMEM accumulators [128 lanes x Npad columns]; tcgen05.commit frees the
//              stage / publishes the accumulator
//   warps 8-11 epilogue: tcgen05.ld (32x32b), bias / ReLU, direct 64-byte row-chunk stores
//   warp 2     TMEM allocation
// The weights are tiny and shared by all tiles: they are split once per call by a prologue kernel
// (also transposing when the caller's matrix is [K, N]).
#include <cuda.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace b200 {
namespace {

constexpr int kBM = 128;          // rows of A per tile = TMEM lanes
constexpr int kBK = 32;           // floats per K block = one 128-byte swizzle row
constexpr int kUmmaK = 8;         // K of one tcgen05.mma.kind::tf32
constexpr int kThreads = 384;
constexpr int kATileBytes = kBM * kBK * 4;  // 16 KB

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// K-major SWIZZLE_128B operand descriptor (cute/arch/mma_sm100_desc.hpp SmemDescriptor;
// canonical layout ((8,n),2):((8,SBO),1) in 16-byte units): start address >> 4, LBO = 1,
// SBO = 1024 B between 8-row groups, version 1 (sm_100), layout type 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t umma_desc_k_sw128(const void* tile) {
  return (uint64_t)((smem_u32(tile) & 0x3FFFF) >> 4) | (1ull << 16) | (64ull << 32) |
         (1ull << 46) | (2ull << 61);
}

// D[tmem] (+)= A[smem] . B[smem]^T, 128 x N x 8, TF32 inputs, fp32 accumulate
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}

struct GemmArgs {
  int M, N, K, Npad;
  int num_m_tiles, num_k_blocks, stages;
  uint32_t stage_bytes, b_tile_bytes, tmem_cols;
  const float* bias;  // [N] or NULL
  int relu;
  float* C;           // [M, ldc]
  int64_t ldc;
  // grouped mode (NULL = one group, tile t covers rows [128 t, 128 t + 128)): tile t covers rows
  // [tile_row0[t], tile_row0[t] + tile_rows[t]) of A / C, all of group tile_group[t], and is
  // contracted with that group's weight (rows [2 Npad g, 2 Npad (g + 1)) of the split-weight map)
  const int32_t *tile_row0, *tile_rows, *tile_group;
};

// Shared-memory stage s: A (raw -> hi) | A_lo | Bt_hi | Bt_lo, every part 1024-byte aligned.
__global__ void __launch_bounds__(kThreads, 1)
gemm_tf32x3_kernel(const __grid_constant__ CUtensorMap map_a,
                   const __grid_constant__ CUtensorMap map_b, const GemmArgs g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic shared memory is only guaranteed 16-byte alignment: align by hand
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int S = g.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * g.stage_bytes);
  uint64_t* full = bars;             // [S] TMA landed
  uint64_t* split = bars + S;        // [S] a_hi / a_lo written
  uint64_t* empty = bars + 2 * S;    // [S] MMAs that read the stage have completed
  uint64_t* acc_full = bars + 3 * S;       // [2] accumulator complete
  uint64_t* acc_empty = bars + 3 * S + 2;  // [2] accumulator drained by the epilogue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * S + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full + s, 1);
      mbar_init(split + s, 128);
      mbar_init(empty + s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(acc_full + a, 1);
      mbar_init(acc_empty + a, 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  } else if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_slot)),
                 "r"(g.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  auto stage_ptr = [&](int s) { return smem + (size_t)s * g.stage_bytes; };
  const uint32_t tx_bytes = kATileBytes + 2 * g.b_tile_bytes;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int s = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < g.num_m_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < g.num_k_blocks; ++kb) {
          mbar_wait(empty + s, phase ^ 1);
          uint8_t* st = stage_ptr(s);
          mbar_arrive_expect_tx(full + s, tx_bytes);
          const int row0 = g.tile_row0 ? __ldg(g.tile_row0 + tile) : tile * kBM;
          const int brow = g.tile_group ? __ldg(g.tile_group + tile) * 2 * g.Npad : 0;
          tma_load_2d(&map_a, full + s, st, kb * kBK, row0);
          tma_load_2d(&map_b, full + s, st + 2 * kATileBytes, kb * kBK, brow);
          tma_load_2d(&map_b, full + s, st + 2 * kATileBytes + g.b_tile_bytes, kb * kBK,
                      brow + g.Npad);
          if (++s == S) { s = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // instruction descriptor (InstrDescriptor): D = F32, A = B = TF32, both K-major, N, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(g.Npad >> 3) << 17) |
                           ((uint32_t)(kBM >> 4) << 24);
    int s = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < g.num_m_tiles; tile += gridDim.x) {
      mbar_wait(acc_empty + acc, acc_phase ^ 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * g.Npad);
      for (int kb = 0; kb < g.num_k_blocks; ++kb) {
        mbar_wait(split + s, phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (lane == 0) {
          uint8_t* st = stage_ptr(s);
          const uint64_t a_hi = umma_desc_k_sw128(st), a_lo = umma_desc_k_sw128(st + kATileBytes);
          const uint64_t b_hi = umma_desc_k_sw128(st + 2 * kATileBytes);
          const uint64_t b_lo = umma_desc_k_sw128(st + 2 * kATileBytes + g.b_tile_bytes);
#pragma unroll
          for (int k = 0; k < kBK / kUmmaK; ++k) {
            const uint64_t adv = (uint64_t)(k * kUmmaK * 4 >> 4);  // 32 bytes per K step
            // small terms first, the leading product last
            umma_tf32(tmem_d, a_lo + adv, b_hi + adv, idesc, (kb | k) != 0);
            umma_tf32(tmem_d, a_hi + adv, b_lo + adv, idesc, 1);
            umma_tf32(tmem_d, a_hi + adv, b_hi + adv, idesc, 1);
          }
          umma_commit(empty + s);                                   // stage reusable when done
          if (kb == g.num_k_blocks - 1) umma_commit(acc_full + acc);  // accumulator ready
        }
        __syncwarp();
        if (++s == S) { s = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===================== splitter =====================
    const int t = threadIdx.x - 128;  // 0..127
    int s = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < g.num_m_tiles; tile += gridDim.x) {
      for (int kb = 0; kb < g.num_k_blocks; ++kb) {
        mbar_wait(full + s, phase);
        uint4* hi = reinterpret_cast<uint4*>(stage_ptr(s));
        uint4* lo = reinterpret_cast<uint4*>(stage_ptr(s) + kATileBytes);
#pragma unroll
        for (int i = 0; i < kATileBytes / 16 / 128; ++i) {  // 8 x 16 bytes per thread
          const int idx = i * 128 + t;
          uint4 v = hi[idx], h, l;
          h.x = v.x & 0xffffe000u; h.y = v.y & 0xffffe000u;
          h.z = v.z & 0xffffe000u; h.w = v.w & 0xffffe000u;
          l.x = __float_as_uint(__fsub_rn(__uint_as_float(v.x), __uint_as_float(h.x)));
          l.y = __float_as_uint(__fsub_rn(__uint_as_float(v.y), __uint_as_float(h.y)));
          l.z = __float_as_uint(__fsub_rn(__uint_as_float(v.z), __uint_as_float(h.z)));
          l.w = __float_as_uint(__fsub_rn(__uint_as_float(v.w), __uint_as_float(h.w)));
          hi[idx] = h;
          lo[idx] = l;
        }
        // generic-proxy writes -> visible to the tensor core's async-proxy reads
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_arrive(split + s);
        if (++s == S) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 8) {
    // ===================== epilogue =====================
    const int q = warp & 3;                 // TMEM lane quarter this warp may read
    const int row_in_tile = q * 32 + lane;  // accumulator row = TMEM lane
    const bool vec_ok = (g.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(g.C) & 15) == 0);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < g.num_m_tiles; tile += gridDim.x) {
      mbar_wait(acc_full + acc, acc_phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int64_t row = (g.tile_row0 ? (int64_t)__ldg(g.tile_row0 + tile) : (int64_t)tile * kBM) +
                          row_in_tile;
      const bool row_on = g.tile_rows ? row_in_tile < __ldg(g.tile_rows + tile) : row < g.M;
      float* crow = g.C + row * g.ldc;
      for (int c0 = 0; c0 < g.Npad; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * g.Npad + c0), v);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (row_on) {
          float f[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            f[j] = __uint_as_float(v[j]);
            if (g.bias && c0 + j < g.N) f[j] = __fadd_rn(f[j], __ldg(g.bias + c0 + j));
            if (g.relu) f[j] = fmaxf(f[j], 0.f);
          }
          if (vec_ok && c0 + 16 <= g.N) {
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              *reinterpret_cast<float4*>(crow + c0 + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c0 + j < g.N) crow[c0 + j] = f[j];
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      mbar_arrive(acc_empty + acc);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 2) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(g.tmem_cols)
                 : "memory");
  }
}


// ---------------------------------------------------------------------------------------------
// Weight gradient  C[Kf, N] = A[M, Kf]^T . G[M, N]  (A = layer input, G = grad of the output):
// the contraction runs over the LONG dimension M, so both operands are MN-major for the tensor
// core (their contiguous dimension is the MMA's M / N, not K). Shared-memory image of a stage
// (kTnRows = 16 rows of A and G): one TMA box [16 rows x 32 floats] per 32-column block,
// 128-byte swizzle with 32-byte atoms, blocks 2 KB apart (LBO = 2048 B: next 32-column block;
// SBO = 512 B: next 4 rows; see umma_desc_mn_sw128). Each CTA owns a
// contiguous chunk of rows, accumulates a full [Kf, N] partial in TMEM (Kf/128 accumulators of
// N columns), writes it to scratch; gemm_tn_fold_kernel adds the partials in CTA order
// (deterministic). Both operands are big, so both are split hi / lo in the kernel.
constexpr int kTnRows = 16;
constexpr int kTnBoxBytes = kTnRows * 128;  // 2 KB

// MN-major operands of 32-bit element types have exactly one legal shared-memory layout
// (cutlass sm100_common.inl: "for mn-major tf32 operands, SW128_32B is the only available smem
// layout"): SWIZZLE_128B_BASE32B (layout type 1, Swizzle<2,5,2>: 32-byte atoms permuted inside a
// 128-byte row by row % 4), canonical form ((8,n),(4,k)):((1,LBO),(8,SBO)) in 16-byte units --
// K groups of FOUR rows (SBO = 512 B), MN blocks of 32 floats (LBO = one TMA box = 2 KB). TMA
// writes the same pattern with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(const void* tile) {
  return (uint64_t)((smem_u32(tile) & 0x3FFFF) >> 4) | ((uint64_t)(kTnBoxBytes >> 4) << 16) |
         (32ull << 32) | (1ull << 46) | (1ull << 61);
}

struct GemmTnArgs {
  int M, Kf, N, Npad;       // Npad = N rounded up to 16 (UMMA N)
  int a_blocks, g_blocks;   // 32-column blocks of A (Kf padded to 128) and of G
  int m_tiles;              // Kf / 128 rounded up
  int rows_per_cta, stages;
  uint32_t stage_bytes, a_bytes, g_bytes, tmem_cols;
  float* partial;           // [gridDim.x, m_tiles * 128, Npad]
};

__global__ void __launch_bounds__(kThreads, 1)
gemm_tn_tf32x3_kernel(const __grid_constant__ CUtensorMap map_a,
                      const __grid_constant__ CUtensorMap map_g, const GemmTnArgs g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int S = g.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * g.stage_bytes);
  uint64_t *full = bars, *split = bars + S, *empty = bars + 2 * S, *acc_full = bars + 3 * S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * S + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full + s, 1);
      mbar_init(split + s, 128);
      mbar_init(empty + s, 1);
    }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  } else if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_slot)),
                 "r"(g.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  const int64_t row_beg = (int64_t)blockIdx.x * g.rows_per_cta;
  const int64_t row_end = min(row_beg + (int64_t)g.rows_per_cta, (int64_t)g.M);
  const int n_steps = row_end > row_beg ? (int)((row_end - row_beg + kTnRows - 1) / kTnRows) : 0;
  // stage: A (raw -> hi) | A_lo | G (raw -> hi) | G_lo
  auto stage_ptr = [&](int s) { return smem + (size_t)s * g.stage_bytes; };

  if (warp == 0) {
    if (lane == 0) {
      int s = 0;
      uint32_t phase = 0;
      for (int it = 0; it < n_steps; ++it) {
        mbar_wait(empty + s, phase ^ 1);
        uint8_t* st = stage_ptr(s);
        mbar_arrive_expect_tx(full + s, g.a_bytes + g.g_bytes);
        const int r0 = (int)(row_beg + (int64_t)it * kTnRows);
        // rows past this CTA's chunk belong to the next CTA: they are zeroed by the splitter
        for (int j = 0; j < g.a_blocks; ++j)
          tma_load_2d(&map_a, full + s, st + j * kTnBoxBytes, j * 32, r0);
        for (int j = 0; j < g.g_blocks; ++j)
          tma_load_2d(&map_g, full + s, st + 2 * g.a_bytes + j * kTnBoxBytes, j * 32, r0);
        if (++s == S) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // D = F32, A = B = TF32, BOTH MN-major (bits 15, 16), N = Npad, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) |
                           ((uint32_t)(g.Npad >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);
    int s = 0;
    uint32_t phase = 0;
    for (int it = 0; it < n_steps; ++it) {
      mbar_wait(split + s, phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (lane == 0) {
        uint8_t* st = stage_ptr(s);
        for (int mt = 0; mt < g.m_tiles; ++mt) {
          const uint32_t tmem_d = tmem_base + (uint32_t)(mt * g.Npad);
          const uint8_t* a_hi_p = st + (size_t)mt * 4 * kTnBoxBytes;  // 128 columns = 4 blocks
          const uint64_t a_hi = umma_desc_mn_sw128(a_hi_p);
          const uint64_t a_lo = umma_desc_mn_sw128(a_hi_p + g.a_bytes);
          const uint64_t g_hi = umma_desc_mn_sw128(st + 2 * g.a_bytes);
          const uint64_t g_lo = umma_desc_mn_sw128(st + 2 * g.a_bytes + g.g_bytes);
#pragma unroll
          for (int k = 0; k < kTnRows / kUmmaK; ++k) {
            const uint64_t adv = (uint64_t)(k * 1024 >> 4);  // next 8 rows (two 4-row K groups)
            umma_tf32(tmem_d, a_lo + adv, g_hi + adv, idesc, (it | k) != 0);
            umma_tf32(tmem_d, a_hi + adv, g_lo + adv, idesc, 1);
            umma_tf32(tmem_d, a_hi + adv, g_hi + adv, idesc, 1);
          }
        }
        umma_commit(empty + s);
        if (it == n_steps - 1) umma_commit(acc_full);
      }
      __syncwarp();
      if (++s == S) { s = 0; phase ^= 1; }
    }
  } else if (warp >= 4 && warp < 8) {
    const int t = threadIdx.x - 128;
    int s = 0;
    uint32_t phase = 0;
    for (int it = 0; it < n_steps; ++it) {
      mbar_wait(full + s, phase);
      const int64_t r0 = row_beg + (int64_t)it * kTnRows;
      const int valid = (int)min((int64_t)kTnRows, row_end - r0);  // rows of this CTA's chunk
      for (int part = 0; part < 2; ++part) {
        uint8_t* base = stage_ptr(s) + (part ? 2 * g.a_bytes : 0);
        const uint32_t bytes = part ? g.g_bytes : g.a_bytes;
        uint4* hi = reinterpret_cast<uint4*>(base);
        uint4* lo = reinterpret_cast<uint4*>(base + bytes);
        for (uint32_t idx = t; idx < bytes / 16; idx += 128) {
          // 16-byte chunk idx: box = idx / 128, row inside the box = (idx % 128) / 8
          const int row = (int)((idx & 127) >> 3);
          uint4 v = hi[idx], h, l;
          if (row >= valid) v = make_uint4(0, 0, 0, 0);
          h.x = v.x & 0xffffe000u; h.y = v.y & 0xffffe000u;
          h.z = v.z & 0xffffe000u; h.w = v.w & 0xffffe000u;
          l.x = __float_as_uint(__fsub_rn(__uint_as_float(v.x), __uint_as_float(h.x)));
          l.y = __float_as_uint(__fsub_rn(__uint_as_float(v.y), __uint_as_float(h.y)));
          l.z = __float_as_uint(__fsub_rn(__uint_as_float(v.z), __uint_as_float(h.z)));
          l.w = __float_as_uint(__fsub_rn(__uint_as_float(v.w), __uint_as_float(h.w)));
          hi[idx] = h;
          lo[idx] = l;
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_arrive(split + s);
      if (++s == S) { s = 0; phase ^= 1; }
    }
  } else if (warp >= 8) {
    const int q = warp & 3;
    float* pbase = g.partial + (size_t)blockIdx.x * g.m_tiles * kBM * g.Npad;
    if (n_steps > 0) {
      mbar_wait(acc_full, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    for (int mt = 0; mt < g.m_tiles; ++mt) {
      float* prow = pbase + (size_t)(mt * kBM + q * 32 + lane) * g.Npad;
      for (int c0 = 0; c0 < g.Npad; c0 += 16) {
        uint32_t v[16];
        if (n_steps > 0) {
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * g.Npad + c0), v);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = 0;
        }
#pragma unroll
        for (int j = 0; j < 16; j += 4)
          *reinterpret_cast<uint4*>(prow + c0 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 2) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(g.tmem_cols)
                 : "memory");
  }
}

// C[i, j] = sum over CTAs (in CTA order) of partial[b][i][j]
__global__ void gemm_tn_fold_kernel(const float* __restrict__ partial, int n_parts, int rows_pad,
                                    int Npad, int Kf, int N, float* __restrict__ C) {
  const int64_t total = (int64_t)Kf * N;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / N), cidx = (int)(i - (int64_t)r * N);
    float acc = 0.f;
    for (int b = 0; b < n_parts; ++b)
      acc = __fadd_rn(acc, __ldg(partial + ((size_t)b * rows_pad + r) * Npad + cidx));
    C[i] = acc;
  }
}

// Bsplit[0][n][k] = hi(Bt[n][k]), Bsplit[1][n][k] = Bt[n][k] - hi; rows N..Npad-1 are zero.
// b_is_kn: the caller's matrix is W[K, N] row-major (Y = X.W), read transposed; otherwise it
// is Bt[N, K] row-major (Y = X.Bt^T, e.g. grad_X = grad_Y.W^T with Bt = W).
// (group gi: b + gi * N * K, out + gi * 2 * Npad * Kld)
__global__ void split_weights_kernel(const float* __restrict__ b, int b_is_kn, int N, int K,
                                     int Npad, int Kld, int groups, float* __restrict__ out) {
  const int64_t per = (int64_t)Npad * Kld, total = per * groups;
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < total;
       j += (int64_t)gridDim.x * blockDim.x) {
    const int gi = (int)(j / per);
    const int64_t i = j - (int64_t)gi * per;
    const int n = (int)(i / Kld), k = (int)(i - (int64_t)n * Kld);
    const float* bg = b + (int64_t)gi * N * K;
    float v = 0.f;
    if (n < N && k < K) v = b_is_kn ? __ldg(bg + (int64_t)k * N + n) : __ldg(bg + (int64_t)n * K + k);
    const float h = __uint_as_float(__float_as_uint(v) & 0xffffe000u);
    out[(int64_t)gi * 2 * per + i] = h;
    out[(int64_t)gi * 2 * per + per + i] = __fsub_rn(v, h);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    B200_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    B200_CHECK(p != nullptr && q == cudaDriverEntryPointSuccess,
               "cuTensorMapEncodeTiled is not available from this driver");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// [rows, cols] fp32 row-major (row pitch ld floats) as a 2-D tiled map with a
// [box_rows x 32] box, SWIZZLE_128B; out-of-bounds elements are delivered as zeros.
CUtensorMap make_map(const float* base, int64_t rows, int64_t cols, int64_t ld, int box_rows,
                     CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  CUtensorMap m;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  const cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = encode_tiled()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                                    const_cast<float*>(base), dims, strides, box, estr,
                                    CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
  return m;
}

}  // namespace

bool gemm_tn_tf32x3_supported(int64_t M, int64_t Kf, int64_t N, const float* A, const float* G) {
  return M >= 1 && M < (1ll << 31) && Kf >= 1 && Kf <= 256 && N >= 1 && N <= 256 && Kf % 4 == 0 &&
         N % 4 == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0 &&
         (reinterpret_cast<uintptr_t>(G) & 15) == 0;
}

// C[Kf, N] = A[M, Kf]^T . G[M, N], all compact row-major fp32 (the weight gradient of Y = X.W)
void gemm_tn_tf32x3(const float* A, const float* G, float* C, int64_t M, int64_t Kf, int64_t N,
                    cudaStream_t s) {
  B200_CHECK(gemm_tn_tf32x3_supported(M, Kf, N, A, G),
             "gemm_tn_tf32x3: needs Kf, N <= 256 and divisible by 4, 16-byte aligned operands");
  GemmTnArgs g;
  g.M = (int)M; g.Kf = (int)Kf; g.N = (int)N;
  g.Npad = (int)((N + 15) / 16 * 16);
  g.m_tiles = (int)((Kf + kBM - 1) / kBM);
  g.a_blocks = g.m_tiles * 4;
  g.g_blocks = (g.Npad + 31) / 32;
  g.a_bytes = (uint32_t)g.a_blocks * kTnBoxBytes;
  g.g_bytes = (uint32_t)g.g_blocks * kTnBoxBytes;
  g.stage_bytes = 2 * (g.a_bytes + g.g_bytes);
  g.stages = (int)std::max<uint32_t>(2, std::min<uint32_t>(8, (200u * 1024) / g.stage_bytes));
  g.tmem_cols = 32;
  while (g.tmem_cols < (uint32_t)(g.m_tiles * g.Npad)) g.tmem_cols <<= 1;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = (int)std::min<int64_t>(sms, (M + kTnRows - 1) / kTnRows);
  g.rows_per_cta = (int)(((M + grid - 1) / grid + kTnRows - 1) / kTnRows * kTnRows);
  Scratch partial((size_t)grid * g.m_tiles * kBM * g.Npad * 4, s);
  g.partial = partial.as<float>();
  const CUtensorMap map_a = make_map(A, M, Kf, Kf, kTnRows, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
  const CUtensorMap map_g = make_map(G, M, N, N, kTnRows, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
  const size_t smem = (size_t)g.stages * g.stage_bytes + (3 * g.stages + 2) * 8 + 16 + 1024;
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(gemm_tn_tf32x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         227 * 1024);
  });
  gemm_tn_tf32x3_kernel<<<grid, kThreads, smem, s>>>(map_a, map_g, g);
  B200_CUDA(cudaGetLastError());
  const int64_t tot = Kf * N;
  gemm_tn_fold_kernel<<<(int)std::min<int64_t>((tot + 255) / 256, 592), 256, 0, s>>>(
      partial.as<float>(), grid, g.m_tiles * kBM, g.Npad, (int)Kf, (int)N, C);
  B200_CUDA(cudaGetLastError());
  note_kernel("gemm_tn_tf32x3", 2);
}

bool gemm_tf32x3_supported(int64_t M, int64_t N, int64_t K, const float* A, int64_t lda) {
  return M >= 1 && N >= 1 && N <= 256 && K >= 1 && K <= 65536 && lda % 4 == 0 && lda >= K &&
         (reinterpret_cast<uintptr_t>(A) & 15) == 0 && M < (1ll << 31);
}

// C[M, N] (row pitch ldc) = A[M, K] (row pitch lda) . B (+ bias) (relu); b_is_kn selects
// B = W[K, N] (1) or B = Bt[N, K] (0), both compact row-major. groups > 1: rows
// [group_ptr[g], group_ptr[g + 1]) of A are contracted with B[g] (B = [groups, K, N] or
// [groups, N, K]); group_ptr is a HOST array of groups + 1 row offsets (group_ptr[groups] = M).
void gemm_tf32x3(const float* A, int64_t lda, const float* B, int b_is_kn, const float* bias,
                 int relu, float* C, int64_t ldc, int64_t M, int64_t N, int64_t K, int groups,
                 const int64_t* group_ptr, cudaStream_t s) {
  B200_CHECK(gemm_tf32x3_supported(M, N, K, A, lda),
             "gemm_tf32x3: needs N <= 256, a 16-byte aligned A and a row pitch divisible by 4");
  B200_CHECK(groups >= 1 && (groups == 1 || group_ptr), "gemm_tf32x3: bad group arguments");
  GemmArgs g;
  g.M = (int)M; g.N = (int)N; g.K = (int)K;
  g.Npad = (int)((N + 15) / 16 * 16);
  B200_CHECK((int64_t)groups * 2 * g.Npad < (1ll << 31), "gemm_tf32x3: too many groups");
  g.num_k_blocks = (int)((K + kBK - 1) / kBK);
  g.b_tile_bytes = (uint32_t)g.Npad * kBK * 4;
  g.stage_bytes = 2 * kATileBytes + 2 * g.b_tile_bytes;
  g.stages = (int)std::max<uint32_t>(2, std::min<uint32_t>(6, (200u * 1024) / g.stage_bytes));
  g.tmem_cols = 32;
  while (g.tmem_cols < 2u * g.Npad) g.tmem_cols <<= 1;
  g.bias = bias; g.relu = relu; g.C = C; g.ldc = ldc;
  g.tile_row0 = g.tile_rows = g.tile_group = nullptr;
  g.num_m_tiles = (int)((M + kBM - 1) / kBM);
  Scratch table;
  if (groups > 1) {
    std::vector<int32_t> t0, tn, tg;
    for (int gi = 0; gi < groups; ++gi) {
      B200_CHECK(group_ptr[gi] <= group_ptr[gi + 1] && group_ptr[gi] >= 0 && group_ptr[gi + 1] <= M,
                 "gemm_tf32x3: group_ptr must be non-decreasing row offsets within [0, M]");
      for (int64_t r = group_ptr[gi]; r < group_ptr[gi + 1]; r += kBM) {
        t0.push_back((int32_t)r);
        tn.push_back((int32_t)std::min<int64_t>(kBM, group_ptr[gi + 1] - r));
        tg.push_back(gi);
      }
    }
    g.num_m_tiles = (int)t0.size();
    if (g.num_m_tiles == 0) return;
    const size_t nb = (size_t)g.num_m_tiles * 4;
    table.alloc(3 * nb, s);
    char* base = table.as<char>();
    B200_CUDA(cudaMemcpyAsync(base, t0.data(), nb, cudaMemcpyHostToDevice, s));
    B200_CUDA(cudaMemcpyAsync(base + nb, tn.data(), nb, cudaMemcpyHostToDevice, s));
    B200_CUDA(cudaMemcpyAsync(base + 2 * nb, tg.data(), nb, cudaMemcpyHostToDevice, s));
    B200_CUDA(cudaStreamSynchronize(s));  // the host vectors go out of scope
    g.tile_row0 = reinterpret_cast<const int32_t*>(base);
    g.tile_rows = reinterpret_cast<const int32_t*>(base + nb);
    g.tile_group = reinterpret_cast<const int32_t*>(base + 2 * nb);
  }
  const int Kld = g.num_k_blocks * kBK;  // K padded to whole blocks (zeros)
  const int64_t tot = (int64_t)groups * g.Npad * Kld;
  Scratch bsplit((size_t)2 * tot * 4, s);
  split_weights_kernel<<<(int)std::min<int64_t>((tot + 255) / 256, 1184), 256, 0, s>>>(
      B, b_is_kn, (int)N, (int)K, g.Npad, Kld, groups, bsplit.as<float>());
  B200_CUDA(cudaGetLastError());
  const CUtensorMap map_a = make_map(A, M, K, lda, kBM);
  const CUtensorMap map_b =
      make_map(bsplit.as<float>(), 2 * (int64_t)groups * g.Npad, Kld, Kld, g.Npad);
  const size_t smem = (size_t)g.stages * g.stage_bytes + (3 * g.stages + 4) * 8 + 16 + 1024;
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(gemm_tf32x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         227 * 1024);
  });
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = std::min(g.num_m_tiles, sms);
  gemm_tf32x3_kernel<<<grid, kThreads, smem, s>>>(map_a, map_b, g);
  B200_CUDA(cudaGetLastError());
  note_kernel("gemm_tf32x3", 2);
}

}  // namespace b200

// Neighbour-access micro-benchmark (our counterpart of the reference's NbAccess harness,
// src/kernel/cuda/binary_reduce_impl.cu:443-900): random row gather + sequential row reduce,
// comparing ways of getting neighbour rows onto the SM on B200. Not part of the product library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bench gather_bench.cu
//   ./gather_bench [N=232965] [D=602] [deg=492]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cmath>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t b) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(b) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t par) {
  uint32_t ok;
  do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(par) : "memory"); } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// ---------------- variant A: register staged, warp per (row, slab), U neighbours in flight
__device__ __forceinline__ float2 ld_na(const float2* p) {
  float2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0, %1}, [%2];" : "=f"(r.x), "=f"(r.y) : "l"(p));
  return r;
}
template <int CH, int U, int NA>
__global__ void __launch_bounds__(256) k_reg(const int* __restrict__ indptr, const int* __restrict__ idx,
                                             const float* __restrict__ X, float* __restrict__ out,
                                             int n_rows, int D, int slabs) {
  const int lane = threadIdx.x & 31;
  const long warp = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long row = warp / slabs; const int slab = warp % slabs;
  if (row >= n_rows) return;
  const int nvec = D / 2;
  const int beg = indptr[row], deg = indptr[row + 1] - beg;
  float2 acc[CH]; int col[CH]; bool ok[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) { acc[c] = make_float2(0.f, 0.f); int vc = slab * 32 * CH + c * 32 + lane; ok[c] = vc < nvec; col[c] = vc * 2; }
  for (int base = 0; base < deg; base += 32) {
    int mine = (base + lane < deg) ? idx[beg + base + lane] : 0;
    int cnt = min(32, deg - base);
    for (int i = 0; i < cnt; i += U) {
      float2 v[U][CH];
#pragma unroll
      for (int k = 0; k < U; ++k) {
        int ia = __shfl_sync(0xffffffffu, mine, (i + k) & 31);
        bool on = i + k < cnt;
#pragma unroll
        for (int c = 0; c < CH; ++c) if (on && ok[c]) v[k][c] = NA ? ld_na((const float2*)(X + (long)ia * D + col[c])) : __ldg((const float2*)(X + (long)ia * D + col[c]));
      }
#pragma unroll
      for (int k = 0; k < U; ++k) if (i + k < cnt) {
#pragma unroll
        for (int c = 0; c < CH; ++c) if (ok[c]) { acc[c].x = __fadd_rn(acc[c].x, v[k][c].x); acc[c].y = __fadd_rn(acc[c].y, v[k][c].y); }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CH; ++c) if (ok[c]) *(float2*)(out + row * D + col[c]) = acc[c];
}

// ---------------- variant B: TMA bulk copy, warp-private ring of S slots, G rows per barrier
template <int CH>
__global__ void k_tma(const int* __restrict__ indptr, const int* __restrict__ idx,
                      const float* __restrict__ X, float* __restrict__ out, int n_rows, int D,
                      int S, int G, int slot_bytes, int warps, int issue_mode) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  unsigned char* ring = smem + (size_t)wid * S * G * slot_bytes;
  const uint32_t ring32 = smem_u32(ring);
  const uint32_t bars = smem_u32(smem + (size_t)warps * S * G * slot_bytes) + wid * S * 8;
  if (lane == 0) { for (int s = 0; s < S; ++s) mbar_init(bars + s * 8, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncwarp();
  const long row = (long)blockIdx.x * warps + wid;
  if (row >= n_rows) return;
  const int nvec = D / 2;
  const int beg = indptr[row], deg = indptr[row + 1] - beg;
  const uint32_t row_bytes = D * 4;
  float2 acc[CH]; int col[CH]; bool ok[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) { acc[c] = make_float2(0.f, 0.f); int vc = c * 32 + lane; ok[c] = vc < nvec; col[c] = vc * 8; }
  const int ngroups = (deg + G - 1) / G;
  // group g -> stage g % S; rows g*G .. g*G+G-1
  auto issue = [&](int g) {
    const int st = g % S;
    const int j0 = g * G;
    const int n = min(G, deg - j0);
    const uint32_t bar = bars + st * 8;
    if (issue_mode == 0) {
      if (lane == 0) {
        uint32_t total = 0;
        for (int k = 0; k < n; ++k) { const float* p = X + (long)idx[beg + j0 + k] * D; uint32_t off = (uintptr_t)p & 15; total += (off + row_bytes + 15) & ~15u; }
        mbar_expect(bar, total);
        for (int k = 0; k < n; ++k) { const float* p = X + (long)idx[beg + j0 + k] * D; uint32_t off = (uintptr_t)p & 15; uint32_t b = (off + row_bytes + 15) & ~15u;
          bulk_g2s(ring32 + (st * G + k) * slot_bytes, (const unsigned char*)p - off, b, bar); }
      }
    } else {
      // lanes 0..n-1 each issue one copy; lane 0 posts the expected bytes first
      uint32_t b = 0; const float* p = nullptr; uint32_t off = 0;
      if (lane < n) { p = X + (long)idx[beg + j0 + lane] * D; off = (uintptr_t)p & 15; b = (off + row_bytes + 15) & ~15u; }
      uint32_t total = b;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
      if (lane == 0) mbar_expect(bar, total);
      __syncwarp();
      if (lane < n) bulk_g2s(ring32 + (st * G + lane) * slot_bytes, (const unsigned char*)p - off, b, bar);
    }
  };
  const int pre = min(S - 1, ngroups);
  for (int g = 0; g < pre; ++g) issue(g);
  for (int g = 0; g < ngroups; ++g) {
    __syncwarp();
    if (g + S - 1 < ngroups) issue(g + S - 1);
    const int st = g % S;
    mbar_wait(bars + st * 8, (g / S) & 1);
    const int n = min(G, deg - g * G);
    for (int k = 0; k < n; ++k) {
      const float* p = X + (long)idx[beg + g * G + k] * D;
      const uint32_t off = (uintptr_t)p & 15;
      const unsigned char* sp = ring + (size_t)(st * G + k) * slot_bytes + off;
#pragma unroll
      for (int c = 0; c < CH; ++c) if (ok[c]) { float2 v = *(const float2*)(sp + col[c]); acc[c].x = __fadd_rn(acc[c].x, v.x); acc[c].y = __fadd_rn(acc[c].y, v.y); }
    }
  }
#pragma unroll
  for (int c = 0; c < CH; ++c) if (ok[c]) *(float2*)(out + row * D + col[c] / 4) = acc[c];
}

int main(int argc, char** argv) {
  int N = argc > 1 ? atoi(argv[1]) : 232965, D = argc > 2 ? atoi(argv[2]) : 602, deg = argc > 3 ? atoi(argv[3]) : 492;
  double alpha = argc > 4 ? atof(argv[4]) : 0.0; int do_tma = argc > 5 ? atoi(argv[5]) : 1;
  long E = (long)N * deg;
  printf("N=%d D=%d deg=%d E=%ld bytes gathered=%.1f GB\n", N, D, deg, E, E * D * 4.0 / 1e9);
  std::vector<int> indptr(N + 1), idx(E);
  uint64_t s = 88172645463325252ull;
  for (int i = 0; i <= N; ++i) indptr[i] = (int)((long)i * deg);
  for (long e = 0; e < E; ++e) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; double u = (s >> 11) * (1.0 / 9007199254740992.0); idx[e] = alpha > 0 ? (int)fmin(N - 1.0, pow((pow((double)N, 1 - alpha) - 1) * u + 1, 1 / (1 - alpha)) - 1) : (int)(s % N); }
  if (alpha > 0) for (long e = 0; e < E; ++e) idx[e] = (int)(((unsigned long)idx[e] * 2654435761ul) % N);
  int *d_ip, *d_idx; float *X, *out;
  CK(cudaMalloc(&d_ip, (N + 1) * 4)); CK(cudaMalloc(&d_idx, E * 4));
  CK(cudaMalloc(&X, (size_t)N * D * 4 + 64)); CK(cudaMalloc(&out, (size_t)N * D * 4));
  CK(cudaMemcpy(d_ip, indptr.data(), (N + 1) * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_idx, idx.data(), E * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(X, 0, (size_t)N * D * 4));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  auto timeit = [&](const char* name, auto launch) {
    launch(); CK(cudaDeviceSynchronize());
    cudaEventRecord(a); for (int r = 0; r < 3; ++r) launch(); cudaEventRecord(b); CK(cudaEventSynchronize(b));
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 3;
    printf("%-44s %8.2f ms  %7.0f GB/s (gather bytes)\n", name, ms, E * D * 4.0 / ms / 1e6);
  };
  char nm[128];
  const int nvec = D / 2;
#define REG(CHV, UV, NA, TPB) { int slabs = (nvec + 32 * CHV - 1) / (32 * CHV); long warps = (long)N * slabs; int wpb = TPB / 32; int blocks = (int)((warps + wpb - 1) / wpb); \
    sprintf(nm, "reg CH=%d U=%d slabs=%d na=%d tpb=%d", CHV, UV, slabs, NA, TPB); timeit(nm, [&] { k_reg<CHV, UV, NA><<<blocks, TPB>>>(d_ip, d_idx, X, out, N, D, slabs); }); }
  if (nvec <= 32 * 10 * 4) {
    REG(10, 1, 0, 256) REG(10, 1, 1, 256) REG(10, 1, 0, 128) REG(10, 2, 1, 256)
    REG(5, 1, 0, 256) REG(5, 1, 1, 256) REG(5, 2, 0, 256) REG(5, 2, 1, 256) REG(5, 4, 1, 256) REG(5, 2, 0, 128)
    REG(3, 1, 0, 256) REG(3, 2, 0, 256) REG(3, 2, 1, 256) REG(3, 4, 1, 256)
    REG(2, 1, 0, 256) REG(2, 2, 0, 256) REG(2, 4, 0, 256) REG(2, 4, 1, 256) REG(2, 8, 1, 256)
    REG(1, 2, 0, 256) REG(1, 4, 0, 256) REG(1, 8, 0, 256) REG(1, 8, 1, 256)
  }
  const int slot = (D * 4 + 30) & ~15;
#define TMA(CHV, SV, GV, WV, MODE) if (nvec <= 32 * CHV) { size_t smem = (size_t)WV * SV * GV * slot + WV * SV * 8; \
    if (smem <= 227 * 1024) { auto kf = k_tma<CHV>; cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    int blocks = (N + WV - 1) / WV; sprintf(nm, "tma CH=%d S=%d G=%d warps=%d mode=%d smem=%zuK", CHV, SV, GV, WV, MODE, smem / 1024); \
    timeit(nm, [&] { kf<<<blocks, WV * 32, smem>>>(d_ip, d_idx, X, out, N, D, SV, GV, slot, WV, MODE); }); } }
  if (do_tma) {
  TMA(10, 9, 1, 4, 0) TMA(10, 4, 1, 4, 0) TMA(10, 2, 1, 4, 0) TMA(10, 16, 1, 4, 0) TMA(10, 8, 1, 8, 0) TMA(10, 4, 1, 16, 0)
  TMA(10, 4, 2, 4, 0) TMA(10, 2, 4, 4, 0) TMA(10, 3, 4, 4, 1) TMA(10, 2, 8, 4, 1) TMA(10, 4, 2, 4, 1) TMA(10, 4, 1, 2, 0) TMA(10, 8, 1, 1, 0)
  TMA(10, 2, 2, 16, 1) TMA(10, 3, 1, 24, 0)
  }
  return 0;
}

// Fused edge_softmax forward / backward (extension; the reference runs it as 4 + 2 builtin kernel
// calls plus torch elementwise ops, python/dgl/nn/pytorch/softmax.py:69-78, 104-110).
//
//   forward : out[e,h]  = exp(z[e,h] - max_{e' -> dst(e)} z[e',h]) / sum_{e' -> dst(e)} exp(...)
//   backward: gz[e,h]   = out*go - out * sum_{e' -> dst(e)} (out*go)[e',h]
//
// One warp owns one destination row and walks its in-edges ONCE from HBM: lanes are laid out as
// (edge slot, head) pairs -- Hp = pow2 >= H lanes per edge, 32/Hp edges in flight -- the first
// kCache rounds of values stay in registers, longer rows re-read from L1/L2. Per-head max / sum are
// reduced across the edge slots with a fixed xor butterfly, so the result is deterministic and
// every output element is written by exactly one lane (no atomics, no identity fill). Rows are
// walked over in_s (in-CSR ordered by (source, edge id)), the order of the reference's sums.
// Bound: 4-byte-per-(edge, head) traffic, z and out indexed by edge id: 2*4*E*H + 8*E + 4*N bytes.
#include <algorithm>

#include "common.cuh"

namespace b200 {

namespace {

constexpr int kCache = 4;  // rounds of (edge, head) values kept in registers per lane

struct SoftmaxArgs {
  const int32_t* indptr;
  const int32_t* eids;
  const float* a;   // forward: z; backward: out
  const float* b;   // backward: grad_out
  float* dst;       // forward: out; backward: grad_z
  int n_rows;
  int H;
};

template <int HP>
__device__ __forceinline__ float slots_max(float v) {
#pragma unroll
  for (int o = 16; o >= HP; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
template <int HP>
__device__ __forceinline__ float slots_sum(float v) {
#pragma unroll
  for (int o = 16; o >= HP; o >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// HP lanes per edge (HP = pow2 >= min(H, 32)); heads beyond 32 are handled in blocks of 32.
template <int HP, bool FWD>
__global__ void __launch_bounds__(256) edge_softmax_kernel(SoftmaxArgs p) {
  constexpr int SLOTS = 32 / HP;
  const int lane = threadIdx.x & 31;
  const int sub = lane % HP, slot = lane / HP;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < p.n_rows; row += warps) {
    const int beg = __ldg(p.indptr + row), end = __ldg(p.indptr + row + 1);
    if (beg == end) continue;
    for (int h0 = 0; h0 < p.H; h0 += HP) {
      const int h = h0 + sub;
      const bool live = h < p.H;
      float cache[kCache];
      int64_t off[kCache];
      // pass 1: load (and cache) values, running max (forward) / running sum of out*go (backward)
      float red = FWD ? -FLT_MAX : 0.f;
#pragma unroll
      for (int k = 0; k < kCache; ++k) {
        const int j = beg + slot + k * SLOTS;
        cache[k] = FWD ? -FLT_MAX : 0.f;
        off[k] = -1;
        if (j < end && live) {
          off[k] = (int64_t)__ldg(p.eids + j) * p.H + h;
          if (FWD) {
            cache[k] = __ldg(p.a + off[k]);
            red = fmaxf(red, cache[k]);
          } else {
            cache[k] = __fmul_rn(__ldg(p.a + off[k]), __ldg(p.b + off[k]));
            red = __fadd_rn(red, cache[k]);
          }
        }
      }
      for (int j = beg + slot + kCache * SLOTS; j < end; j += SLOTS) {
        if (!live) break;
        const int64_t o = (int64_t)__ldg(p.eids + j) * p.H + h;
        if (FWD) red = fmaxf(red, __ldg(p.a + o));
        else red = __fadd_rn(red, __fmul_rn(__ldg(p.a + o), __ldg(p.b + o)));
      }
      red = FWD ? slots_max<HP>(red) : slots_sum<HP>(red);
      if (FWD) {
        // pass 2: exp and its sum; long rows park exp in `out` and rescale in pass 3
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < kCache; ++k) {
          if (off[k] >= 0) {
            cache[k] = expf(__fsub_rn(cache[k], red));
            s = __fadd_rn(s, cache[k]);
          }
        }
        for (int j = beg + slot + kCache * SLOTS; j < end; j += SLOTS) {
          if (!live) break;
          const int64_t o = (int64_t)__ldg(p.eids + j) * p.H + h;
          const float e = expf(__fsub_rn(__ldg(p.a + o), red));
          p.dst[o] = e;
          s = __fadd_rn(s, e);
        }
        s = slots_sum<HP>(s);
#pragma unroll
        for (int k = 0; k < kCache; ++k)
          if (off[k] >= 0) p.dst[off[k]] = __fdiv_rn(cache[k], s);
        for (int j = beg + slot + kCache * SLOTS; j < end; j += SLOTS) {
          if (!live) break;
          const int64_t o = (int64_t)__ldg(p.eids + j) * p.H + h;
          p.dst[o] = __fdiv_rn(p.dst[o], s);  // written by this same lane above
        }
      } else {
        // grad_z = sds - out * accum   (softmax.py:104-110: grad_s - out * accum)
#pragma unroll
        for (int k = 0; k < kCache; ++k)
          if (off[k] >= 0)
            p.dst[off[k]] = __fsub_rn(cache[k], __fmul_rn(__ldg(p.a + off[k]), red));
        for (int j = beg + slot + kCache * SLOTS; j < end; j += SLOTS) {
          if (!live) break;
          const int64_t o = (int64_t)__ldg(p.eids + j) * p.H + h;
          const float a = __ldg(p.a + o);
          p.dst[o] = __fsub_rn(__fmul_rn(a, __ldg(p.b + o)), __fmul_rn(a, red));
        }
      }
    }
  }
}

template <bool FWD>
void launch(const SoftmaxArgs& p, cudaStream_t s) {
  if (p.n_rows == 0) return;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int blocks = (int)std::min<int64_t>(ceil_div(p.n_rows, 8), (int64_t)sms * 64);
#define B200_SM(HP) edge_softmax_kernel<HP, FWD><<<blocks, 256, 0, s>>>(p)
  if (p.H <= 1) B200_SM(1);
  else if (p.H <= 2) B200_SM(2);
  else if (p.H <= 4) B200_SM(4);
  else if (p.H <= 8) B200_SM(8);
  else if (p.H <= 16) B200_SM(16);
  else B200_SM(32);
#undef B200_SM
  B200_CUDA(cudaGetLastError());
}

}  // namespace

void edge_softmax_forward(DGLB200Graph* g, const float* z, float* out, int64_t H, cudaStream_t s) {
  SoftmaxArgs p{g->in_s.indptr, g->in_s.eids, z, nullptr, out, (int)g->num_dst, (int)H};
  launch<true>(p, s);
  note_kernel("edge_softmax_fwd", 1);
}

void edge_softmax_backward(DGLB200Graph* g, const float* out, const float* gout, float* gz,
                           int64_t H, cudaStream_t s) {
  SoftmaxArgs p{g->in_s.indptr, g->in_s.eids, out, gout, gz, (int)g->num_dst, (int)H};
  launch<false>(p, s);
  note_kernel("edge_softmax_bwd", 1);
}

}  // namespace b200

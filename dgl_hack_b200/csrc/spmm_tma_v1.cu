#include "spmm_tma_kernels.cuh"
namespace b200 {
template void launch_tma_vec<1>(const TmaArgs&, int, bool, bool, int, int, size_t, cudaStream_t);
}

#include "spmm_kernels.cuh"
namespace b200 {
template void launch_vec<2>(const SpmmArgs&, int, int, bool, int, cudaStream_t);
}

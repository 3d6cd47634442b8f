// Edge-output (SDDMM-style) fast paths. Declines everything for now: generic_edge_kernel in
// generic.cu serves reducer "none" and edge-target gradients until the specialised kernels land.
#include "common.cuh"

namespace b200 {
bool try_sddmm(const KernelCall&) { return false; }
}  // namespace b200

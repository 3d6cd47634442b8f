// TEST INFRASTRUCTURE ONLY (oracle build recipe) -- not product code.
//
// The fork references seven functions unconditionally from src/kernel/binary_reduce.cc:395-449,548
// (declared in src/kernel/binary_reduce_impl_decl.h:150-203) but defines them only in the CUDA
// file src/kernel/cuda/binary_reduce_impl.cu. A CPU-only libdgl.so therefore fails to load with
// undefined symbols. This translation unit is the single non-reference file linked into
// oracle/_ref/libdgl.so: each stub refuses loudly, exactly as a CPU-only build should.
#include <dgl/graph_interface.h>
#include <dgl/runtime/ndarray.h>
#include <dmlc/logging.h>
#include "kernel/binary_reduce_impl_decl.h"

namespace dgl {
namespace kernel {

using runtime::NDArray;

void FusedGatKernelImpl(const CSRWrapper&, NDArray, NDArray, NDArray, NDArray, NDArray,
                        NDArray, float) {
  LOG(FATAL) << "FusedGatKernelImpl is CUDA-only in the reference";
}
void NbAccessImpl(const CSRWrapper&, NDArray, NDArray, NDArray) {
  LOG(FATAL) << "NbAccessImpl is CUDA-only in the reference";
}
void RgcnLayer0Impl(GraphRef, NDArray, NDArray, NDArray) {
  LOG(FATAL) << "RgcnLayer0Impl is CUDA-only in the reference";
}
void RgcnLayer1Impl(GraphRef, NDArray, NDArray, NDArray, NDArray) {
  LOG(FATAL) << "RgcnLayer1Impl is CUDA-only in the reference";
}
void RgcnLayer0BackwardImpl(GraphRef, NDArray, NDArray, NDArray) {
  LOG(FATAL) << "RgcnLayer0BackwardImpl is CUDA-only in the reference";
}
void RgcnLayer1BackwardImpl(GraphRef, NDArray, NDArray, NDArray, NDArray, NDArray, NDArray) {
  LOG(FATAL) << "RgcnLayer1BackwardImpl is CUDA-only in the reference";
}
void BackwardFusedGatKernelImpl(const CSRWrapper&, NDArray, NDArray, NDArray, NDArray, NDArray,
                                NDArray, NDArray, NDArray, NDArray, NDArray, float) {
  LOG(FATAL) << "BackwardFusedGatKernelImpl is CUDA-only in the reference";
}

}  // namespace kernel
}  // namespace dgl

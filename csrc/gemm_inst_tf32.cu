// Explicit instantiation of the tcgen05 GEMM kernels for one operand / output type pair (see gemm_kernel.cuh).
#include "gemm_kernel.cuh"

namespace b200 {
template int dispatch_bn<float, float>(int, bool, bool, bool, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                 const CUtensorMap&, const CUtensorMap&, const GemmArgs&, cudaStream_t);
}  // namespace b200

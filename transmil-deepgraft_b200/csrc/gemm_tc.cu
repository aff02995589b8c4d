// tcgen05 / TMA GEMM family -- placeholder until the tensor-core kernels land.
#include "common.cuh"
namespace tmil {
bool tc_available() { return false; }
int gemm_tc(const GemmParams&, cudaStream_t) {
  set_error("tcgen05 GEMM not built");
  return TM_ERR_UNSUPPORTED_ARCH;
}
}  // namespace tmil

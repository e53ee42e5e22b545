// placeholder: replaced by the tcgen05 + TMA GEMM
#include "kernels.h"
namespace gp {
bool gemm_tc_supported(int, int, int) { return false; }
cudaError_t init_gemm_tc() { return cudaSuccess; }
cudaError_t launch_gemm_tc(const void*, int, const void*, const float*, const float*, int, void*, int, int, int, int,
                           uint32_t, cudaStream_t) { return cudaErrorNotSupported; }
}  // namespace gp

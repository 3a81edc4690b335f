// placeholder until the tcgen05 kernels land
#include "common.cuh"
namespace nsg {
int conv_gemm_tc(const nsg_conv_geom*, const void*, const float*, const nsg_prologue*, const nsg_epilogue*, void*,
                 cudaStream_t) {
  set_error("conv_gemm: tensor-core implementation not built");
  return NSG_E_UNSUPPORTED;
}
int conv_wgrad_tc(const nsg_conv_geom*, const void*, const nsg_prologue*, const void*, const nsg_prologue*, float*,
                  float*, cudaStream_t) {
  set_error("conv_wgrad: tensor-core implementation not built");
  return NSG_E_UNSUPPORTED;
}
}  // namespace nsg

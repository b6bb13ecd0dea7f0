// tcgen05 / TMEM implicit-GEMM kernels (FNET_MATH_TF32, FNET_MATH_BF16).  Placeholder until the
// tensor-core kernels land: nothing is "supported", so every call takes the fp32 FFMA path.
#include "igemm.cuh"

namespace fnet {

bool tc_supported(const ConvGeom &, int, int) { return false; }
size_t tc_conv_workspace(const ConvGeom &, int) { return 0; }
int tc_conv_fwd(const ConvGeom &, int, int, const float *, const float *, const float *, float *, void *, size_t,
                cudaStream_t) { set_error("tensor-core path not built"); return FNET_ERR_UNSUPPORTED; }
int tc_conv_dgrad(const ConvGeom &, int, const float *, const float *, float *, int, void *, size_t, cudaStream_t) {
    set_error("tensor-core path not built"); return FNET_ERR_UNSUPPORTED; }
int tc_conv_wgrad(const ConvGeom &, int, const float *, const float *, float *, float *, void *, size_t,
                  cudaStream_t) { set_error("tensor-core path not built"); return FNET_ERR_UNSUPPORTED; }

}  // namespace fnet

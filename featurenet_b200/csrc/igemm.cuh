// Internal interface between the C-ABI wrappers (conv_api.cu) and the GEMM-shaped kernels.
#pragma once
#include "common.cuh"

namespace fnet {

struct ConvGeom {
    int B, H, W, C, Ho, Wo, K, kh, kw, sh, sw, pt, pl, dh, dw;
};

// What a data gradient does with its result: dx = (accumulate ? dx + v : v) * act'(y).  act / y describe the
// activation of the layer that PRODUCED x (y = that layer's output, same shape as dx): folding its derivative into
// this epilogue saves the separate act_bwd pass over the tensor.  act == FNET_ACT_NONE: plain store.
struct DxEpilogue {
    int accumulate;
    int act;
    const float *y;
};

int geom_checked(const fnet_conv2d_desc *d, ConvGeom &g, const char *what);
ConvGeom dense_geom(long long m, long long k, long long n);

// FNET_MATH_FP32: FFMA implicit GEMM (igemm_simt.cu)
size_t simt_conv_workspace(const ConvGeom &g);
int simt_conv_fwd(const ConvGeom &g, int act, const float *x, const float *w, const float *bias, float *y, void *ws,
                  size_t ws_bytes, cudaStream_t stream);
int simt_conv_dgrad(const ConvGeom &g, const float *dz, const float *w, float *dx, DxEpilogue ep, void *ws,
                    size_t ws_bytes, cudaStream_t stream);
int simt_conv_wgrad(const ConvGeom &g, const float *x, const float *dz, float *dw, float *dbias, void *ws,
                    size_t ws_bytes, cudaStream_t stream);

// FNET_MATH_TF32 / FNET_MATH_BF16: tcgen05 implicit GEMM (igemm_tc.cu)
bool tc_supported(const ConvGeom &g, int math, int product);   // product: 0 fwd, 1 dgrad, 2 wgrad
size_t tc_conv_workspace(const ConvGeom &g, int math);
int tc_conv_fwd(const ConvGeom &g, int math, int act, const float *x, const float *w, const float *bias, float *y,
                void *ws, size_t ws_bytes, cudaStream_t stream);
int tc_conv_dgrad(const ConvGeom &g, int math, const float *dz, const float *w, float *dx, DxEpilogue ep, void *ws,
                  size_t ws_bytes, cudaStream_t stream);
int tc_conv_wgrad(const ConvGeom &g, int math, const float *x, const float *dz, float *dw, float *dbias, void *ws,
                  size_t ws_bytes, cudaStream_t stream);

// FNET_MATH_TF32 / FNET_MATH_TF32X3, layers whose operands are boxes (Dense, 1x1 and k x k stride-1 convolutions): TMA-fed persistent
// kernel (gemm_tma.cu)
bool tma_gemm_supported(const ConvGeom &g, int math, int product);   // product: 0 fwd, 1 dgrad, 2 wgrad
int tma_gemm_fwd(const ConvGeom &g, int math, int act, const float *x, const float *w, const float *bias, float *y,
                 cudaStream_t stream);
int tma_gemm_dgrad(const ConvGeom &g, int math, const float *dz, const float *w, float *dx, DxEpilogue ep, cudaStream_t stream);
size_t tma_gemm_workspace(const ConvGeom &g, int math);
int tma_gemm_wgrad(const ConvGeom &g, int math, const float *x, const float *dz, float *dw, float *dbias, void *ws,
                   size_t ws_bytes, cudaStream_t stream);
int simt_reduce_wgrad_partials(const ConvGeom &g, float *dw, float *dbias, int M, int splits, const float *partial,
                               cudaStream_t stream);

// direct kernels for the classifier head and for small-channel convolutions (skinny.cu); fp32, any math mode
bool skinny_supported(const ConvGeom &g, int product, bool with_bias);
size_t skinny_workspace(const ConvGeom &g);
int skinny_fwd(const ConvGeom &g, int act, const float *x, const float *w, const float *bias, float *y, void *ws,
               size_t ws_bytes, cudaStream_t s);
// y = act(bias + sum of `splits` partial [M x N] products), fixed order (igemm_simt.cu)
int simt_reduce_fwd_partials(const ConvGeom &g, int act, const float *bias, float *y, int splits, const float *partial,
                             cudaStream_t stream);
int skinny_dgrad(const ConvGeom &g, const float *dz, const float *w, float *dx, DxEpilogue ep, cudaStream_t s);
int skinny_wgrad(const ConvGeom &g, const float *x, const float *dz, float *dw, float *dbias, void *ws, size_t ws_bytes,
                 cudaStream_t s);

// data gradient for few input channels: dense GEMM into T (workspace) + col2im (col2im.cu)
size_t col2im_t_bytes(const ConvGeom &g);                    // 0 = not applicable
int col2im_scatter(const ConvGeom &g, const float *t, float *dx, DxEpilogue ep, cudaStream_t stream);

// very few input channels, up to 64 output channels: direct kernels over groups of eight output channels (fewc.cu)
bool fewc_supported(const ConvGeom &g, int product, bool with_bias);
size_t fewc_workspace(const ConvGeom &g);
int fewc_fwd(const ConvGeom &g, int act, const float *x, const float *w, const float *bias, float *y, cudaStream_t s);
int fewc_dgrad(const ConvGeom &g, const float *dz, const float *w, float *dx, DxEpilogue ep, cudaStream_t s);
int fewc_wgrad(const ConvGeom &g, const float *x, const float *dz, float *dw, float *dbias, void *ws, size_t ws_bytes, cudaStream_t s);

// 1x1 convolutions with a stride as "keep every s-th pixel" + a plain GEMM (strided.cu)
bool strided_pointwise_ok(const ConvGeom &g);
ConvGeom strided_inner(const ConvGeom &g);                   // the stride-1 layer over the kept pixels
size_t strided_sub_bytes(const ConvGeom &g);                 // bytes of sub(x) / dx_sub in the workspace
int strided_subsample(const ConvGeom &g, const float *x, float *x_sub, cudaStream_t stream);
int strided_scatter(const ConvGeom &g, const float *dx_sub, float *dx, DxEpilogue ep, cudaStream_t stream);

// k x k stride-1 convolutions, forward and data gradient, with the input tile resident in shared memory (dconv.cu)
bool dconv_supported(const ConvGeom &g, int math, int product);
size_t dconv_workspace(const ConvGeom &g, int math);
int dconv_fwd(const ConvGeom &g, int math, int act, const float *x, const float *w, const float *bias, float *y, void *ws,
              size_t ws_bytes, cudaStream_t stream);
int dconv_dgrad(const ConvGeom &g, int math, const float *dz, const float *w, float *dx, DxEpilogue ep, void *ws, size_t ws_bytes,
                cudaStream_t stream);

int run_softmax_rows(int64_t rows, int64_t cols, const float *x, float *y, cudaStream_t stream);

}  // namespace fnet

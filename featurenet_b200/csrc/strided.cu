// 1x1 convolutions with a stride (most convolutions of FeatureNet's NAS spaces: `Cout x Cin, kernel 1x1, stride 2`).
//
// A 1x1 kernel never needs padding, so the layer is "keep every s-th pixel, then a plain GEMM over the kept pixels":
//   forward          y      = sub(x) . w
//   filter gradient  dw     = sub(x)^T . dz
//   data gradient    dx_sub = dz . w^T,  dx = dx_sub at the kept pixels and 0 elsewhere
// Treated as a strided implicit GEMM the data gradient walks ALL input pixels, three quarters of which receive nothing
// (16x8x192 -> 1536 at batch 128: 268 us); as a plain GEMM on the kept pixels every product takes the kernels the
// stride-1 layers take (TMA-fed / pointwise / FFMA), plus one small copy: sub(x) into the workspace, or dx_sub out of it
// with the zero fill, the accumulate flag and the activation derivative of the producing layer in the same pass.
#include "igemm.cuh"

namespace fnet {

bool strided_pointwise_ok(const ConvGeom &g) {
    static const int mode = env_switch("FNET_STRIDED_POINTWISE", 1);     // 0 disables (A/B measurements)
    // from 16 input channels on: below that the copy of sub(x) costs a launch that the GEMM does not win back
    // (3 -> 24 on 32x32, stride 2: 15.4 -> 17.3 us forward, 16.0 -> 19.9 us filter gradient with the extra launch)
    // ... and from half a million input elements on: a layer smaller than that is one launch either way, and the copy
    // would be a second one (the deep 2x1 / 1x1 layers of the sampled products: 128 pixels x 96..768 channels)
    return mode && g.C >= 16 && (long long)g.B * g.H * g.W * g.C >= (1LL << 19) && g.kh == 1 && g.kw == 1 &&
           (g.sh > 1 || g.sw > 1) && g.pt == 0 && g.pl == 0 &&
           (long long)g.B * g.H * g.W * g.C < 0x7fffffffLL && (long long)g.B * g.Ho * g.Wo * g.C < 0x7fffffffLL;
}

ConvGeom strided_inner(const ConvGeom &g) {
    return ConvGeom{g.B, g.Ho, g.Wo, g.C, g.Ho, g.Wo, g.K, 1, 1, 1, 1, 0, 0, 1, 1};
}

size_t strided_sub_bytes(const ConvGeom &g) {
    return (((size_t)g.B * g.Ho * g.Wo * g.C * sizeof(float)) + 255) & ~(size_t)255;
}

// x_sub[b][ho][wo][:] = x[b][ho*sh][wo*sw][:]   (T = float4 when C % 4 == 0: c counts quads)
template <typename T>
__global__ void __launch_bounds__(256) subsample_kernel(int total, int c, int Ho, int Wo, int H, int W, int sh, int sw,
                                                        const T *__restrict__ x, T *__restrict__ x_sub) {
    pdl_prologue();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int ch = i % c, p = i / c, wo = p % Wo, t = p / Wo, ho = t % Ho, b = t / Ho;
        x_sub[i] = __ldg(x + ((long long)(b * H + ho * sh) * W + wo * sw) * c + ch);
    }
}

__device__ __forceinline__ float4 dx_finish(float4 v, const float4 *dx, const float4 *y, int i, int accumulate, int act) {
    if (accumulate) { const float4 o = dx[i]; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
    if (act) {
        const float4 yy = __ldg(y + i);
        v.x *= act_grad(yy.x, act); v.y *= act_grad(yy.y, act); v.z *= act_grad(yy.z, act); v.w *= act_grad(yy.w, act);
    }
    return v;
}
__device__ __forceinline__ float dx_finish(float v, const float *dx, const float *y, int i, int accumulate, int act) {
    if (accumulate) v += dx[i];
    if (act) v *= act_grad(__ldg(y + i), act);
    return v;
}
__device__ __forceinline__ void zero_of(float4 &v) { v = make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ void zero_of(float &v) { v = 0.f; }

// dx[b][h][w][:] = (accumulate ? dx : 0) + (kept pixel ? dx_sub[b][h/sh][w/sw][:] : 0), times act'(y)
template <typename T>
__global__ void __launch_bounds__(256) strided_scatter_kernel(int total, int c, int Ho, int Wo, int H, int W, int sh, int sw,
                                                              const T *__restrict__ dx_sub, T *dx, int accumulate, int act,
                                                              const T *__restrict__ y) {
    pdl_prologue();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int ch = i % c, p = i / c, w = p % W, t = p / W, h = t % H, b = t / H;
        const int ho = h / sh, wo = w / sw;
        T v;
        zero_of(v);
        if (ho * sh == h && wo * sw == w && ho < Ho && wo < Wo) v = __ldg(dx_sub + ((long long)(b * Ho + ho) * Wo + wo) * c + ch);
        dx[i] = dx_finish(v, dx, y, i, accumulate, act);
    }
}

static bool quads_ok(const ConvGeom &g, const void *a, const void *b, const void *c) {
    return !(g.C & 3) && !((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c)) & 15);
}

int strided_subsample(const ConvGeom &g, const float *x, float *x_sub, cudaStream_t stream) {
    if (quads_ok(g, x, x_sub, nullptr)) {
        const int c = g.C >> 2, total = g.B * g.Ho * g.Wo * c;
        launch_pdl(subsample_kernel<float4>, dim3(stream_grid(total, 256)), dim3(256), 0, stream, total, c, g.Ho, g.Wo, g.H, g.W, g.sh,
                   g.sw, reinterpret_cast<const float4 *>(x), reinterpret_cast<float4 *>(x_sub));
    } else {
        const int total = g.B * g.Ho * g.Wo * g.C;
        launch_pdl(subsample_kernel<float>, dim3(stream_grid(total, 256)), dim3(256), 0, stream, total, g.C, g.Ho, g.Wo, g.H, g.W, g.sh,
                   g.sw, x, x_sub);
    }
    return check_launch("conv2d[1x1 strided]: subsample");
}

int strided_scatter(const ConvGeom &g, const float *dx_sub, float *dx, DxEpilogue ep, cudaStream_t stream) {
    if (quads_ok(g, dx_sub, dx, ep.act ? ep.y : nullptr)) {
        const int c = g.C >> 2, total = g.B * g.H * g.W * c;
        launch_pdl(strided_scatter_kernel<float4>, dim3(stream_grid(total, 256)), dim3(256), 0, stream, total, c, g.Ho, g.Wo, g.H, g.W,
                   g.sh, g.sw, reinterpret_cast<const float4 *>(dx_sub), reinterpret_cast<float4 *>(dx), ep.accumulate, ep.act,
                   reinterpret_cast<const float4 *>(ep.y));
    } else {
        const int total = g.B * g.H * g.W * g.C;
        launch_pdl(strided_scatter_kernel<float>, dim3(stream_grid(total, 256)), dim3(256), 0, stream, total, g.C, g.Ho, g.Wo, g.H, g.W,
                   g.sh, g.sw, dx_sub, dx, ep.accumulate, ep.act, ep.y);
    }
    return check_launch("conv2d_bwd_data[1x1 strided]: scatter");
}

}  // namespace fnet

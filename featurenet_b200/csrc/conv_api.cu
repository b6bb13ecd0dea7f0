// C-ABI entry points of the GEMM-shaped ops (Conv2D, Dense): argument checks and the choice
// between the fp32 FFMA kernels and the tcgen05 tensor-core kernels.
#include "igemm.cuh"

using namespace fnet;

namespace {

inline size_t max_sz(size_t a, size_t b) { return a > b ? a : b; }

int check_math(int math, const char *what) {
    if (math == FNET_MATH_BF16) {
        // refused rather than served by the fp32 kernels under another name: a bf16 mode means bf16 STORAGE of the
        // activations (half the HBM traffic of these bandwidth-bound layers), which every kernel of the library would
        // have to read and write; it is not built
        set_error("%s: FNET_MATH_BF16 is not built (use FNET_MATH_TF32 for the reduced-precision tensor-core mode)", what);
        return FNET_ERR_UNSUPPORTED;
    }
    if (math != FNET_MATH_FP32 && math != FNET_MATH_TF32 && math != FNET_MATH_TF32X3 && math != FNET_MATH_AUTO) {
        set_error("%s: unknown math mode %d", what, math);
        return FNET_ERR_ARG;
    }
    return FNET_OK;
}

// FNET_MATH_AUTO: both candidates meet the fp32 bar, so the choice is speed alone.  Measured on B200 at batch 128
// (scripts/mode_sweep.sh): the three-MMA tensor-core kernels win once the output tile is not tiny -- 1.2x to 3.7x on
// 24..192-channel convolutions and Dense layers -- and lose on the narrow ones (6 -> 12 5x5: 94 vs 56 us on the data
// gradient, 101 vs 69 us on the filter gradient; 8-wide data gradients; 16 x 16 filter gradients).
// product: 0 forward (N = K, reduction taps*C), 1 data gradient (N = C, reduction taps*K), 2 filter gradient
// (rows taps*C, N = K, reduction over the pixels).
static int resolve_math(const ConvGeom &g, int math, int product) {
    if (math != FNET_MATH_AUTO) return math;
    const long long taps = (long long)g.kh * g.kw;
    // pixels of the product: the rows of the forward / data-gradient GEMM, the reduction of the filter gradient
    const long long pixels = product == 1 ? (long long)g.B * g.H * g.W : (long long)g.B * g.Ho * g.Wo;
    bool tensor;
    if (product == 2) {
        // 1x1 layers from 16 input channels on reach the TMA-fed filter gradient (24 -> 192 on 262 144 pixels: 112 us against
        // 188 us for the FFMA kernel); the 32-row bound is for the gathered kernels, whose A tile is mostly padding below it
        tensor = g.K >= 24 && (taps * g.C >= 32 || (taps == 1 && g.C >= 16 && !(g.C & 3) && !(g.K & 3)));
        // the deep cells of a sampled candidate: a few thousand pixels at most.  The tensor-core kernels' fixed costs
        // (TMEM, barriers, the split pass, a separate reduce) are then the whole launch (scripts/conv_sweep.py, warm L2:
        // 192 -> 24 on 4096 pixels 24.4 vs 13.9 us, on 2048 pixels 23.6 vs 11.2 us, 192 -> 1536 on 256 pixels 22.7 vs 14.7 us)
        // ... unless the product is large anyway (192 -> 1536 on 4096 pixels, 1.2 G multiply-adds: 80 vs 92 us)
        if (pixels <= 4096 && pixels * taps * g.C * g.K <= (1LL << 27)) tensor = false;
    } else {
        const long long n = product == 0 ? g.K : g.C, kg = taps * (product == 0 ? g.C : g.K);
        tensor = (n >= 32 && kg >= 8) || (n >= 16 && kg >= 128) || (n >= 8 && kg >= 1024);
        // few rows, a narrow output and a short reduction: FFMA (same sweep; data gradients 192 -> 24 on 1024 pixels 16.1 vs
        // 9.9 us, 768 -> 96 on 128 pixels 24.6 vs 13.5 us; forward 192 -> 24 on 2048 pixels 13.9 vs 11.6 us).  Wide outputs
        // (192 columns from 24) and long reductions (1536) stay on the tensor cores: 11.8 vs 14.7 us, 28.3 vs 34.0 us.
        if (product == 1 && pixels <= 8192 && n <= 96 && kg <= 1024) tensor = false;
        if (product == 0 && pixels <= 4096 && n <= 32 && kg <= 1024) tensor = false;
    }
    return tensor ? FNET_MATH_TF32X3 : FNET_MATH_FP32;
}

static inline bool al16(const void *a, const void *b = nullptr, const void *c = nullptr) {
    return !((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c)) & 15);
}

int conv_fwd(const ConvGeom &g, int math_in, int act, const float *x, const float *w, const float *bias, float *y,
             void *ws, size_t ws_bytes, cudaStream_t s) {
    const int math = resolve_math(g, math_in, 0);
    int rc;
    if (strided_pointwise_ok(g) && !skinny_supported(g, 0, bias != nullptr)) {
        const size_t sub = strided_sub_bytes(g);
        if (!ws || ws_bytes < sub) { set_error("conv2d_fwd[1x1 strided]: workspace of %zu bytes needed, %zu given", sub, ws_bytes); return FNET_ERR_WORKSPACE; }
        float *x_sub = reinterpret_cast<float *>(ws);
        if ((rc = strided_subsample(g, x, x_sub, s)) != FNET_OK) return rc;
        return conv_fwd(strided_inner(g), math_in, act, x_sub, w, bias, y, reinterpret_cast<char *>(ws) + sub, ws_bytes - sub, s);
    }
    if (fewc_supported(g, 0, bias != nullptr) && al16(y)) {
        rc = fewc_fwd(g, act, x, w, bias, y, s);
        if (rc == FNET_OK && act == FNET_ACT_SOFTMAX) rc = run_softmax_rows((int64_t)g.B * g.Ho * g.Wo, g.K, y, y, s);
        return rc;
    }
    if (skinny_supported(g, 0, bias != nullptr)) rc = skinny_fwd(g, act, x, w, bias, y, ws, ws_bytes, s);
    else if (dconv_supported(g, math, 0)) rc = dconv_fwd(g, math, act, x, w, bias, y, ws, ws_bytes, s);
    else if (tma_gemm_supported(g, math, 0)) rc = tma_gemm_fwd(g, math, act, x, w, bias, y, s);
    else if (math != FNET_MATH_FP32 && tc_supported(g, math, 0)) rc = tc_conv_fwd(g, math, act, x, w, bias, y, ws, ws_bytes, s);
    else rc = simt_conv_fwd(g, act, x, w, bias, y, ws, ws_bytes, s);
    if (rc == FNET_OK && act == FNET_ACT_SOFTMAX)
        rc = run_softmax_rows((int64_t)g.B * g.Ho * g.Wo, g.K, y, y, s);
    return rc;
}

int conv_dgrad(const ConvGeom &g, int math_in, const float *dz, const float *w, float *dx, DxEpilogue ep, void *ws,
               size_t ws_bytes, cudaStream_t s) {
    const int math = resolve_math(g, math_in, 1);
    if (strided_pointwise_ok(g) && !skinny_supported(g, 1, false)) {
        const size_t sub = strided_sub_bytes(g);
        if (!ws || ws_bytes < sub) { set_error("conv2d_bwd_data[1x1 strided]: workspace of %zu bytes needed, %zu given", sub, ws_bytes); return FNET_ERR_WORKSPACE; }
        float *dx_sub = reinterpret_cast<float *>(ws);
        int rc = conv_dgrad(strided_inner(g), math_in, dz, w, dx_sub, DxEpilogue{0, 0, nullptr}, reinterpret_cast<char *>(ws) + sub,
                            ws_bytes - sub, s);
        return rc == FNET_OK ? strided_scatter(g, dx_sub, dx, ep, s) : rc;
    }
    if (fewc_supported(g, 1, false) && al16(dz)) return fewc_dgrad(g, dz, w, dx, ep, s);
    // few channels on BOTH sides: the direct kernel beats GEMM + col2im (28x28x6 -> 12, 5x5: 54 vs 112 us)
    if (skinny_supported(g, 1, false)) return skinny_dgrad(g, dz, w, dx, ep, s);
    if (size_t t_bytes = col2im_t_bytes(g)) {
        // T = dz x w^T as the data gradient of a dense layer [taps*C] -> [K], then dx = col2im(T)
        if (!ws || ws_bytes < t_bytes) {
            set_error("conv2d_bwd_data: workspace of %zu bytes needed, %zu given", t_bytes, ws_bytes);
            return FNET_ERR_WORKSPACE;
        }
        float *t = reinterpret_cast<float *>(ws);
        ConvGeom gd = dense_geom((long long)g.B * g.Ho * g.Wo, (long long)g.kh * g.kw * g.C, g.K);
        int rc = conv_dgrad(gd, math_in, dz, w, t, DxEpilogue{0, 0, nullptr}, reinterpret_cast<char *>(ws) + t_bytes, ws_bytes - t_bytes, s);
        return rc == FNET_OK ? col2im_scatter(g, t, dx, ep, s) : rc;
    }
    if (dconv_supported(g, math, 1)) return dconv_dgrad(g, math, dz, w, dx, ep, ws, ws_bytes, s);
    if (tma_gemm_supported(g, math, 1)) return tma_gemm_dgrad(g, math, dz, w, dx, ep, s);
    if (math != FNET_MATH_FP32 && tc_supported(g, math, 1)) return tc_conv_dgrad(g, math, dz, w, dx, ep, ws, ws_bytes, s);
    return simt_conv_dgrad(g, dz, w, dx, ep, ws, ws_bytes, s);
}

// Does the data gradient's epilogue run on a tensor-core kernel (where folding act'(y) into it does not pay)?
// Mirrors conv_dgrad's dispatch; the col2im path applies the epilogue in its scatter pass (cheap).
static int dgrad_on_tensor_core(const ConvGeom &g, int math_in) {
    const int math = resolve_math(g, math_in, 1);
    if (strided_pointwise_ok(g)) return 0;                   // the scatter pass carries the epilogue
    if (fewc_supported(g, 1, false)) return 0;
    // a small dx (< 16 MB) is a launch-bound layer: one launch less beats the extra read of y in the epilogue (the
    // population's candidates spent 5 % of their steps in separate 7 us act_bwd launches)
    if ((long long)g.B * g.H * g.W * g.C < (1LL << 22)) return 0;
    if (skinny_supported(g, 1, false) || col2im_t_bytes(g)) return 0;
    if (dconv_supported(g, math, 1) || tma_gemm_supported(g, math, 1)) return 1;
    return (math != FNET_MATH_FP32 && tc_supported(g, math, 1)) ? 1 : 0;
}

int conv_wgrad(const ConvGeom &g, int math_in, const float *x, const float *dz, float *dw, float *dbias, void *ws,
               size_t ws_bytes, cudaStream_t s) {
    const int math = resolve_math(g, math_in, 2);
    if (strided_pointwise_ok(g) && !skinny_supported(g, 2, dbias != nullptr)) {
        const size_t sub = strided_sub_bytes(g);
        if (!ws || ws_bytes < sub) { set_error("conv2d_bwd_filter[1x1 strided]: workspace of %zu bytes needed, %zu given", sub, ws_bytes); return FNET_ERR_WORKSPACE; }
        float *x_sub = reinterpret_cast<float *>(ws);
        int rc = strided_subsample(g, x, x_sub, s);
        return rc == FNET_OK ? conv_wgrad(strided_inner(g), math_in, x_sub, dz, dw, dbias, reinterpret_cast<char *>(ws) + sub, ws_bytes - sub, s) : rc;
    }
    if (fewc_supported(g, 2, dbias != nullptr) && al16(dz)) return fewc_wgrad(g, x, dz, dw, dbias, ws, ws_bytes, s);
    if (skinny_supported(g, 2, dbias != nullptr)) return skinny_wgrad(g, x, dz, dw, dbias, ws, ws_bytes, s);
    if (tma_gemm_supported(g, math, 2)) return tma_gemm_wgrad(g, math, x, dz, dw, dbias, ws, ws_bytes, s);
    if (math != FNET_MATH_FP32 && tc_supported(g, math, 2)) return tc_conv_wgrad(g, math, x, dz, dw, dbias, ws, ws_bytes, s);
    return simt_conv_wgrad(g, x, dz, dw, dbias, ws, ws_bytes, s);
}

int check_dense(int64_t m, int64_t k, int64_t n, const char *what) {
    if (m < 0 || k <= 0 || n <= 0 || m > 0x7fffffffLL || k > 0x7fffffffLL || n > 0x7fffffffLL) {
        set_error("%s: bad GEMM shape m=%lld k=%lld n=%lld", what, (long long)m, (long long)k, (long long)n);
        return FNET_ERR_ARG;
    }
    return FNET_OK;
}

}  // namespace

extern "C" {

size_t fnet_conv2d_workspace_bytes(const fnet_conv2d_desc *d) {
    ConvGeom g;
    if (geom_checked(d, g, "fnet_conv2d_workspace_bytes") != FNET_OK) return 0;
    size_t front = 0;
    if (strided_pointwise_ok(g)) {                           // sub(x) / dx_sub first, then whatever the stride-1 layer needs
        front = strided_sub_bytes(g);
        g = strided_inner(g);
    }
    size_t need = max_sz(max_sz(simt_conv_workspace(g), skinny_workspace(g)), fewc_workspace(g));
    if (d->math != FNET_MATH_FP32) {
        const int tensor_math = d->math == FNET_MATH_AUTO ? FNET_MATH_TF32X3 : d->math;     // covers whatever AUTO resolves to
        need = max_sz(need, max_sz(tc_conv_workspace(g, tensor_math), tma_gemm_workspace(g, tensor_math)));
        need = max_sz(need, max_sz(dconv_workspace(g, FNET_MATH_TF32X3), dconv_workspace(g, FNET_MATH_TF32)));
    }
    if (size_t t_bytes = col2im_t_bytes(g)) {
        ConvGeom gd = dense_geom((long long)g.B * g.Ho * g.Wo, (long long)g.kh * g.kw * g.C, g.K);
        need = max_sz(need, t_bytes + max_sz(simt_conv_workspace(gd), tc_conv_workspace(gd, FNET_MATH_TF32)));
    }
    return front + need;
}

int fnet_conv2d_fwd(const fnet_conv2d_desc *d, const float *x, const float *w, const float *bias, float *y, void *ws,
                    size_t ws_bytes, void *stream) {
    ConvGeom g;
    int rc = geom_checked(d, g, "fnet_conv2d_fwd");
    if (rc) return rc;
    if ((rc = check_math(d->math, "fnet_conv2d_fwd"))) return rc;
    FNET_REQUIRE(x && w && y, "fnet_conv2d_fwd: null pointer");
    FNET_REQUIRE(d->act >= FNET_ACT_NONE && d->act <= FNET_ACT_SOFTMAX, "fnet_conv2d_fwd: unknown activation");
    return conv_fwd(g, d->math, d->act, x, w, bias, y, ws, ws_bytes, as_stream(stream));
}

int fnet_conv2d_bwd_data(const fnet_conv2d_desc *d, const float *dz, const float *w, float *dx, int accumulate,
                         void *ws, size_t ws_bytes, void *stream) {
    ConvGeom g;
    int rc = geom_checked(d, g, "fnet_conv2d_bwd_data");
    if (rc) return rc;
    if ((rc = check_math(d->math, "fnet_conv2d_bwd_data"))) return rc;
    FNET_REQUIRE(dz && w && dx, "fnet_conv2d_bwd_data: null pointer");
    return conv_dgrad(g, d->math, dz, w, dx, DxEpilogue{accumulate, 0, nullptr}, ws, ws_bytes, as_stream(stream));
}

int fnet_conv2d_bwd_data_act(const fnet_conv2d_desc *d, const float *dz, const float *w, float *dx, int accumulate,
                             int act, const float *y, void *ws, size_t ws_bytes, void *stream) {
    ConvGeom g;
    int rc = geom_checked(d, g, "fnet_conv2d_bwd_data_act");
    if (rc) return rc;
    if ((rc = check_math(d->math, "fnet_conv2d_bwd_data_act"))) return rc;
    FNET_REQUIRE(dz && w && dx, "fnet_conv2d_bwd_data_act: null pointer");
    FNET_REQUIRE(act >= FNET_ACT_NONE && act <= FNET_ACT_SIGMOID && (act == FNET_ACT_NONE || y != nullptr),
                 "fnet_conv2d_bwd_data_act: relu / tanh / sigmoid with the producing layer's output, or none");
    return conv_dgrad(g, d->math, dz, w, dx, DxEpilogue{accumulate, act, y}, ws, ws_bytes, as_stream(stream));
}

int fnet_conv2d_bwd_filter(const fnet_conv2d_desc *d, const float *x, const float *dz, float *dw, float *dbias,
                           void *ws, size_t ws_bytes, void *stream) {
    ConvGeom g;
    int rc = geom_checked(d, g, "fnet_conv2d_bwd_filter");
    if (rc) return rc;
    if ((rc = check_math(d->math, "fnet_conv2d_bwd_filter"))) return rc;
    FNET_REQUIRE(x && dz && dw, "fnet_conv2d_bwd_filter: null pointer");
    return conv_wgrad(g, d->math, x, dz, dw, dbias, ws, ws_bytes, as_stream(stream));
}

int fnet_conv2d_bwd_data_on_tensor_core(const fnet_conv2d_desc *d) {
    ConvGeom g;
    if (geom_checked(d, g, "fnet_conv2d_bwd_data_on_tensor_core") != FNET_OK) return -1;
    return dgrad_on_tensor_core(g, d->math);
}

int fnet_dense_bwd_data_on_tensor_core(int64_t m, int64_t k, int64_t n, int math) {
    if (check_dense(m, k, n, "fnet_dense_bwd_data_on_tensor_core") != FNET_OK || check_math(math, "fnet_dense_bwd_data_on_tensor_core")) return -1;
    return m == 0 ? 0 : dgrad_on_tensor_core(dense_geom(m, k, n), math);
}

size_t fnet_dense_workspace_bytes(int64_t m, int64_t k, int64_t n) {
    if (check_dense(m, k, n, "fnet_dense_workspace_bytes") != FNET_OK || m == 0) return 0;
    ConvGeom g = dense_geom(m, k, n);
    return max_sz(max_sz(simt_conv_workspace(g), skinny_workspace(g)),
                  max_sz(tc_conv_workspace(g, FNET_MATH_TF32), tma_gemm_workspace(g, FNET_MATH_TF32)));
}

int fnet_dense_fwd(int64_t m, int64_t k, int64_t n, int act, int math, const float *x, const float *w,
                   const float *bias, float *y, void *ws, size_t ws_bytes, void *stream) {
    int rc = check_dense(m, k, n, "fnet_dense_fwd");
    if (rc) return rc;
    if ((rc = check_math(math, "fnet_dense_fwd"))) return rc;
    if (m == 0) return FNET_OK;
    FNET_REQUIRE(x && w && y, "fnet_dense_fwd: null pointer");
    FNET_REQUIRE(act >= FNET_ACT_NONE && act <= FNET_ACT_SOFTMAX, "fnet_dense_fwd: unknown activation");
    return conv_fwd(dense_geom(m, k, n), math, act, x, w, bias, y, ws, ws_bytes, as_stream(stream));
}

int fnet_dense_bwd_data(int64_t m, int64_t k, int64_t n, int math, const float *dz, const float *w, float *dx,
                        int accumulate, void *ws, size_t ws_bytes, void *stream) {
    int rc = check_dense(m, k, n, "fnet_dense_bwd_data");
    if (rc) return rc;
    if ((rc = check_math(math, "fnet_dense_bwd_data"))) return rc;
    if (m == 0) return FNET_OK;
    FNET_REQUIRE(dz && w && dx, "fnet_dense_bwd_data: null pointer");
    return conv_dgrad(dense_geom(m, k, n), math, dz, w, dx, DxEpilogue{accumulate, 0, nullptr}, ws, ws_bytes, as_stream(stream));
}

int fnet_dense_bwd_data_act(int64_t m, int64_t k, int64_t n, int math, const float *dz, const float *w, float *dx,
                            int accumulate, int act, const float *y, void *ws, size_t ws_bytes, void *stream) {
    int rc = check_dense(m, k, n, "fnet_dense_bwd_data_act");
    if (rc) return rc;
    if ((rc = check_math(math, "fnet_dense_bwd_data_act"))) return rc;
    if (m == 0) return FNET_OK;
    FNET_REQUIRE(dz && w && dx, "fnet_dense_bwd_data_act: null pointer");
    FNET_REQUIRE(act >= FNET_ACT_NONE && act <= FNET_ACT_SIGMOID && (act == FNET_ACT_NONE || y != nullptr),
                 "fnet_dense_bwd_data_act: relu / tanh / sigmoid with the producing layer's output, or none");
    return conv_dgrad(dense_geom(m, k, n), math, dz, w, dx, DxEpilogue{accumulate, act, y}, ws, ws_bytes, as_stream(stream));
}

int fnet_dense_bwd_filter(int64_t m, int64_t k, int64_t n, int math, const float *x, const float *dz, float *dw,
                          float *dbias, void *ws, size_t ws_bytes, void *stream) {
    int rc = check_dense(m, k, n, "fnet_dense_bwd_filter");
    if (rc) return rc;
    if ((rc = check_math(math, "fnet_dense_bwd_filter"))) return rc;
    FNET_REQUIRE(x && dz && dw, "fnet_dense_bwd_filter: null pointer");
    if (m == 0) {
        rc = fnet_fill(k * n, 0.f, dw, stream);
        if (rc == FNET_OK && dbias) rc = fnet_fill(n, 0.f, dbias, stream);
        return rc;
    }
    return conv_wgrad(dense_geom(m, k, n), math, x, dz, dw, dbias, ws, ws_bytes, as_stream(stream));
}

}  // extern "C"

// Convolutions with very few INPUT channels and a moderate number of output channels (C <= 4, 16 < K <= 64): the
// first layer of almost every sampled product (3 -> 24, 1x1 / 2x2 / 3x1, stride 1 or 2) and whatever follows a
// `Reg_` aligner back to 3 channels.  Their reduction has 3..16 terms, so as implicit GEMMs they are all tile
// overhead (38 / 53 / 27 us forward / filter / data gradient on 128 x 64 x 32 pixels for 28 MB of traffic), and the
// direct kernels of skinny.cu stop at 16 output channels because a thread holds all K sums.  Here the K outputs are
// cut into groups of eight: a thread owns (pixel, group) in the forward pass and the filter gradient, a pixel in the
// data gradient; dz / y rows are read and written 16 bytes at a time.
#include "igemm.cuh"

namespace fnet {

constexpr int kFewcRows = 16;       // taps * C of the forward pass / the data gradient
constexpr int kFewcWgradRows = 8;   // taps * C + 1 (bias row) of the filter gradient: 8 x 8 sums per thread
constexpr int kFewcBlocks = 4 * kNumSMs;

static bool fewc_geom(const ConvGeom &g) {
    static const int mode = env_switch("FNET_FEWC", 1);      // 0 disables (A/B measurements)
    return mode && g.C <= 4 && g.K > 16 && g.K <= 64 && !(g.K & 3) && g.kh * g.kw * g.C <= kFewcRows &&
           (long long)g.B * g.H * g.W < 0x7fffffffLL / 64 && (long long)g.B * g.Ho * g.Wo < 0x7fffffffLL / 64;
}

bool fewc_supported(const ConvGeom &g, int product, bool with_bias) {
    if (!fewc_geom(g)) return false;
    if (product == 2) return g.kh * g.kw * g.C + (with_bias ? 1 : 0) <= kFewcWgradRows;
    return true;
}

size_t fewc_workspace(const ConvGeom &g) {
    return fewc_supported(g, 2, true) ? (size_t)kFewcBlocks * (g.kh * g.kw * g.C + 1) * g.K * sizeof(float) : 0;
}

// a thread = (output pixel, group of eight output channels); the filter with its bias row sits in shared memory,
// rows padded to a multiple of eight columns with zeros
__global__ void __launch_bounds__(256) fewc_fwd_kernel(ConvGeom g, int act, const float *__restrict__ x, const float *__restrict__ w,
                                                       const float *__restrict__ bias, float *__restrict__ y) {
    pdl_prologue();
    extern __shared__ __align__(16) float wsm[];
    const int K = g.K, C = g.C, rows = g.kh * g.kw * C, groups = (K + 7) >> 3, K8 = groups << 3;
    for (int i = threadIdx.x; i < (rows + 1) * K8; i += 256) {
        const int r = i / K8, k = i - r * K8;
        wsm[i] = k < K ? (r < rows ? __ldg(w + r * K + k) : (bias ? __ldg(bias + k) : 0.f)) : 0.f;
    }
    __syncthreads();
    const int M = g.B * g.Ho * g.Wo, total = M * groups;
    for (int idx = blockIdx.x * 256 + threadIdx.x; idx < total; idx += gridDim.x * 256) {
        const int p = idx / groups, k0 = (idx - p * groups) << 3;
        const int wo = p % g.Wo, t = p / g.Wo, ho = t % g.Ho, b = t / g.Ho;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = wsm[rows * K8 + k0 + j];
        int row = 0;
        for (int r = 0; r < g.kh; ++r) {
            const int h = ho * g.sh - g.pt + r * g.dh;
            for (int s = 0; s < g.kw; ++s) {
                const int ww = wo * g.sw - g.pl + s * g.dw;
                const bool ok = (unsigned)h < (unsigned)g.H && (unsigned)ww < (unsigned)g.W;
                const float *px = x + ((long long)(b * g.H + (ok ? h : 0)) * g.W + (ok ? ww : 0)) * C;
                for (int c = 0; c < C; ++c, ++row) {
                    const float xv = ok ? __ldg(px + c) : 0.f;
                    const float4 w0 = *reinterpret_cast<const float4 *>(wsm + row * K8 + k0);
                    const float4 w1 = *reinterpret_cast<const float4 *>(wsm + row * K8 + k0 + 4);
                    acc[0] = fmaf(xv, w0.x, acc[0]); acc[1] = fmaf(xv, w0.y, acc[1]); acc[2] = fmaf(xv, w0.z, acc[2]); acc[3] = fmaf(xv, w0.w, acc[3]);
                    acc[4] = fmaf(xv, w1.x, acc[4]); acc[5] = fmaf(xv, w1.y, acc[5]); acc[6] = fmaf(xv, w1.z, acc[6]); acc[7] = fmaf(xv, w1.w, acc[7]);
                }
            }
        }
        float *py = y + (long long)p * K + k0;                // K is a multiple of 4: every half of a group is whole or absent
        *reinterpret_cast<float4 *>(py) = make_float4(act_apply(acc[0], act), act_apply(acc[1], act), act_apply(acc[2], act), act_apply(acc[3], act));
        if (k0 + 4 < K)
            *reinterpret_cast<float4 *>(py + 4) = make_float4(act_apply(acc[4], act), act_apply(acc[5], act), act_apply(acc[6], act), act_apply(acc[7], act));
    }
}

// a thread = an INPUT pixel; it gathers the output pixels whose windows cover it and reads their dz rows 16 bytes at a
// time; the filter sits in shared memory as [tap][c][K]
__global__ void __launch_bounds__(256) fewc_dgrad_kernel(ConvGeom g, const float *__restrict__ dz, const float *__restrict__ w, float *dx,
                                                         int accumulate, int act, const float *__restrict__ yact) {
    pdl_prologue();
    extern __shared__ __align__(16) float wsm[];
    const int K = g.K, C = g.C, taps = g.kh * g.kw;
    for (int i = threadIdx.x; i < taps * C * K; i += 256) wsm[i] = __ldg(w + i);
    __syncthreads();
    const int M = g.B * g.H * g.W;
    for (int m = blockIdx.x * 256 + threadIdx.x; m < M; m += gridDim.x * 256) {
        const int wi = m % g.W, t = m / g.W, hi = t % g.H, b = t / g.H;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int r = 0; r < g.kh; ++r) {
            const int hn = hi + g.pt - r * g.dh;
            if (hn < 0) continue;
            const int ho = hn / g.sh;
            if (ho * g.sh != hn || ho >= g.Ho) continue;
            for (int s = 0; s < g.kw; ++s) {
                const int wn = wi + g.pl - s * g.dw;
                if (wn < 0) continue;
                const int wo = wn / g.sw;
                if (wo * g.sw != wn || wo >= g.Wo) continue;
                const float4 *pz = reinterpret_cast<const float4 *>(dz + ((long long)(b * g.Ho + ho) * g.Wo + wo) * K);
                const float *wr = wsm + (r * g.kw + s) * C * K;
                for (int k4 = 0; k4 < (K >> 2); ++k4) {
                    const float4 gz = __ldg(pz + k4);
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        if (c < C) {
                            const float4 wv = *reinterpret_cast<const float4 *>(wr + c * K + (k4 << 2));
                            acc[c] = fmaf(gz.x, wv.x, fmaf(gz.y, wv.y, fmaf(gz.z, wv.z, fmaf(gz.w, wv.w, acc[c]))));
                        }
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (c < C) store_dx(dx, (long long)m * C + c, acc[c], accumulate, act, yact);
    }
}

// a thread = (its share of the output pixels, group of eight output channels): rows x 8 sums in registers, folded by the
// CTA (shuffle tree, warp slots in order) into partial[block][row][K]; conv_small_wgrad_reduce_kernel sums the blocks
template <int RT>
__global__ void __launch_bounds__(256) fewc_wgrad_kernel(ConvGeom g, int rows, const float *__restrict__ x, const float *__restrict__ dz,
                                                         float *__restrict__ partial) {
    pdl_prologue();
    __shared__ float red[8][RT * 8];
    const int K = g.K, C = g.C, taps_c = g.kh * g.kw * C, k0 = blockIdx.y << 3;
    const bool second = k0 + 4 < K;
    float acc[RT][8];
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    const int M = g.B * g.Ho * g.Wo;
    for (int m = blockIdx.x * 256 + threadIdx.x; m < M; m += gridDim.x * 256) {
        const int wo = m % g.Wo, t = m / g.Wo, ho = t % g.Ho, b = t / g.Ho;
        const float *img = x + (long long)b * g.H * g.W * C;
        const float4 g0 = __ldg(reinterpret_cast<const float4 *>(dz + (long long)m * K + k0));
        const float4 g1 = second ? __ldg(reinterpret_cast<const float4 *>(dz + (long long)m * K + k0 + 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float gz[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        int r = 0, s = 0, c = 0;
#pragma unroll
        for (int row = 0; row < RT; ++row) {
            if (row < rows) {
                float xv = 1.f;                     // the bias row
                if (row < taps_c) {
                    const int h = ho * g.sh - g.pt + r * g.dh, ww = wo * g.sw - g.pl + s * g.dw;
                    xv = ((unsigned)h < (unsigned)g.H && (unsigned)ww < (unsigned)g.W) ? __ldg(img + ((long long)h * g.W + ww) * C + c) : 0.f;
                    if (++c == C) { c = 0; if (++s == g.kw) { s = 0; ++r; } }
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[row][j] = fmaf(xv, gz[j], acc[row][j]);
            }
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float v = warp_sum(acc[i][j]);
            if (lane == 0) red[warp][i * 8 + j] = v;
        }
    __syncthreads();
    if (threadIdx.x < rows * 8) {
        const int row = threadIdx.x >> 3, j = threadIdx.x & 7;
        if (k0 + j < K) {
            float v = 0.f;
#pragma unroll
            for (int wv = 0; wv < 8; ++wv) v += red[wv][row * 8 + j];
            partial[((long long)blockIdx.x * rows + row) * K + k0 + j] = v;
        }
    }
}

static int fewc_grid(long long items) {
    long long blocks = ceil_div<long long>(items, 256);
    const long long cap = (long long)kNumSMs * 8;
    return (int)(blocks > cap ? cap : (blocks < 1 ? 1 : blocks));
}

int fewc_fwd(const ConvGeom &g, int act, const float *x, const float *w, const float *bias, float *y, cudaStream_t s) {
    if (act == FNET_ACT_SOFTMAX) act = FNET_ACT_NONE;      // the caller runs the softmax pass
    const int groups = (g.K + 7) >> 3;
    const size_t smem = (size_t)(g.kh * g.kw * g.C + 1) * (groups << 3) * sizeof(float);
    launch_pdl(fewc_fwd_kernel, dim3(fewc_grid((long long)g.B * g.Ho * g.Wo * groups)), dim3(256), smem, s, g, act, x, w, bias, y);
    return check_launch("conv2d_fwd[few input channels]");
}

int fewc_dgrad(const ConvGeom &g, const float *dz, const float *w, float *dx, DxEpilogue ep, cudaStream_t s) {
    const size_t smem = (size_t)g.kh * g.kw * g.C * g.K * sizeof(float);
    launch_pdl(fewc_dgrad_kernel, dim3(fewc_grid((long long)g.B * g.H * g.W)), dim3(256), smem, s, g, dz, w, dx, ep.accumulate, ep.act, ep.y);
    return check_launch("conv2d_bwd_data[few input channels]");
}

// defined in skinny.cu
void launch_small_wgrad_reduce(int blocks, int taps_c, int K, int total, const float *partial, float *dw, float *dbias, cudaStream_t s);

int fewc_wgrad(const ConvGeom &g, const float *x, const float *dz, float *dw, float *dbias, void *ws, size_t ws_bytes, cudaStream_t s) {
    const int taps_c = g.kh * g.kw * g.C, rows = taps_c + (dbias ? 1 : 0), total = rows * g.K;
    long long M = (long long)g.B * g.Ho * g.Wo;
    int blocks = (int)ceil_div<long long>(M, 256);
    if (blocks > kFewcBlocks) blocks = kFewcBlocks;
    if (blocks < 1) blocks = 1;
    const size_t need = (size_t)blocks * total * sizeof(float);
    if (!ws || ws_bytes < need) {
        set_error("conv2d_bwd_filter[few input channels]: workspace of %zu bytes needed, %zu given", need, ws_bytes);
        return FNET_ERR_WORKSPACE;
    }
    float *partial = reinterpret_cast<float *>(ws);
    const dim3 grid(blocks, (g.K + 7) >> 3);
    if (rows <= 4) launch_pdl(fewc_wgrad_kernel<4>, grid, dim3(256), 0, s, g, rows, x, dz, partial);
    else launch_pdl(fewc_wgrad_kernel<8>, grid, dim3(256), 0, s, g, rows, x, dz, partial);
    int rc = check_launch("conv2d_bwd_filter[few input channels]");
    if (rc) return rc;
    launch_small_wgrad_reduce(blocks, taps_c, g.K, total, partial, dw, dbias, s);
    return check_launch("conv2d_bwd_filter[few input channels]");
}

}  // namespace fnet

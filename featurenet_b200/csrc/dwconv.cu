// Depthwise convolution (depth multiplier 1), NHWC, forward / backward-data / backward-filter.
//
// Not a dense contraction: 2*kh*kw flops per element moved, so these are HBM-bound streaming
// kernels, not tensor-core work.  One thread per output (forward) or input
// (backward data) element with channels fastest => coalesced; the filter (kh*kw*C floats)
// stays in L1/L2.  The filter gradient is a two-stage deterministic reduction over pixels.
#include "common.cuh"

namespace fnet {

__global__ void __launch_bounds__(256) dw_fwd_kernel(fnet_conv2d_desc d, const float *__restrict__ x,
                                                     const float *__restrict__ w, const float *__restrict__ bias,
                                                     float *__restrict__ y, int act) {
    pdl_prologue();
    int C = d.in_c;
    int64_t total = (int64_t)d.batch * d.out_h * d.out_w * C;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int c, wo, ho;
        int64_t b;
        decode4(i, d.out_h, d.out_w, C, b, ho, wo, c);
        int h0 = ho * d.stride_h - d.pad_top, w0 = wo * d.stride_w - d.pad_left;
        const float *img = x + b * d.in_h * d.in_w * C + c;
        float acc = 0.f;
        for (int r = 0; r < d.kh; ++r) {
            int h = h0 + r;
            if ((unsigned)h >= (unsigned)d.in_h) continue;
            for (int s = 0; s < d.kw; ++s) {
                int ww = w0 + s;
                if ((unsigned)ww >= (unsigned)d.in_w) continue;
                acc = fmaf(__ldg(img + ((int64_t)h * d.in_w + ww) * C), __ldg(w + (r * d.kw + s) * C + c), acc);
            }
        }
        if (bias) acc += __ldg(bias + c);
        y[i] = act_apply(acc, act);
    }
}

__global__ void __launch_bounds__(256) dw_dgrad_kernel(fnet_conv2d_desc d, const float *__restrict__ dz,
                                                       const float *__restrict__ w, float *__restrict__ dx,
                                                       int accumulate) {
    pdl_prologue();
    int C = d.in_c;
    int64_t total = (int64_t)d.batch * d.in_h * d.in_w * C;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int c, wi, hi;
        int64_t b;
        decode4(i, d.in_h, d.in_w, C, b, hi, wi, c);
        const float *img = dz + b * d.out_h * d.out_w * C + c;
        float acc = 0.f;
        for (int r = 0; r < d.kh; ++r) {
            int hh = hi + d.pad_top - r;
            if (hh < 0) break;
            int ho = hh / d.stride_h;
            if (ho * d.stride_h != hh || ho >= d.out_h) continue;
            for (int s = 0; s < d.kw; ++s) {
                int ww = wi + d.pad_left - s;
                if (ww < 0) break;
                int wo = ww / d.stride_w;
                if (wo * d.stride_w != ww || wo >= d.out_w) continue;
                acc = fmaf(__ldg(img + ((int64_t)ho * d.out_w + wo) * C), __ldg(w + (r * d.kw + s) * C + c), acc);
            }
        }
        dx[i] = accumulate ? dx[i] + acc : acc;
    }
}

// stage 1: CTA (cgroup, split) walks its share of output pixels; thread = (channel lane, pixel lane).
// partial[split][tap or bias][c]
constexpr int kDwChan = 32;   // channels per CTA (x dimension of the thread block)
constexpr int kDwPix = 8;     // pixel lanes per CTA (y dimension)

template <int MAXROWS>
__global__ void __launch_bounds__(kDwChan * kDwPix) dw_wgrad_partial_kernel(fnet_conv2d_desc d, int pix_per_split,
                                                                           const float *__restrict__ x,
                                                                           const float *__restrict__ dz,
                                                                           float *__restrict__ partial, int rows) {
    pdl_prologue();
    __shared__ float red[kDwPix][kDwChan + 1];
    int C = d.in_c;
    int c = blockIdx.x * kDwChan + threadIdx.x;
    int split = blockIdx.y;
    int64_t npix = (int64_t)d.batch * d.out_h * d.out_w;
    int64_t p0 = (int64_t)split * pix_per_split;
    int64_t p1 = p0 + pix_per_split < npix ? p0 + pix_per_split : npix;
    int taps = d.kh * d.kw;
    float acc[MAXROWS];                 // one accumulator per tap (+ bias): x and dz are read once
#pragma unroll
    for (int t = 0; t < MAXROWS; ++t) acc[t] = 0.f;
    if (c < C) {
        for (int64_t p = p0 + threadIdx.y; p < p1; p += kDwPix) {
            float g = __ldg(dz + p * C + c);
            int wo = (int)(p % d.out_w);
            int64_t q = p / d.out_w;
            int ho = (int)(q % d.out_h);
            int64_t b = q / d.out_h;
            int h0 = ho * d.stride_h - d.pad_top, w0 = wo * d.stride_w - d.pad_left;
            const float *img = x + b * d.in_h * d.in_w * C + c;
            int r = 0, s = 0;
#pragma unroll
            for (int t = 0; t < MAXROWS; ++t) {
                if (t < taps) {
                    int h = h0 + r, ww = w0 + s;
                    if ((unsigned)h < (unsigned)d.in_h && (unsigned)ww < (unsigned)d.in_w)
                        acc[t] = fmaf(g, __ldg(img + ((int64_t)h * d.in_w + ww) * C), acc[t]);
                    if (++s == d.kw) { s = 0; ++r; }
                } else if (t == taps) {
                    acc[t] += g;            // bias row
                }
            }
        }
    }
#pragma unroll
    for (int t = 0; t < MAXROWS; ++t) {
        if (t < rows) {
            red[threadIdx.y][threadIdx.x] = acc[t];
            __syncthreads();
            if (threadIdx.y == 0 && c < C) {
                float sum = 0.f;
#pragma unroll
                for (int j = 0; j < kDwPix; ++j) sum += red[j][threadIdx.x];
                partial[((int64_t)split * rows + t) * C + c] = sum;
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(256) dw_wgrad_reduce_kernel(int splits, int rows, int taps, int C,
                                                              const float *__restrict__ partial, float *__restrict__ dw,
                                                              float *__restrict__ dbias) {
    pdl_prologue();
    int total = rows * C;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int z = 0; z < splits; ++z) s += partial[(int64_t)z * total + i];
        int row = i / C, c = i - row * C;
        if (row < taps) dw[i] = s;
        else dbias[c] = s;
    }
}

static int dw_splits(const fnet_conv2d_desc *d) {
    int64_t npix = (int64_t)d->batch * d->out_h * d->out_w;
    int cgroups = ceil_div(d->in_c, kDwChan);
    int64_t want = ceil_div(4 * kNumSMs, cgroups);
    int64_t most = ceil_div<int64_t>(npix, 4 * kDwPix);
    if (want > most) want = most;
    if (want < 1) want = 1;
    if (want > 1024) want = 1024;
    return (int)want;
}

static int check_dw(const fnet_conv2d_desc *d, const char *what) {
    if (!d) { set_error("%s: null descriptor", what); return FNET_ERR_ARG; }
    if (d->batch < 0 || d->in_h <= 0 || d->in_w <= 0 || d->in_c <= 0 || d->out_h <= 0 || d->out_w <= 0 ||
        d->out_c != d->in_c || d->kh <= 0 || d->kw <= 0 || d->stride_h <= 0 || d->stride_w <= 0 || d->pad_top < 0 ||
        d->pad_left < 0 || d->dil_h != 1 || d->dil_w != 1) {
        set_error("%s: bad depthwise geometry (out_c must equal in_c, dilation 1)", what);
        return FNET_ERR_ARG;
    }
    return FNET_OK;
}

}  // namespace fnet

using namespace fnet;

extern "C" {

size_t fnet_dwconv2d_workspace_bytes(const fnet_conv2d_desc *d) {
    if (!d || d->in_c <= 0) return 0;
    return (size_t)dw_splits(d) * (d->kh * d->kw + 1) * d->in_c * sizeof(float);
}

int fnet_dwconv2d_fwd(const fnet_conv2d_desc *d, const float *x, const float *w, const float *bias, float *y,
                      void *stream) {
    int rc = check_dw(d, "fnet_dwconv2d_fwd");
    if (rc) return rc;
    FNET_REQUIRE(x && w && y, "fnet_dwconv2d_fwd: null pointer");
    int64_t total = (int64_t)d->batch * d->out_h * d->out_w * d->in_c;
    if (total == 0) return FNET_OK;
    int act = d->act == FNET_ACT_SOFTMAX ? FNET_ACT_NONE : d->act;
    launch_pdl(dw_fwd_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), *d, x, w, bias, y, act);
    rc = check_launch("fnet_dwconv2d_fwd");
    if (rc == FNET_OK && d->act == FNET_ACT_SOFTMAX)
        rc = fnet_act_fwd(FNET_ACT_SOFTMAX, total / d->in_c, d->in_c, y, y, stream);
    return rc;
}

int fnet_dwconv2d_bwd_data(const fnet_conv2d_desc *d, const float *dz, const float *w, float *dx, int accumulate,
                           void *stream) {
    int rc = check_dw(d, "fnet_dwconv2d_bwd_data");
    if (rc) return rc;
    FNET_REQUIRE(dz && w && dx, "fnet_dwconv2d_bwd_data: null pointer");
    int64_t total = (int64_t)d->batch * d->in_h * d->in_w * d->in_c;
    if (total == 0) return FNET_OK;
    launch_pdl(dw_dgrad_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), *d, dz, w, dx, accumulate);
    return check_launch("fnet_dwconv2d_bwd_data");
}

int fnet_dwconv2d_bwd_filter(const fnet_conv2d_desc *d, const float *x, const float *dz, float *dw, float *dbias,
                             void *ws, size_t ws_bytes, void *stream) {
    int rc = check_dw(d, "fnet_dwconv2d_bwd_filter");
    if (rc) return rc;
    FNET_REQUIRE(x && dz && dw, "fnet_dwconv2d_bwd_filter: null pointer");
    size_t need = fnet_dwconv2d_workspace_bytes(d);
    if (!ws || ws_bytes < need) {
        set_error("fnet_dwconv2d_bwd_filter: workspace of %zu bytes needed, %zu given", need, ws_bytes);
        return FNET_ERR_WORKSPACE;
    }
    int splits = dw_splits(d);
    int taps = d->kh * d->kw;
    int rows = taps + (dbias ? 1 : 0);
    int64_t npix = (int64_t)d->batch * d->out_h * d->out_w;
    if (npix == 0) return fnet_fill((int64_t)taps * d->in_c, 0.f, dw, stream);
    int pix_per_split = (int)ceil_div<int64_t>(npix, splits);
    splits = (int)ceil_div<int64_t>(npix, pix_per_split);
    float *partial = reinterpret_cast<float *>(ws);
    dim3 grid(ceil_div(d->in_c, kDwChan), splits), block(kDwChan, kDwPix);
    if (taps + 1 <= 10)
        launch_pdl(dw_wgrad_partial_kernel<10>, dim3(grid), dim3(block), 0, as_stream(stream), *d, pix_per_split, x, dz, partial, rows);
    else if (taps + 1 <= 26)
        launch_pdl(dw_wgrad_partial_kernel<26>, dim3(grid), dim3(block), 0, as_stream(stream), *d, pix_per_split, x, dz, partial, rows);
    else if (taps + 1 <= 50)
        launch_pdl(dw_wgrad_partial_kernel<50>, dim3(grid), dim3(block), 0, as_stream(stream), *d, pix_per_split, x, dz, partial, rows);
    else {
        set_error("fnet_dwconv2d_bwd_filter: kernels above 7x7 are outside the FeatureNet vocabulary");
        return FNET_ERR_UNSUPPORTED;
    }
    rc = check_launch("fnet_dwconv2d_bwd_filter");
    if (rc) return rc;
    launch_pdl(dw_wgrad_reduce_kernel, dim3(stream_grid((int64_t)rows * d->in_c, 256)), dim3(256), 0, as_stream(stream), 
        splits, rows, taps, d->in_c, partial, dw, dbias);
    return check_launch("fnet_dwconv2d_bwd_filter");
}

}  // extern "C"

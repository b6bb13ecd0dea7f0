// Max / average / global-average pooling, NHWC, forward and backward.
// One thread per output (forward) or per input (backward) element with the channel index
// fastest, so every warp access is a contiguous run of channels.  The backward is a gather
// (each input element visits the <= kh*kw windows that contain it): no atomics, so the
// result is deterministic.  HBM-bound: forward reads in + writes out (+1 byte/elem argmax).
#include "common.cuh"

namespace fnet {

// I = index type: 32-bit whenever the tensor has fewer than 2^31 elements (always, in practice) -- 64-bit
// division is a long software routine and these kernels do four of them per element to decode an index.
template <typename I>
__global__ void __launch_bounds__(256) pool_fwd_kernel(fnet_pool2d_desc d, const float *__restrict__ x,
                                                       float *__restrict__ y, uint8_t *__restrict__ argmax) {
    pdl_prologue();
    I total = (I)d.batch * d.out_h * d.out_w * d.channels;
    for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (I)gridDim.x * blockDim.x) {
        int c = (int)(i % (I)d.channels);
        I t = i / (I)d.channels;
        int wo = (int)(t % (I)d.out_w);
        t /= (I)d.out_w;
        int ho = (int)(t % (I)d.out_h);
        I b = t / (I)d.out_h;
        int h0 = ho * d.stride_h - d.pad_top, w0 = wo * d.stride_w - d.pad_left;
        const float *img = x + (int64_t)b * d.in_h * d.in_w * d.channels + c;
        if (d.kind == FNET_POOL_MAX) {
            float best = -INFINITY;
            int arg = 0;
            for (int r = 0; r < d.kh; ++r) {
                int h = h0 + r;
                if ((unsigned)h >= (unsigned)d.in_h) continue;
                for (int s = 0; s < d.kw; ++s) {
                    int w = w0 + s;
                    if ((unsigned)w >= (unsigned)d.in_w) continue;
                    float v = __ldg(img + (h * d.in_w + w) * d.channels);
                    if (v > best) { best = v; arg = r * d.kw + s; }   // strict: first maximum wins
                }
            }
            y[i] = best;
            if (argmax) argmax[i] = (uint8_t)arg;
        } else {
            float sum = 0.f;
            int cnt = 0;
            for (int r = 0; r < d.kh; ++r) {
                int h = h0 + r;
                if ((unsigned)h >= (unsigned)d.in_h) continue;
                for (int s = 0; s < d.kw; ++s) {
                    int w = w0 + s;
                    if ((unsigned)w >= (unsigned)d.in_w) continue;
                    sum += __ldg(img + (h * d.in_w + w) * d.channels);
                    ++cnt;
                }
            }
            y[i] = sum / (float)cnt;
        }
    }
}

__device__ __forceinline__ int taps_in_bounds(int o, int stride, int pad, int k, int extent) {
    int lo = o * stride - pad, hi = lo + k;
    if (lo < 0) lo = 0;
    if (hi > extent) hi = extent;
    return hi - lo;
}

// v / stride when exact and below extent, else -1 (strides are 1 or 2 in practice: no division)
__device__ __forceinline__ int pool_exact_div(int v, int stride, int extent) {
    int o;
    if (stride == 1) o = v;
    else if (stride == 2) { if (v & 1) return -1; o = v >> 1; }
    else { o = v / stride; if (o * stride != v) return -1; }
    return o < extent ? o : -1;
}

template <typename I>
__global__ void __launch_bounds__(256) pool_bwd_kernel(fnet_pool2d_desc d, const float *__restrict__ dy,
                                                       const uint8_t *__restrict__ argmax, float *__restrict__ dx,
                                                       int accumulate) {
    pdl_prologue();
    I total = (I)d.batch * d.in_h * d.in_w * d.channels;
    for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (I)gridDim.x * blockDim.x) {
        int c = (int)(i % (I)d.channels);
        I t = i / (I)d.channels;
        int w = (int)(t % (I)d.in_w);
        t /= (I)d.in_w;
        int h = (int)(t % (I)d.in_h);
        I b = t / (I)d.in_h;
        I out_img = b * d.out_h * d.out_w;
        float g = 0.f;
        for (int r = 0; r < d.kh; ++r) {
            int hh = h + d.pad_top - r;
            if (hh < 0) break;
            int ho = pool_exact_div(hh, d.stride_h, d.out_h);
            if (ho < 0) continue;
            for (int s = 0; s < d.kw; ++s) {
                int ww = w + d.pad_left - s;
                if (ww < 0) break;
                int wo = pool_exact_div(ww, d.stride_w, d.out_w);
                if (wo < 0) continue;
                I o = ((out_img + (I)ho * d.out_w + wo) * d.channels) + c;
                if (d.kind == FNET_POOL_MAX) {
                    if (argmax[o] == (uint8_t)(r * d.kw + s)) g += __ldg(dy + o);
                } else {
                    int cnt = taps_in_bounds(ho, d.stride_h, d.pad_top, d.kh, d.in_h) *
                              taps_in_bounds(wo, d.stride_w, d.pad_left, d.kw, d.in_w);
                    g += __ldg(dy + o) / (float)cnt;
                }
            }
        }
        dx[i] = accumulate ? dx[i] + g : g;
    }
}

// one warp per (b, 32-channel group): lanes walk channels, the warp walks the hw pixels
__global__ void __launch_bounds__(256) gap_fwd_kernel(int batch, int hw, int channels, const float *__restrict__ x,
                                                      float *__restrict__ y) {
    pdl_prologue();
    int64_t total = (int64_t)batch * channels;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int c = (int)(i % channels);
        int64_t b = i / channels;
        const float *p = x + b * hw * channels + c;
        float s = 0.f;
        for (int j = 0; j < hw; ++j) s += __ldg(p + (int64_t)j * channels);
        y[i] = s / (float)hw;
    }
}

__global__ void __launch_bounds__(256) gap_bwd_kernel(int batch, int hw, int channels, const float *__restrict__ dy,
                                                      float *__restrict__ dx, int accumulate) {
    pdl_prologue();
    int64_t total = (int64_t)batch * hw * channels;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int c = (int)(i % channels);
        int64_t b = i / ((int64_t)hw * channels);
        float g = __ldg(dy + b * channels + c) / (float)hw;
        dx[i] = accumulate ? dx[i] + g : g;
    }
}

static int check_pool(const fnet_pool2d_desc *d, const char *what) {
    if (!d) { set_error("%s: null descriptor", what); return FNET_ERR_ARG; }
    if (d->batch < 0 || d->in_h <= 0 || d->in_w <= 0 || d->channels <= 0 || d->out_h <= 0 || d->out_w <= 0 ||
        d->kh <= 0 || d->kw <= 0 || d->stride_h <= 0 || d->stride_w <= 0 || d->pad_top < 0 || d->pad_left < 0 ||
        d->kh * d->kw > 255 || (d->kind != FNET_POOL_MAX && d->kind != FNET_POOL_AVG)) {
        set_error("%s: bad pooling geometry", what);
        return FNET_ERR_ARG;
    }
    // every window needs at least one in-bounds tap (true for TF SAME/VALID geometry)
    if ((d->out_h - 1) * d->stride_h - d->pad_top >= d->in_h || (d->out_w - 1) * d->stride_w - d->pad_left >= d->in_w ||
        d->pad_top >= d->kh || d->pad_left >= d->kw) {
        set_error("%s: a pooling window lies entirely in the padding", what);
        return FNET_ERR_ARG;
    }
    return FNET_OK;
}

}  // namespace fnet

using namespace fnet;

extern "C" {

int fnet_pool2d_fwd(const fnet_pool2d_desc *d, const float *x, float *y, uint8_t *argmax, void *stream) {
    int rc = check_pool(d, "fnet_pool2d_fwd");
    if (rc) return rc;
    FNET_REQUIRE(x && y, "fnet_pool2d_fwd: null pointer");
    int64_t total = (int64_t)d->batch * d->out_h * d->out_w * d->channels;
    if (total == 0) return FNET_OK;
    int64_t in_total = (int64_t)d->batch * d->in_h * d->in_w * d->channels;
    if (total < 0x7fffffffLL - (1 << 26) && in_total < 0x7fffffffLL)      // the grid-stride increment must not overflow either
        launch_pdl(pool_fwd_kernel<int>, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), *d, x, y, argmax);
    else
        launch_pdl(pool_fwd_kernel<int64_t>, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), *d, x, y, argmax);
    return check_launch("fnet_pool2d_fwd");
}

int fnet_pool2d_bwd(const fnet_pool2d_desc *d, const float *dy, const uint8_t *argmax, float *dx, int accumulate,
                    void *stream) {
    int rc = check_pool(d, "fnet_pool2d_bwd");
    if (rc) return rc;
    FNET_REQUIRE(dy && dx, "fnet_pool2d_bwd: null pointer");
    FNET_REQUIRE(d->kind != FNET_POOL_MAX || argmax != nullptr, "fnet_pool2d_bwd: max pooling needs the argmax of the forward");
    int64_t total = (int64_t)d->batch * d->in_h * d->in_w * d->channels;
    if (total == 0) return FNET_OK;
    if (total < 0x7fffffffLL - (1 << 26))
        launch_pdl(pool_bwd_kernel<int>, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), *d, dy, argmax, dx, accumulate);
    else
        launch_pdl(pool_bwd_kernel<int64_t>, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), *d, dy, argmax, dx, accumulate);
    return check_launch("fnet_pool2d_bwd");
}

int fnet_gap_fwd(int32_t batch, int32_t hw, int32_t channels, const float *x, float *y, void *stream) {
    FNET_REQUIRE(batch >= 0 && hw > 0 && channels > 0 && x && y, "fnet_gap_fwd: bad arguments");
    if (batch == 0) return FNET_OK;
    launch_pdl(gap_fwd_kernel, dim3(stream_grid((int64_t)batch * channels, 256)), dim3(256), 0, as_stream(stream), batch, hw, channels, x, y);
    return check_launch("fnet_gap_fwd");
}

int fnet_gap_bwd(int32_t batch, int32_t hw, int32_t channels, const float *dy, float *dx, int accumulate,
                 void *stream) {
    FNET_REQUIRE(batch >= 0 && hw > 0 && channels > 0 && dy && dx, "fnet_gap_bwd: bad arguments");
    if (batch == 0) return FNET_OK;
    launch_pdl(gap_bwd_kernel, dim3(stream_grid((int64_t)batch * hw * channels, 256)), dim3(256), 0, as_stream(stream), 
        batch, hw, channels, dy, dx, accumulate);
    return check_launch("fnet_gap_bwd");
}

}  // extern "C"

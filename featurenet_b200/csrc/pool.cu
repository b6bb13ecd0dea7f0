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

// Four channels per thread (channels % 4 == 0, 16-byte aligned tensors): the index decode and the window loop are
// paid once per 16 bytes instead of once per 4.  The scalar kernel above spends ~250 instructions per element and is
// issue-bound at a tenth of the HBM rate (2x2 stride-1 SAME pooling of 128 x 32 x 32 x 32: 65 us for 38 MB).
// KH, KW > 0: window known at compile time (2x2 and 3x3, the sizes of the NAS space): loops unroll, loads are issued
// together with a clamped address.  0: runtime window.
template <int KH, int KW>
__global__ void __launch_bounds__(256) pool_fwd_vec4_kernel(fnet_pool2d_desc d, const float4 *__restrict__ x,
                                                            float4 *__restrict__ y, uchar4 *__restrict__ argmax) {
    pdl_prologue();
    if (KH > 0) { d.kh = KH; d.kw = KW; }
    const int quads = d.channels >> 2;
    const int total = d.batch * d.out_h * d.out_w * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        int wo = t % d.out_w;
        t /= d.out_w;
        int ho = t % d.out_h, b = t / d.out_h;
        const int h0 = ho * d.stride_h - d.pad_top, w0 = wo * d.stride_w - d.pad_left;
        const float4 *img = x + (int64_t)b * d.in_h * d.in_w * quads + q;
        if (d.kind == FNET_POOL_MAX) {
            float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
            uchar4 arg = make_uchar4(0, 0, 0, 0);
            if (KH > 0) {
                constexpr int T = KH * KW > 0 ? KH * KW : 1;
                float4 v[T];
                bool ok[T];
#pragma unroll
                for (int t2 = 0; t2 < T; ++t2) {
                    const int h = h0 + t2 / (KW > 0 ? KW : 1), w = w0 + t2 % (KW > 0 ? KW : 1);
                    ok[t2] = (unsigned)h < (unsigned)d.in_h && (unsigned)w < (unsigned)d.in_w;
                    v[t2] = __ldg(img + (ok[t2] ? (h * d.in_w + w) * quads : 0));
                }
#pragma unroll
                for (int t2 = 0; t2 < T; ++t2)
                    if (ok[t2]) {
                        if (v[t2].x > best.x) { best.x = v[t2].x; arg.x = (unsigned char)t2; }
                        if (v[t2].y > best.y) { best.y = v[t2].y; arg.y = (unsigned char)t2; }
                        if (v[t2].z > best.z) { best.z = v[t2].z; arg.z = (unsigned char)t2; }
                        if (v[t2].w > best.w) { best.w = v[t2].w; arg.w = (unsigned char)t2; }
                    }
            } else
            for (int r = 0; r < d.kh; ++r) {
                const int h = h0 + r;
                if ((unsigned)h >= (unsigned)d.in_h) continue;
                for (int s = 0; s < d.kw; ++s) {
                    const int w = w0 + s;
                    if ((unsigned)w >= (unsigned)d.in_w) continue;
                    const float4 v = __ldg(img + (h * d.in_w + w) * quads);
                    const unsigned char tap = (unsigned char)(r * d.kw + s);
                    if (v.x > best.x) { best.x = v.x; arg.x = tap; }       // strict: first maximum wins
                    if (v.y > best.y) { best.y = v.y; arg.y = tap; }
                    if (v.z > best.z) { best.z = v.z; arg.z = tap; }
                    if (v.w > best.w) { best.w = v.w; arg.w = tap; }
                }
            }
            y[i] = best;
            if (argmax) argmax[i] = arg;
        } else {
            float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
            int cnt = 0;
            for (int r = 0; r < d.kh; ++r) {
                const int h = h0 + r;
                if ((unsigned)h >= (unsigned)d.in_h) continue;
                for (int s = 0; s < d.kw; ++s) {
                    const int w = w0 + s;
                    if ((unsigned)w >= (unsigned)d.in_w) continue;
                    const float4 v = __ldg(img + (h * d.in_w + w) * quads);
                    sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
                    ++cnt;
                }
            }
            const float n = (float)cnt;
            y[i] = make_float4(sum.x / n, sum.y / n, sum.z / n, sum.w / n);
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

__global__ void __launch_bounds__(256) pool_bwd_vec4_kernel(fnet_pool2d_desc d, const float4 *__restrict__ dy,
                                                            const uchar4 *__restrict__ argmax, float4 *__restrict__ dx,
                                                            int accumulate) {
    pdl_prologue();
    const int quads = d.channels >> 2;
    const int total = d.batch * d.in_h * d.in_w * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        int w = t % d.in_w;
        t /= d.in_w;
        int h = t % d.in_h, b = t / d.in_h;
        const int out_img = b * d.out_h * d.out_w;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < d.kh; ++r) {
            const int hh = h + d.pad_top - r;
            if (hh < 0) break;
            const int ho = pool_exact_div(hh, d.stride_h, d.out_h);
            if (ho < 0) continue;
            for (int s = 0; s < d.kw; ++s) {
                const int ww = w + d.pad_left - s;
                if (ww < 0) break;
                const int wo = pool_exact_div(ww, d.stride_w, d.out_w);
                if (wo < 0) continue;
                const int o = (out_img + ho * d.out_w + wo) * quads + q;
                const float4 v = __ldg(dy + o);
                if (d.kind == FNET_POOL_MAX) {
                    const uchar4 a = argmax[o];
                    const unsigned char tap = (unsigned char)(r * d.kw + s);
                    if (a.x == tap) g.x += v.x;
                    if (a.y == tap) g.y += v.y;
                    if (a.z == tap) g.z += v.z;
                    if (a.w == tap) g.w += v.w;
                } else {
                    const float n = (float)(taps_in_bounds(ho, d.stride_h, d.pad_top, d.kh, d.in_h) *
                                            taps_in_bounds(wo, d.stride_w, d.pad_left, d.kw, d.in_w));
                    g.x += v.x / n; g.y += v.y / n; g.z += v.z / n; g.w += v.w / n;
                }
            }
        }
        if (accumulate) {
            const float4 old = dx[i];
            g.x += old.x; g.y += old.y; g.z += old.z; g.w += old.w;
        }
        dx[i] = g;
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
    const bool aligned = !((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) &&
                         !(reinterpret_cast<uintptr_t>(argmax) & 3);
    if (!(d->channels & 3) && aligned && total < 0x7fffffffLL - (1 << 26) && in_total < 0x7fffffffLL)
    {
        const float4 *x4 = reinterpret_cast<const float4 *>(x);
        float4 *y4 = reinterpret_cast<float4 *>(y);
        uchar4 *a4 = reinterpret_cast<uchar4 *>(argmax);
        const dim3 grid(stream_grid(total / 4, 256));
        if (d->kh == 2 && d->kw == 2) launch_pdl(pool_fwd_vec4_kernel<2, 2>, grid, dim3(256), 0, as_stream(stream), *d, x4, y4, a4);
        else if (d->kh == 3 && d->kw == 3) launch_pdl(pool_fwd_vec4_kernel<3, 3>, grid, dim3(256), 0, as_stream(stream), *d, x4, y4, a4);
        else launch_pdl(pool_fwd_vec4_kernel<0, 0>, grid, dim3(256), 0, as_stream(stream), *d, x4, y4, a4);
    }
    else if (total < 0x7fffffffLL - (1 << 26) && in_total < 0x7fffffffLL)      // the grid-stride increment must not overflow either
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
    const int64_t out_total = (int64_t)d->batch * d->out_h * d->out_w * d->channels;
    const bool aligned = !((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx)) & 15) &&
                         !(reinterpret_cast<uintptr_t>(argmax) & 3);
    if (!(d->channels & 3) && aligned && total < 0x7fffffffLL - (1 << 26) && out_total < 0x7fffffffLL)
        launch_pdl(pool_bwd_vec4_kernel, dim3(stream_grid(total / 4, 256)), dim3(256), 0, as_stream(stream), *d,
                   reinterpret_cast<const float4 *>(dy), reinterpret_cast<const uchar4 *>(argmax), reinterpret_cast<float4 *>(dx), accumulate);
    else if (total < 0x7fffffffLL - (1 << 26))
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

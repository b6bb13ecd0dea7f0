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

// ---------------------------------------------------------------------------------------------------------------
// Strip kernels for the window sizes of the NAS space (2x2 and 3x3, stride 1 or 2, channels % 4 == 0): one thread
// produces WP = 4 neighbouring pixels of one row for four channels.  The index decode is paid once per 64 bytes of
// output, and a row of the window is loaded once for all four pixels ((WP - 1) * S + KW loads instead of WP * KW:
// 5 instead of 8 for 2x2 stride 1, 6 instead of 12 for 3x3 stride 1).  Same arithmetic in the same order as the
// kernels above, so results (and argmax ties) are identical.
constexpr int POOL_WP = 4;

template <int KH, int KW, int S, bool MAX>
__global__ void __launch_bounds__(256, 4) pool_fwd_strip_kernel(fnet_pool2d_desc d, const float4 *__restrict__ x,
                                                             float4 *__restrict__ y, uchar4 *__restrict__ argmax) {
    pdl_prologue();
    constexpr int WP = POOL_WP, NCOL = (WP - 1) * S + KW;
    const int quads = d.channels >> 2, strips = (d.out_w + WP - 1) / WP;
    const int total = d.batch * d.out_h * strips * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        const int st = t % strips;
        t /= strips;
        const int ho = t % d.out_h, b = t / d.out_h;
        const int wo0 = st * WP, h0 = ho * S - d.pad_top, w0 = wo0 * S - d.pad_left;
        const float4 *img = x + (int64_t)b * d.in_h * d.in_w * quads + q;
        const float fill = MAX ? -INFINITY : 0.f;
        float4 acc[WP];
        uchar4 arg[WP];
#pragma unroll
        for (int p = 0; p < WP; ++p) { acc[p] = make_float4(fill, fill, fill, fill); arg[p] = make_uchar4(0, 0, 0, 0); }
#pragma unroll(KH * NCOL <= 12 ? KH : 1)                    // both rows of a 2x2 window in flight together: ten loads per thread
        for (int r = 0; r < KH; ++r) {
            const int h = h0 + r;
            if ((unsigned)h >= (unsigned)d.in_h) continue;
            float4 v[NCOL];
#pragma unroll
            for (int c = 0; c < NCOL; ++c) {
                const int w = w0 + c;
                const bool ok = (unsigned)w < (unsigned)d.in_w;
                v[c] = __ldg(img + (h * d.in_w + (ok ? w : 0)) * quads);
                if (!ok) v[c] = make_float4(fill, fill, fill, fill);
            }
#pragma unroll
            for (int p = 0; p < WP; ++p) {
#pragma unroll
                for (int s2 = 0; s2 < KW; ++s2) {
                    const float4 u = v[p * S + s2];
                    if (MAX) {
                        const unsigned char tap = (unsigned char)(r * KW + s2);
                        if (u.x > acc[p].x) { acc[p].x = u.x; arg[p].x = tap; }     // strict: first maximum wins
                        if (u.y > acc[p].y) { acc[p].y = u.y; arg[p].y = tap; }
                        if (u.z > acc[p].z) { acc[p].z = u.z; arg[p].z = tap; }
                        if (u.w > acc[p].w) { acc[p].w = u.w; arg[p].w = tap; }
                    } else {
                        acc[p].x += u.x; acc[p].y += u.y; acc[p].z += u.z; acc[p].w += u.w;
                    }
                }
            }
        }
        const int rows_in = MAX ? 1 : taps_in_bounds(ho, S, d.pad_top, KH, d.in_h);
        const int out = ((b * d.out_h + ho) * d.out_w + wo0) * quads + q;
#pragma unroll
        for (int p = 0; p < WP; ++p) {
            if (wo0 + p < d.out_w) {
                if (MAX) {
                    y[out + p * quads] = acc[p];
                    if (argmax) argmax[out + p * quads] = arg[p];
                } else {
                    const float n = (float)(rows_in * taps_in_bounds(wo0 + p, S, d.pad_left, KW, d.in_w));
                    y[out + p * quads] = make_float4(acc[p].x / n, acc[p].y / n, acc[p].z / n, acc[p].w / n);
                }
            }
        }
    }
}

// backward: one thread gathers the gradient of WP = 4 neighbouring input pixels of one row; a row of dy (and of the
// argmax bytes) that reaches the strip is loaded once for all four
template <int KH, int KW, int S, bool MAX>
__global__ void __launch_bounds__(256, 4) pool_bwd_strip_kernel(fnet_pool2d_desc d, const float4 *__restrict__ dy,
                                                             const uchar4 *__restrict__ argmax, float4 *__restrict__ dx,
                                                             int accumulate) {
    pdl_prologue();
    constexpr int WP = POOL_WP, NC = (WP + KW - 2) / S + 1;      // output columns whose windows can touch the strip
    const int quads = d.channels >> 2, strips = (d.in_w + WP - 1) / WP;
    const int total = d.batch * d.in_h * strips * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        const int st = t % strips;
        t /= strips;
        const int h = t % d.in_h, b = t / d.in_h;
        const int w0 = st * WP;
        // first output column that can reach pixel w0 through its last tap: ceil((w0 + pad_left - (KW - 1)) / S), may be < 0
        const int lo = w0 + d.pad_left - (KW - 1);
        const int wo_base = (lo + S * 8 + S - 1) / S - 8;
        const int off = w0 + d.pad_left - wo_base * S;            // tap of pixel w0 in window wo_base
        float4 g[WP];
#pragma unroll
        for (int p = 0; p < WP; ++p) g[p] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
        for (int r = 0; r < KH; ++r) {
            const int hh = h + d.pad_top - r;
            if (hh < 0) break;
            const int ho = pool_exact_div(hh, S, d.out_h);
            if (ho < 0) continue;
            float4 v[NC];
            uchar4 a[NC];
            const int row = (b * d.out_h + ho) * d.out_w;
            const float rows_in = MAX ? 1.f : (float)taps_in_bounds(ho, S, d.pad_top, KH, d.in_h);
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const int wo = wo_base + c;
                const bool ok = (unsigned)wo < (unsigned)d.out_w;
                const int o = (row + (ok ? wo : 0)) * quads + q;
                v[c] = __ldg(dy + o);
                if (MAX) a[c] = argmax[o];
                else {
                    const float n = rows_in * (float)taps_in_bounds(ok ? wo : 0, S, d.pad_left, KW, d.in_w);
                    v[c] = make_float4(v[c].x / n, v[c].y / n, v[c].z / n, v[c].w / n);
                }
                if (!ok) { v[c] = make_float4(0.f, 0.f, 0.f, 0.f); if (MAX) a[c] = make_uchar4(255, 255, 255, 255); }
            }
#pragma unroll
            for (int p = 0; p < WP; ++p) {
#pragma unroll
                for (int c = NC - 1; c >= 0; --c) {               // descending window = ascending tap: the order of the kernels above
                    const int s2 = off + p - c * S;
                    if (s2 >= 0 && s2 < KW) {
                        if (MAX) {
                            const unsigned char tap = (unsigned char)(r * KW + s2);
                            if (a[c].x == tap) g[p].x += v[c].x;
                            if (a[c].y == tap) g[p].y += v[c].y;
                            if (a[c].z == tap) g[p].z += v[c].z;
                            if (a[c].w == tap) g[p].w += v[c].w;
                        } else {
                            g[p].x += v[c].x; g[p].y += v[c].y; g[p].z += v[c].z; g[p].w += v[c].w;
                        }
                    }
                }
            }
        }
        const int at = ((b * d.in_h + h) * d.in_w + w0) * quads + q;
#pragma unroll
        for (int p = 0; p < WP; ++p) {
            if (w0 + p < d.in_w) {
                float4 o = g[p];
                if (accumulate) {
                    const float4 old = dx[at + p * quads];
                    o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                }
                dx[at + p * quads] = o;
            }
        }
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
        // forward: the strips pay with stride 1 only (2x2 stride 2 on 128 x 32 x 32 x 64: 12.4 us per-pixel, 18.4 us in strips --
        // the windows do not overlap, so a strip saves no loads and has a quarter of the threads)
        const bool square = d->stride_h == 1 && d->stride_w == 1 && d->kh == d->kw && (d->kh == 2 || d->kh == 3);
        if (square) {
            const int64_t strips = (int64_t)d->batch * d->out_h * ((d->out_w + POOL_WP - 1) / POOL_WP) * (d->channels / 4);
            const dim3 sgrid(stream_grid(strips, 256));
            const bool is_max = d->kind == FNET_POOL_MAX;
#define FNET_POOL_FWD(KH_, S_)                                                                                                   \
    if (d->kh == KH_ && d->stride_h == S_) {                                                                                     \
        if (is_max) launch_pdl(pool_fwd_strip_kernel<KH_, KH_, S_, true>, sgrid, dim3(256), 0, as_stream(stream), *d, x4, y4, a4); \
        else launch_pdl(pool_fwd_strip_kernel<KH_, KH_, S_, false>, sgrid, dim3(256), 0, as_stream(stream), *d, x4, y4, a4);       \
    }
            FNET_POOL_FWD(2, 1) FNET_POOL_FWD(3, 1)
#undef FNET_POOL_FWD
            return check_launch("fnet_pool2d_fwd");
        }
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
    const bool square = d->stride_h == d->stride_w && (d->stride_h == 1 || d->stride_h == 2) && d->kh == d->kw && (d->kh == 2 || d->kh == 3);
    if (!(d->channels & 3) && aligned && total < 0x7fffffffLL - (1 << 26) && out_total < 0x7fffffffLL && square) {
        const int64_t strips = (int64_t)d->batch * d->in_h * ((d->in_w + POOL_WP - 1) / POOL_WP) * (d->channels / 4);
        const dim3 sgrid(stream_grid(strips, 256));
        const bool is_max = d->kind == FNET_POOL_MAX;
        const float4 *dy4 = reinterpret_cast<const float4 *>(dy);
        const uchar4 *a4 = reinterpret_cast<const uchar4 *>(argmax);
        float4 *dx4 = reinterpret_cast<float4 *>(dx);
#define FNET_POOL_BWD(KH_, S_)                                                                                                                    \
    if (d->kh == KH_ && d->stride_h == S_) {                                                                                                      \
        if (is_max) launch_pdl(pool_bwd_strip_kernel<KH_, KH_, S_, true>, sgrid, dim3(256), 0, as_stream(stream), *d, dy4, a4, dx4, accumulate);    \
        else launch_pdl(pool_bwd_strip_kernel<KH_, KH_, S_, false>, sgrid, dim3(256), 0, as_stream(stream), *d, dy4, a4, dx4, accumulate);          \
    }
        FNET_POOL_BWD(2, 1) FNET_POOL_BWD(2, 2) FNET_POOL_BWD(3, 1) FNET_POOL_BWD(3, 2)
#undef FNET_POOL_BWD
        return check_launch("fnet_pool2d_bwd");
    }
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

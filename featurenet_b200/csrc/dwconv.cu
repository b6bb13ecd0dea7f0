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

// Four channels per thread (C % 4 == 0, 16-byte aligned tensors, < 2^31 elements): 32-bit index arithmetic and one
// decode, one bounds check and two LDG.128 per tap for four outputs.  The scalar kernels above are issue-bound at
// ~350 GB/s (3x3 on 128 x 32 x 32 x 64: 191 us for 67 MB).
__global__ void __launch_bounds__(256) dw_fwd_vec4_kernel(fnet_conv2d_desc d, const float4 *__restrict__ x,
                                                          const float4 *__restrict__ w, const float4 *__restrict__ bias,
                                                          float4 *__restrict__ y, int act) {
    pdl_prologue();
    const int quads = d.in_c >> 2;
    const int total = d.batch * d.out_h * d.out_w * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        int wo = t % d.out_w;
        t /= d.out_w;
        int ho = t % d.out_h, b = t / d.out_h;
        const int h0 = ho * d.stride_h - d.pad_top, w0 = wo * d.stride_w - d.pad_left;
        const float4 *img = x + (int64_t)b * d.in_h * d.in_w * quads + q;
        float4 acc = bias ? __ldg(bias + q) : make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < d.kh; ++r) {
            const int h = h0 + r;
            if ((unsigned)h >= (unsigned)d.in_h) continue;
            for (int s = 0; s < d.kw; ++s) {
                const int ww = w0 + s;
                if ((unsigned)ww >= (unsigned)d.in_w) continue;
                const float4 xv = __ldg(img + (h * d.in_w + ww) * quads), wv = __ldg(w + (r * d.kw + s) * quads + q);
                acc.x = fmaf(xv.x, wv.x, acc.x);
                acc.y = fmaf(xv.y, wv.y, acc.y);
                acc.z = fmaf(xv.z, wv.z, acc.z);
                acc.w = fmaf(xv.w, wv.w, acc.w);
            }
        }
        y[i] = make_float4(act_apply(acc.x, act), act_apply(acc.y, act), act_apply(acc.z, act), act_apply(acc.w, act));
    }
}

__global__ void __launch_bounds__(256) dw_dgrad_vec4_kernel(fnet_conv2d_desc d, const float4 *__restrict__ dz,
                                                            const float4 *__restrict__ w, float4 *__restrict__ dx,
                                                            int accumulate) {
    pdl_prologue();
    const int quads = d.in_c >> 2;
    const int total = d.batch * d.in_h * d.in_w * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        int wi = t % d.in_w;
        t /= d.in_w;
        int hi = t % d.in_h, b = t / d.in_h;
        const float4 *img = dz + (int64_t)b * d.out_h * d.out_w * quads + q;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < d.kh; ++r) {
            const int hh = hi + d.pad_top - r;
            if (hh < 0) break;
            const int ho = hh / d.stride_h;
            if (ho * d.stride_h != hh || ho >= d.out_h) continue;
            for (int s = 0; s < d.kw; ++s) {
                const int ww = wi + d.pad_left - s;
                if (ww < 0) break;
                const int wo = ww / d.stride_w;
                if (wo * d.stride_w != ww || wo >= d.out_w) continue;
                const float4 gv = __ldg(img + (ho * d.out_w + wo) * quads), wv = __ldg(w + (r * d.kw + s) * quads + q);
                acc.x = fmaf(gv.x, wv.x, acc.x);
                acc.y = fmaf(gv.y, wv.y, acc.y);
                acc.z = fmaf(gv.z, wv.z, acc.z);
                acc.w = fmaf(gv.w, wv.w, acc.w);
            }
        }
        if (accumulate) {
            const float4 old = dx[i];
            acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
        }
        dx[i] = acc;
    }
}

// Kernel size known at compile time (the NAS space uses 1, 3 and 5 per axis): the tap loops unroll, a row's loads are
// issued together with a clamped address and a zero for padding instead of a branch per tap -- the runtime-bound
// loops above are chains of (bounds check -> load -> FMA) that leave one load in flight per thread.
template <int KH, int KW>
__global__ void __launch_bounds__(256) dw_fwd_vec4_fixed_kernel(fnet_conv2d_desc d, const float4 *__restrict__ x,
                                                                const float4 *__restrict__ w, const float4 *__restrict__ bias,
                                                                float4 *__restrict__ y, int act) {
    pdl_prologue();
    const int quads = d.in_c >> 2;
    const int total = d.batch * d.out_h * d.out_w * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        int wo = t % d.out_w;
        t /= d.out_w;
        int ho = t % d.out_h, b = t / d.out_h;
        const int h0 = ho * d.stride_h - d.pad_top, w0 = wo * d.stride_w - d.pad_left;
        const float4 *img = x + (int64_t)b * d.in_h * d.in_w * quads + q;
        float4 acc = bias ? __ldg(bias + q) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < KH; ++r) {
            const int h = h0 + r;
            const bool row_ok = (unsigned)h < (unsigned)d.in_h;
            float4 xv[KW], wv[KW];
#pragma unroll
            for (int s = 0; s < KW; ++s) {
                const int ww = w0 + s;
                const bool ok = row_ok && (unsigned)ww < (unsigned)d.in_w;
                xv[s] = __ldg(img + (ok ? (h * d.in_w + ww) * quads : 0));
                if (!ok) xv[s] = make_float4(0.f, 0.f, 0.f, 0.f);
                wv[s] = __ldg(w + (r * KW + s) * quads + q);
            }
#pragma unroll
            for (int s = 0; s < KW; ++s) {
                acc.x = fmaf(xv[s].x, wv[s].x, acc.x);
                acc.y = fmaf(xv[s].y, wv[s].y, acc.y);
                acc.z = fmaf(xv[s].z, wv[s].z, acc.z);
                acc.w = fmaf(xv[s].w, wv[s].w, acc.w);
            }
        }
        y[i] = make_float4(act_apply(acc.x, act), act_apply(acc.y, act), act_apply(acc.z, act), act_apply(acc.w, act));
    }
}

template <int KH, int KW>
__global__ void __launch_bounds__(256) dw_dgrad_vec4_fixed_kernel(fnet_conv2d_desc d, const float4 *__restrict__ dz,
                                                                  const float4 *__restrict__ w, float4 *__restrict__ dx,
                                                                  int accumulate) {
    pdl_prologue();
    const int quads = d.in_c >> 2;
    const int total = d.batch * d.in_h * d.in_w * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        int wi = t % d.in_w;
        t /= d.in_w;
        int hi = t % d.in_h, b = t / d.in_h;
        const float4 *img = dz + (int64_t)b * d.out_h * d.out_w * quads + q;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < KH; ++r) {
            const int hh = hi + d.pad_top - r;
            const int ho = hh / d.stride_h;
            const bool row_ok = hh >= 0 && ho * d.stride_h == hh && ho < d.out_h;
            float4 gv[KW], wv[KW];
#pragma unroll
            for (int s = 0; s < KW; ++s) {
                const int ww = wi + d.pad_left - s;
                const int wo = ww / d.stride_w;
                const bool ok = row_ok && ww >= 0 && wo * d.stride_w == ww && wo < d.out_w;
                gv[s] = __ldg(img + (ok ? (ho * d.out_w + wo) * quads : 0));
                if (!ok) gv[s] = make_float4(0.f, 0.f, 0.f, 0.f);
                wv[s] = __ldg(w + (r * KW + s) * quads + q);
            }
#pragma unroll
            for (int s = 0; s < KW; ++s) {
                acc.x = fmaf(gv[s].x, wv[s].x, acc.x);
                acc.y = fmaf(gv[s].y, wv[s].y, acc.y);
                acc.z = fmaf(gv[s].z, wv[s].z, acc.z);
                acc.w = fmaf(gv[s].w, wv[s].w, acc.w);
            }
        }
        if (accumulate) {
            const float4 old = dx[i];
            acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
        }
        dx[i] = acc;
    }
}

// Stride-1 depthwise correlation, four consecutive destination pixels per thread.  dst[h][w] = sum over taps of
// src[h + off_h + r][w + off_w + s] * w[tap(r, s)] serves both the forward pass (src = x, off = -pad) and the data
// gradient (src = dz, off = pad - (k - 1), taps flipped).  One pixel per thread loads k*k source quads and k*k weight
// quads per output quad (18 loads for 3x3); a strip of four shares its window: per filter row k + 3 source loads and k
// weight loads feed four outputs (6.75 loads per output quad for 3x3), all of a row's loads issued before the first FMA.
template <int KH, int KW, int WP>
__global__ void __launch_bounds__(256, 3) dw_strip_kernel(int batch, int hs, int ws_, int hd, int wd, int quads, int off_h, int off_w,
                                                       int flip, const float4 *__restrict__ src, const float4 *__restrict__ w,
                                                       const float4 *__restrict__ bias, float4 *dst, int act, int accumulate) {
    pdl_prologue();
    const int strips = (wd + WP - 1) / WP;
    const int total = batch * hd * strips * quads;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int q = i % quads, t = i / quads;
        const int st = t % strips;
        t /= strips;
        const int h = t % hd, b = t / hd, w0 = st * WP;
        const float4 *img = src + (int64_t)b * hs * ws_ * quads + q;
        float4 acc[WP];
        const float4 b4 = bias ? __ldg(bias + q) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int p = 0; p < WP; ++p) acc[p] = b4;
#pragma unroll 1                                   // one filter row's loads at a time: all rows at once is 128 registers
        for (int r = 0; r < KH; ++r) {
            const int hh = h + off_h + r;
            const bool row_ok = (unsigned)hh < (unsigned)hs;
            float4 xv[KW + WP - 1], wv[KW];
#pragma unroll
            for (int c = 0; c < KW + WP - 1; ++c) {
                const int ww = w0 + off_w + c;
                const bool ok = row_ok && (unsigned)ww < (unsigned)ws_;
                xv[c] = __ldg(img + (ok ? (hh * ws_ + ww) * quads : 0));
                if (!ok) xv[c] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int sx = 0; sx < KW; ++sx) wv[sx] = __ldg(w + (flip ? (KH - 1 - r) * KW + (KW - 1 - sx) : r * KW + sx) * quads + q);
#pragma unroll
            for (int p = 0; p < WP; ++p)
#pragma unroll
                for (int sx = 0; sx < KW; ++sx) {
                    acc[p].x = fmaf(xv[p + sx].x, wv[sx].x, acc[p].x);
                    acc[p].y = fmaf(xv[p + sx].y, wv[sx].y, acc[p].y);
                    acc[p].z = fmaf(xv[p + sx].z, wv[sx].z, acc[p].z);
                    acc[p].w = fmaf(xv[p + sx].w, wv[sx].w, acc[p].w);
                }
        }
        float4 *out = dst + ((int64_t)(b * hd + h) * wd + w0) * quads + q;
#pragma unroll
        for (int p = 0; p < WP; ++p)
            if (w0 + p < wd) {
                float4 v = acc[p];
                if (accumulate) { const float4 o = out[p * quads]; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
                out[p * quads] = make_float4(act_apply(v.x, act), act_apply(v.y, act), act_apply(v.z, act), act_apply(v.w, act));
            }
    }
}

static bool dw_strip_ok(const fnet_conv2d_desc *d) {
    return d->stride_h == 1 && d->stride_w == 1 && d->dil_h == 1 && d->dil_w == 1 && d->kh * d->kw <= 25 && d->kh * d->kw > 1;
}

template <int KH, int KW>
static void dw_strip_launch(const fnet_conv2d_desc *d, bool dgrad, const float4 *src, const float4 *w, const float4 *bias,
                            float4 *dst, int act, int accumulate, cudaStream_t stream) {
    constexpr int WP = 4;
    const int quads = d->in_c >> 2;
    const int hs = dgrad ? d->out_h : d->in_h, ws_ = dgrad ? d->out_w : d->in_w, hd = dgrad ? d->in_h : d->out_h,
              wd = dgrad ? d->in_w : d->out_w;
    const int off_h = dgrad ? d->pad_top - (KH - 1) : -d->pad_top, off_w = dgrad ? d->pad_left - (KW - 1) : -d->pad_left;
    const int64_t total = (int64_t)d->batch * hd * ((wd + WP - 1) / WP) * quads;
    launch_pdl(dw_strip_kernel<KH, KW, WP>, dim3(stream_grid(total, 256)), dim3(256), 0, stream, d->batch, hs, ws_, hd, wd, quads, off_h,
               off_w, dgrad ? 1 : 0, src, w, bias, dst, act, accumulate);
}

// stage 1: CTA (cgroup, split) walks its share of output pixels; thread = (channel lane, pixel lane).
// partial[split][tap or bias][c]
constexpr int kDwChan = 32;   // channels per CTA (x dimension of the thread block)
constexpr int kDwPix = 8;     // pixel lanes per CTA (y dimension)

// I = type of the pixel index: int whenever the tensor allows it (64-bit division is a long software routine and the
// loop decodes the pixel index three times per trip)
template <int MAXROWS, typename I>
__global__ void __launch_bounds__(kDwChan * kDwPix) dw_wgrad_partial_kernel(fnet_conv2d_desc d, int pix_per_split,
                                                                           const float *__restrict__ x,
                                                                           const float *__restrict__ dz,
                                                                           float *__restrict__ partial, int rows) {
    pdl_prologue();
    // blockDim = (channel lanes, pixel lanes), 256 threads: 32 x 8 for wide layers, 4 x 64 for 3-channel ones (with a
    // fixed 32 x 8 split a 3-channel layer left 29 of 32 lanes idle: 186 us for a 5x5 on 128 x 32 x 32 x 3)
    __shared__ float red[kDwChan * kDwPix + kDwPix * 8];
    const int lanes_c = blockDim.x, lanes_p = blockDim.y, pitch = lanes_c + 1;
    int C = d.in_c;
    int c = blockIdx.x * lanes_c + threadIdx.x;
    int split = blockIdx.y;
    I npix = (I)d.batch * d.out_h * d.out_w;
    I p0 = (I)split * pix_per_split;
    I p1 = p0 + pix_per_split < npix ? p0 + pix_per_split : npix;
    int taps = d.kh * d.kw;
    float acc[MAXROWS];                 // one accumulator per tap (+ bias): x and dz are read once
#pragma unroll
    for (int t = 0; t < MAXROWS; ++t) acc[t] = 0.f;
    if (c < C) {
        for (I p = p0 + threadIdx.y; p < p1; p += lanes_p) {
            float g = __ldg(dz + (int64_t)p * C + c);
            int wo = (int)(p % (I)d.out_w);
            I q = p / (I)d.out_w;
            int ho = (int)(q % (I)d.out_h);
            I b = q / (I)d.out_h;
            int h0 = ho * d.stride_h - d.pad_top, w0 = wo * d.stride_w - d.pad_left;
            const float *img = x + (int64_t)b * d.in_h * d.in_w * C + c;
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
            // pixel lanes of one warp first (shuffles), then the eight warp sums through shared memory: the serial sum
            // over up to 64 pixel lanes per accumulator row used to be the longest part of a CTA's life
            float v = acc[t];
            for (int o = lanes_c; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            const int tid = threadIdx.y * lanes_c + threadIdx.x;
            if ((tid & 31) < lanes_c) red[(tid >> 5) * pitch + (tid & 31)] = v;
            __syncthreads();
            if (threadIdx.y == 0 && c < C) {
                float sum = 0.f;
#pragma unroll
                for (int j = 0; j < kDwChan * kDwPix / 32; ++j) sum += red[j * pitch + threadIdx.x];
                partial[((int64_t)split * rows + t) * C + c] = sum;
            }
            __syncthreads();
        }
    }
}

// The same with four channels per thread (C % 4 == 0): one LDG.128 of dz and one of x per tap feed four sums, and
// the 256 threads cover four times as many pixels per trip (the scalar kernel is bound by the latency of its
// per-tap loads: 150 us for a 3x3 on 128 x 32 x 32 x 64).
template <int MAXROWS>
__global__ void __launch_bounds__(kDwChan * kDwPix) dw_wgrad_partial_vec4_kernel(fnet_conv2d_desc d, int pix_per_split,
                                                                                const float4 *__restrict__ x,
                                                                                const float4 *__restrict__ dz,
                                                                                float4 *__restrict__ partial, int rows) {
    pdl_prologue();
    __shared__ float4 red[kDwChan * kDwPix + 64];
    const int lanes_q = blockDim.x, lanes_p = blockDim.y, pitch = lanes_q + 1;
    const int quads = d.in_c >> 2;
    const int q = blockIdx.x * lanes_q + threadIdx.x;
    const int split = blockIdx.y;
    const int npix = d.batch * d.out_h * d.out_w;
    const int p0 = split * pix_per_split;
    const int p1 = min(p0 + pix_per_split, npix);
    const int taps = d.kh * d.kw;
    float4 acc[MAXROWS];
#pragma unroll
    for (int t = 0; t < MAXROWS; ++t) acc[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q < quads) {
        for (int p = p0 + threadIdx.y; p < p1; p += lanes_p) {
            const float4 g = __ldg(dz + (int64_t)p * quads + q);
            const int wo = p % d.out_w, t2 = p / d.out_w;
            const int ho = t2 % d.out_h, b = t2 / d.out_h;
            const int h0 = ho * d.stride_h - d.pad_top, w0 = wo * d.stride_w - d.pad_left;
            const float4 *img = x + (int64_t)b * d.in_h * d.in_w * quads + q;
            int r = 0, s = 0;
#pragma unroll
            for (int t = 0; t < MAXROWS; ++t) {
                if (t < taps) {
                    const int h = h0 + r, ww = w0 + s;
                    if ((unsigned)h < (unsigned)d.in_h && (unsigned)ww < (unsigned)d.in_w) {
                        const float4 xv = __ldg(img + (h * d.in_w + ww) * quads);
                        acc[t].x = fmaf(g.x, xv.x, acc[t].x);
                        acc[t].y = fmaf(g.y, xv.y, acc[t].y);
                        acc[t].z = fmaf(g.z, xv.z, acc[t].z);
                        acc[t].w = fmaf(g.w, xv.w, acc[t].w);
                    }
                    if (++s == d.kw) { s = 0; ++r; }
                } else if (t == taps) {
                    acc[t].x += g.x; acc[t].y += g.y; acc[t].z += g.z; acc[t].w += g.w;      // bias row
                }
            }
        }
    }
#pragma unroll
    for (int t = 0; t < MAXROWS; ++t) {
        if (t < rows) {
            float4 v = acc[t];                                 // as in the scalar kernel: warp shuffles, then 8 warp sums
            for (int o = lanes_q; o < 32; o <<= 1) {
                v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
                v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
                v.z += __shfl_xor_sync(0xffffffffu, v.z, o);
                v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
            }
            const int tid = threadIdx.y * lanes_q + threadIdx.x;
            if ((tid & 31) < lanes_q) red[(tid >> 5) * pitch + (tid & 31)] = v;
            __syncthreads();
            if (threadIdx.y == 0 && q < quads) {
                float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int j = 0; j < kDwChan * kDwPix / 32; ++j) {
                    const float4 w8 = red[j * pitch + threadIdx.x];
                    sum.x += w8.x; sum.y += w8.y; sum.z += w8.z; sum.w += w8.w;
                }
                partial[((int64_t)split * rows + t) * quads + q] = sum;
            }
            __syncthreads();
        }
    }
}

// Stride 1, four channels per thread, FOUR neighbouring output pixels of one row per trip: a row of the window is
// loaded once for the four pixels (KH * (3 + KW) loads of x instead of 4 * KH * KW: 18 instead of 36 for 3x3) and the
// pixel index is decoded once per strip.  Same CTA shape, partial layout and reduction as the kernels above.
template <int KH, int KW>
__global__ void __launch_bounds__(kDwChan * kDwPix) dw_wgrad_strip_kernel(fnet_conv2d_desc d, int strips_per_split,
                                                                         const float4 *__restrict__ x,
                                                                         const float4 *__restrict__ dz,
                                                                         float4 *__restrict__ partial, int rows) {
    pdl_prologue();
    constexpr int WP = 4, NCOL = WP - 1 + KW, TAPS = KH * KW;
    __shared__ float4 red[kDwChan * kDwPix + 64];
    const int lanes_q = blockDim.x, lanes_p = blockDim.y, pitch = lanes_q + 1;
    const int quads = d.in_c >> 2;
    const int q = blockIdx.x * lanes_q + threadIdx.x;
    const int split = blockIdx.y;
    const int strips_w = (d.out_w + WP - 1) / WP;
    const int nstrips = d.batch * d.out_h * strips_w;
    const int s0 = split * strips_per_split, s1 = min(s0 + strips_per_split, nstrips);
    float4 acc[TAPS + 1];
#pragma unroll
    for (int t = 0; t <= TAPS; ++t) acc[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q < quads) {
        for (int si = s0 + threadIdx.y; si < s1; si += lanes_p) {
            const int ws_ = si % strips_w, t2 = si / strips_w;
            const int ho = t2 % d.out_h, b = t2 / d.out_h;
            const int wo0 = ws_ * WP, w0 = wo0 - d.pad_left, h0 = ho - d.pad_top;
            const float4 *gp = dz + ((int64_t)(b * d.out_h + ho) * d.out_w + wo0) * quads + q;
            float4 g[WP];
#pragma unroll
            for (int j = 0; j < WP; ++j) g[j] = wo0 + j < d.out_w ? __ldg(gp + j * quads) : make_float4(0.f, 0.f, 0.f, 0.f);
            acc[TAPS].x += (g[0].x + g[1].x) + (g[2].x + g[3].x);      // bias row
            acc[TAPS].y += (g[0].y + g[1].y) + (g[2].y + g[3].y);
            acc[TAPS].z += (g[0].z + g[1].z) + (g[2].z + g[3].z);
            acc[TAPS].w += (g[0].w + g[1].w) + (g[2].w + g[3].w);
            const float4 *img = x + (int64_t)b * d.in_h * d.in_w * quads + q;
#pragma unroll
            for (int r = 0; r < KH; ++r) {
                const int h = h0 + r;
                if ((unsigned)h >= (unsigned)d.in_h) continue;
                float4 xv[NCOL];
#pragma unroll
                for (int c = 0; c < NCOL; ++c) {
                    const int w = w0 + c;
                    const bool ok = (unsigned)w < (unsigned)d.in_w;
                    xv[c] = __ldg(img + (h * d.in_w + (ok ? w : 0)) * quads);
                    if (!ok) xv[c] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int s2 = 0; s2 < KW; ++s2) {
#pragma unroll
                    for (int j = 0; j < WP; ++j) {
                        float4 &a = acc[r * KW + s2];
                        a.x = fmaf(g[j].x, xv[j + s2].x, a.x);
                        a.y = fmaf(g[j].y, xv[j + s2].y, a.y);
                        a.z = fmaf(g[j].z, xv[j + s2].z, a.z);
                        a.w = fmaf(g[j].w, xv[j + s2].w, a.w);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int t = 0; t <= TAPS; ++t) {
        if (t < rows) {
            float4 v = acc[t];
            for (int o = lanes_q; o < 32; o <<= 1) {
                v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
                v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
                v.z += __shfl_xor_sync(0xffffffffu, v.z, o);
                v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
            }
            const int tid = threadIdx.y * lanes_q + threadIdx.x;
            if ((tid & 31) < lanes_q) red[(tid >> 5) * pitch + (tid & 31)] = v;
            __syncthreads();
            if (threadIdx.y == 0 && q < quads) {
                float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int j = 0; j < kDwChan * kDwPix / 32; ++j) {
                    const float4 w8 = red[j * pitch + threadIdx.x];
                    sum.x += w8.x; sum.y += w8.y; sum.z += w8.z; sum.w += w8.w;
                }
                partial[((int64_t)split * rows + t) * quads + q] = sum;
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(256) dw_wgrad_reduce_kernel(int splits, int rows, int taps, int C,
                                                              const float *__restrict__ partial, float *__restrict__ dw,
                                                              float *__restrict__ dbias) {
    pdl_prologue();
    // one warp per output: lanes over the splits (coalescing does not matter for a few hundred KB, the depth of the
    // dependent chain does: one thread per output walked up to 1024 partials in sequence), fixed shuffle tree
    const int total = rows * C, lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < total; i += warps) {
        float s = 0.f;
        for (int z = lane; z < splits; z += 32) s += partial[(int64_t)z * total + i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) {
            int row = i / C, c = i - row * C;
            if (row < taps) dw[i] = s;
            else dbias[c] = s;
        }
    }
}

// channel lanes of the filter-gradient CTA: the smallest power of two >= C, between 4 and 32
static int dw_chan_lanes(int c) {
    int lanes = 4;
    while (lanes < c && lanes < kDwChan) lanes *= 2;
    return lanes;
}

// four channels per thread in the filter gradient: channel count, kernel size (register budget) and index range allow it
static bool dw_wgrad_vec4(const fnet_conv2d_desc *d) {
    return !(d->in_c & 3) && d->kh * d->kw + 1 <= 26 && (int64_t)d->batch * d->in_h * d->in_w * d->in_c < 0x7fffffffLL &&
           (int64_t)d->batch * d->out_h * d->out_w * d->in_c < 0x7fffffffLL;
}

static int dw_splits(const fnet_conv2d_desc *d) {
    int64_t npix = (int64_t)d->batch * d->out_h * d->out_w;
    const int units = dw_wgrad_vec4(d) ? d->in_c / 4 : d->in_c;      // what a thread's x lane stands for
    const int lanes_c = dw_chan_lanes(units), lanes_p = kDwChan * kDwPix / lanes_c;
    int cgroups = ceil_div(units, lanes_c);
    int64_t want = ceil_div(4 * kNumSMs, cgroups);
    int64_t most = ceil_div<int64_t>(npix, 4 * lanes_p);
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
    const int64_t in_total = (int64_t)d->batch * d->in_h * d->in_w * d->in_c;
    const bool vec = !(d->in_c & 3) && total < 0x7fffffffLL - (1 << 26) && in_total < 0x7fffffffLL &&
                     !((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(bias) |
                        reinterpret_cast<uintptr_t>(y)) & 15);
    if (vec) {
        const float4 *x4 = reinterpret_cast<const float4 *>(x), *w4 = reinterpret_cast<const float4 *>(w);
        const float4 *b4 = reinterpret_cast<const float4 *>(bias);
        float4 *y4 = reinterpret_cast<float4 *>(y);
        const dim3 grid(stream_grid(total / 4, 256));
#define FNET_DW_FWD(KH_, KW_) \
    else if (d->kh == KH_ && d->kw == KW_) launch_pdl(dw_fwd_vec4_fixed_kernel<KH_, KW_>, grid, dim3(256), 0, as_stream(stream), *d, x4, w4, b4, y4, act)
        if (dw_strip_ok(d) && d->kh == 3 && d->kw == 3) dw_strip_launch<3, 3>(d, false, x4, w4, b4, y4, act, 0, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 1 && d->kw == 3) dw_strip_launch<1, 3>(d, false, x4, w4, b4, y4, act, 0, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 3 && d->kw == 1) dw_strip_launch<3, 1>(d, false, x4, w4, b4, y4, act, 0, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 2 && d->kw == 2) dw_strip_launch<2, 2>(d, false, x4, w4, b4, y4, act, 0, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 5 && d->kw == 5) dw_strip_launch<5, 5>(d, false, x4, w4, b4, y4, act, 0, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 1 && d->kw == 5) dw_strip_launch<1, 5>(d, false, x4, w4, b4, y4, act, 0, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 5 && d->kw == 1) dw_strip_launch<5, 1>(d, false, x4, w4, b4, y4, act, 0, as_stream(stream));
        FNET_DW_FWD(1, 1); FNET_DW_FWD(1, 3); FNET_DW_FWD(3, 1); FNET_DW_FWD(3, 3); FNET_DW_FWD(1, 5); FNET_DW_FWD(5, 1);
        // up to nine taps; 5x5 measured slower unrolled (65 vs 44 us on 128 x 32 x 32 x 24) and keeps the loops
        else launch_pdl(dw_fwd_vec4_kernel, grid, dim3(256), 0, as_stream(stream), *d, x4, w4, b4, y4, act);
#undef FNET_DW_FWD
    } else
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
    const int64_t out_total = (int64_t)d->batch * d->out_h * d->out_w * d->in_c;
    const bool vec = !(d->in_c & 3) && total < 0x7fffffffLL - (1 << 26) && out_total < 0x7fffffffLL &&
                     !((reinterpret_cast<uintptr_t>(dz) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(dx)) & 15);
    if (vec) {
        const float4 *g4 = reinterpret_cast<const float4 *>(dz), *w4 = reinterpret_cast<const float4 *>(w);
        float4 *dx4 = reinterpret_cast<float4 *>(dx);
        const dim3 grid(stream_grid(total / 4, 256));
#define FNET_DW_DGRAD(KH_, KW_) \
    else if (d->kh == KH_ && d->kw == KW_) launch_pdl(dw_dgrad_vec4_fixed_kernel<KH_, KW_>, grid, dim3(256), 0, as_stream(stream), *d, g4, w4, dx4, accumulate)
        if (dw_strip_ok(d) && d->kh == 3 && d->kw == 3) dw_strip_launch<3, 3>(d, true, g4, w4, nullptr, dx4, 0, accumulate, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 1 && d->kw == 3) dw_strip_launch<1, 3>(d, true, g4, w4, nullptr, dx4, 0, accumulate, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 3 && d->kw == 1) dw_strip_launch<3, 1>(d, true, g4, w4, nullptr, dx4, 0, accumulate, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 2 && d->kw == 2) dw_strip_launch<2, 2>(d, true, g4, w4, nullptr, dx4, 0, accumulate, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 5 && d->kw == 5) dw_strip_launch<5, 5>(d, true, g4, w4, nullptr, dx4, 0, accumulate, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 1 && d->kw == 5) dw_strip_launch<1, 5>(d, true, g4, w4, nullptr, dx4, 0, accumulate, as_stream(stream));
        else if (dw_strip_ok(d) && d->kh == 5 && d->kw == 1) dw_strip_launch<5, 1>(d, true, g4, w4, nullptr, dx4, 0, accumulate, as_stream(stream));
        FNET_DW_DGRAD(1, 1); FNET_DW_DGRAD(1, 3); FNET_DW_DGRAD(3, 1); FNET_DW_DGRAD(3, 3); FNET_DW_DGRAD(1, 5); FNET_DW_DGRAD(5, 1);
        else launch_pdl(dw_dgrad_vec4_kernel, grid, dim3(256), 0, as_stream(stream), *d, g4, w4, dx4, accumulate);
#undef FNET_DW_DGRAD
    } else
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
    const bool vec = dw_wgrad_vec4(d) && !((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dz) | reinterpret_cast<uintptr_t>(ws)) & 15);
    const int units = dw_wgrad_vec4(d) ? d->in_c / 4 : d->in_c;     // the split plan (workspace) follows the shape alone
    const int lanes_c = dw_chan_lanes(vec ? units : d->in_c);
    dim3 grid(ceil_div(vec ? units : d->in_c, lanes_c), splits), block(lanes_c, kDwChan * kDwPix / lanes_c);
    const bool strip = vec && d->stride_h == 1 && d->stride_w == 1 &&
                       ((d->kh == 3 && d->kw == 3) || (d->kh == 1 && d->kw == 3) || (d->kh == 3 && d->kw == 1) || (d->kh == 2 && d->kw == 2) ||
                        (d->kh == 5 && d->kw == 5));
    if (strip) {
        // the same number of CTAs per channel group, each over a share of the four-pixel strips
        const float4 *x4 = reinterpret_cast<const float4 *>(x), *dz4 = reinterpret_cast<const float4 *>(dz);
        float4 *p4 = reinterpret_cast<float4 *>(ws);
        const int64_t nstrips = (int64_t)d->batch * d->out_h * ((d->out_w + 3) / 4);
        const int strips_per_split = (int)ceil_div<int64_t>(nstrips, splits);
        splits = (int)ceil_div<int64_t>(nstrips, strips_per_split);
        grid.y = splits;
#define FNET_DW_WSTRIP(KH_, KW_) \
    if (d->kh == KH_ && d->kw == KW_) launch_pdl(dw_wgrad_strip_kernel<KH_, KW_>, dim3(grid), dim3(block), 0, as_stream(stream), *d, strips_per_split, x4, dz4, p4, rows)
        FNET_DW_WSTRIP(3, 3); FNET_DW_WSTRIP(1, 3); FNET_DW_WSTRIP(3, 1); FNET_DW_WSTRIP(2, 2); FNET_DW_WSTRIP(5, 5);
#undef FNET_DW_WSTRIP
    } else if (vec) {
        const float4 *x4 = reinterpret_cast<const float4 *>(x), *dz4 = reinterpret_cast<const float4 *>(dz);
        float4 *p4 = reinterpret_cast<float4 *>(ws);
        if (taps + 1 <= 10)
            launch_pdl(dw_wgrad_partial_vec4_kernel<10>, dim3(grid), dim3(block), 0, as_stream(stream), *d, pix_per_split, x4, dz4, p4, rows);
        else
            launch_pdl(dw_wgrad_partial_vec4_kernel<26>, dim3(grid), dim3(block), 0, as_stream(stream), *d, pix_per_split, x4, dz4, p4, rows);
    } else {
    const bool small = npix < 0x7fffffffLL - (1 << 20);
#define FNET_DW_WGRAD(R_)                                                                                                              \
    do {                                                                                                                               \
        if (small) launch_pdl(dw_wgrad_partial_kernel<R_, int>, dim3(grid), dim3(block), 0, as_stream(stream), *d, pix_per_split, x, dz, partial, rows); \
        else launch_pdl(dw_wgrad_partial_kernel<R_, int64_t>, dim3(grid), dim3(block), 0, as_stream(stream), *d, pix_per_split, x, dz, partial, rows);   \
    } while (0)
    if (taps + 1 <= 10) FNET_DW_WGRAD(10);
    else if (taps + 1 <= 26) FNET_DW_WGRAD(26);
    else if (taps + 1 <= 50) FNET_DW_WGRAD(50);
    else {
        set_error("fnet_dwconv2d_bwd_filter: kernels above 7x7 are outside the FeatureNet vocabulary");
        return FNET_ERR_UNSUPPORTED;
    }
#undef FNET_DW_WGRAD
    }
    rc = check_launch("fnet_dwconv2d_bwd_filter");
    if (rc) return rc;
    launch_pdl(dw_wgrad_reduce_kernel, dim3(stream_grid((int64_t)rows * d->in_c * 32, 256)), dim3(256), 0, as_stream(stream), 
        splits, rows, taps, d->in_c, partial, dw, dbias);
    return check_launch("fnet_dwconv2d_bwd_filter");
}

}  // extern "C"

// The fused step: ONE persistent kernel runs a whole optimisation step of a narrow candidate.
//
// Why: the products of main_1block_nas.xml are a few 1x1 / 2x2 convolutions between 3 and 6 channels, a concatenation
// and the softmax head -- 3 MB of activations per batch of 128.  As separate launches their step is 13 kernels of ~4 us
// each for ~7 us worth of memory traffic: launch latency, not bandwidth.  Here one CTA per SM walks the step's list of
// operations; operations that depend on one another are separated by a grid-wide barrier (~1 us) instead of a kernel
// boundary, independent ones (a layer's filter gradient and its data gradient) share a phase, and every tensor of the
// step stays in L2 from the operation that writes it to the one that reads it.
//
// Rules of this file:
//  * tensors written inside the kernel are read with plain loads (never __ldg / ld.global.nc: the non-coherent path
//    may serve a line fetched before another SM rewrote it); the barrier's fences make plain loads see them;
//  * every reduction has a fixed order (per-thread walk, shuffle tree, per-warp slots, per-CTA partials summed by
//    index): same inputs, same bits, whatever the scheduling;
//  * an operation strides over (blockIdx.x, gridDim.x) itself, so several operations of one phase simply follow one
//    another without idling anybody.
#include "common.cuh"
#include "loss_row.cuh"

namespace fnet {

constexpr int FS_THREADS = 512;
constexpr int FS_SMEM_FLOATS = 10240;          // 40 KB: a layer's filter (+ bias row), or the head's dz block
constexpr int FS_MAX_OPS = 48;                 // operations of one step (7.5 KB of shared memory)
constexpr int FS_WGRAD_ACC = 64;               // partial sums of a small filter gradient a thread keeps in registers

// ------------------------------------------------------------------ grid-wide barrier
// Every CTA adds 1, CTA 0 adds what makes the sum flip bit 31 once all have arrived; a CTA waits for the flip of the
// value it saw.  The counter needs no reset between barriers or launches (the scheme of cooperative groups' grid sync).
__device__ __forceinline__ void grid_barrier(unsigned *counter) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned add = blockIdx.x == 0 ? 0x80000000u - (gridDim.x - 1) : 1u;
        unsigned seen, now;
        // release: everything this CTA wrote (ordered before by the __syncthreads above) is visible to whoever
        // acquires the counter afterwards; acquire: ... and what the others wrote is visible to this CTA after the
        // __syncthreads below
        asm volatile("atom.add.release.gpu.global.u32 %0, [%1], %2;" : "=r"(seen) : "l"(counter), "r"(add) : "memory");
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(now) : "l"(counter) : "memory");
        } while (((seen ^ now) & 0x80000000u) == 0);
    }
    __syncthreads();
}

struct Geom {
    int B, H, W, C, Ho, Wo, K, kh, kw, sh, sw, pt, pl, dh, dw;
};

__device__ __forceinline__ Geom geom_of(const fnet_fused_op &op) {
    Geom g;
    g.B = op.d[0]; g.H = op.d[1]; g.W = op.d[2]; g.C = op.d[3]; g.Ho = op.d[4]; g.Wo = op.d[5]; g.K = op.d[6];
    g.kh = op.d[7]; g.kw = op.d[8]; g.sh = op.d[9]; g.sw = op.d[10]; g.pt = op.d[11]; g.pl = op.d[12]; g.dh = op.d[13];
    g.dw = op.d[14];
    return g;
}

__device__ __forceinline__ void lr_t_update(int32_t *state, float beta1, float beta2) {
    // Keras Adam: lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t), in double, once per step (fnet_step_begin)
    double t = (double)(state[0] + 1);
    double lr = (double)__int_as_float(state[2]);
    state[3] = __float_as_int((float)(lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t))));
}

// ------------------------------------------------------------------ batch in place
__device__ void fs_gather(const fnet_fused_op &op) {
    const int batch = op.d[0];
    const long long x_row = op.d[1], y_row = op.d[2];
    const int32_t *perm = static_cast<const int32_t *>(op.p[0]);
    int32_t *state = static_cast<int32_t *>(op.p[1]);
    const float *x_src = static_cast<const float *>(op.p[2]), *y_src = static_cast<const float *>(op.p[3]);
    float *x_dst = static_cast<float *>(op.p[4]), *y_dst = static_cast<float *>(op.p[5]);
    const int cursor = state[1];
    if (op.d[3] && blockIdx.x == 0 && threadIdx.x == 0) lr_t_update(state, op.f[0], op.f[1]);
    const bool vec = (x_row & 3) == 0;
    for (int row = blockIdx.x; row < batch; row += gridDim.x) {
        const long long from = __ldg(perm + cursor + row);          // the permutation and the dataset are never written here
        const float *xs = x_src + from * x_row;
        float *xd = x_dst + (long long)row * x_row;
        if (vec) {
            const float4 *s4 = reinterpret_cast<const float4 *>(xs);
            float4 *d4 = reinterpret_cast<float4 *>(xd);
            for (int i = threadIdx.x; i < (int)(x_row >> 2); i += FS_THREADS) d4[i] = __ldg(s4 + i);
        } else {
            for (int i = threadIdx.x; i < (int)x_row; i += FS_THREADS) xd[i] = __ldg(xs + i);
        }
        for (int i = threadIdx.x; i < (int)y_row; i += FS_THREADS) y_dst[(long long)row * y_row + i] = __ldg(y_src + from * y_row + i);
    }
}

__device__ void fs_feed_u8(const fnet_fused_op &op) {
    const long long x_elems = (long long)op.d[0] * op.d[1], y_elems = (long long)op.d[0] * op.d[2];
    int32_t *state = static_cast<int32_t *>(op.p[0]);
    const uint8_t *x_src = static_cast<const uint8_t *>(op.p[1]);
    const float *y_src = static_cast<const float *>(op.p[2]);
    float *x_dst = static_cast<float *>(op.p[3]), *y_dst = static_cast<float *>(op.p[4]);
    const float scale = op.f[2];
    const long long tid = (long long)blockIdx.x * FS_THREADS + threadIdx.x, stride = (long long)gridDim.x * FS_THREADS;
    if (op.d[3] && tid == 0) lr_t_update(state, op.f[0], op.f[1]);
    const long long vec = x_elems >> 4;
    const uint4 *s16 = reinterpret_cast<const uint4 *>(x_src);
    float4 *d4 = reinterpret_cast<float4 *>(x_dst);
    for (long long i = tid; i < vec; i += stride) {
        const uint4 q = s16[i];                                      // uploaded by a copy that finished before the launch
        const uint32_t words[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t wd = words[k];
            d4[i * 4 + k] = make_float4((float)(wd & 255u) / scale, (float)((wd >> 8) & 255u) / scale,
                                        (float)((wd >> 16) & 255u) / scale, (float)(wd >> 24) / scale);
        }
    }
    for (long long i = (vec << 4) + tid; i < x_elems; i += stride) x_dst[i] = (float)x_src[i] / scale;
    for (long long i = tid; i < y_elems; i += stride) y_dst[i] = y_src[i];
}

// ------------------------------------------------------------------ merges, activation derivative
template <typename T>
__device__ __forceinline__ void concat_fwd_body(int outer, int blk_a, int blk_b, const T *a, const T *b, T *y) {
    const int blk = blk_a + blk_b, total = outer * blk;
    for (int i = blockIdx.x * FS_THREADS + threadIdx.x; i < total; i += gridDim.x * FS_THREADS) {
        const int o = i / blk, e = i - o * blk;
        y[i] = e < blk_a ? a[o * blk_a + e] : b[o * blk_b + (e - blk_a)];
    }
}

__device__ void fs_concat_fwd(const fnet_fused_op &op) {
    const int outer = op.d[0], blk_a = op.d[1], blk_b = op.d[2];
    const float *a = static_cast<const float *>(op.p[0]), *b = static_cast<const float *>(op.p[1]);
    float *y = static_cast<float *>(op.p[2]);
    if (!((blk_a | blk_b) & 3) && !((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(y)) & 15))
        concat_fwd_body(outer, blk_a >> 2, blk_b >> 2, reinterpret_cast<const float4 *>(a), reinterpret_cast<const float4 *>(b),
                        reinterpret_cast<float4 *>(y));
    else
        concat_fwd_body(outer, blk_a, blk_b, a, b, y);
}

__device__ __forceinline__ float vsum(float a, float b) { return a + b; }
__device__ __forceinline__ float4 vsum(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

template <typename T>
__device__ __forceinline__ void concat_bwd_body(int outer, int blk_a, int blk_b, const T *dy, T *da, T *db, int acc_a, int acc_b) {
    const int blk = blk_a + blk_b, total = outer * blk;
    for (int i = blockIdx.x * FS_THREADS + threadIdx.x; i < total; i += gridDim.x * FS_THREADS) {
        const int o = i / blk, e = i - o * blk;
        if (e < blk_a) {
            if (da) { T *d = da + o * blk_a + e; *d = acc_a ? vsum(*d, dy[i]) : dy[i]; }
        } else if (db) {
            T *d = db + o * blk_b + (e - blk_a);
            *d = acc_b ? vsum(*d, dy[i]) : dy[i];
        }
    }
}

__device__ void fs_concat_bwd(const fnet_fused_op &op) {
    const int outer = op.d[0], blk_a = op.d[1], blk_b = op.d[2], acc_a = op.flags & 1, acc_b = op.flags & 2;
    const float *dy = static_cast<const float *>(op.p[0]);
    float *da = static_cast<float *>(op.p[1]), *db = static_cast<float *>(op.p[2]);
    if (!((blk_a | blk_b) & 3) && !((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(da) | reinterpret_cast<uintptr_t>(db)) & 15))
        concat_bwd_body(outer, blk_a >> 2, blk_b >> 2, reinterpret_cast<const float4 *>(dy), reinterpret_cast<float4 *>(da),
                        reinterpret_cast<float4 *>(db), acc_a, acc_b);
    else
        concat_bwd_body(outer, blk_a, blk_b, dy, da, db, acc_a, acc_b);
}

__device__ void fs_act_bwd(const fnet_fused_op &op) {
    const long long n = op.d[0];
    const float *y = static_cast<const float *>(op.p[0]), *dy = static_cast<const float *>(op.p[1]);
    float *dx = static_cast<float *>(op.p[2]);
    const int act = op.act, acc = op.flags & 1;
    for (long long i = (long long)blockIdx.x * FS_THREADS + threadIdx.x; i < n; i += (long long)gridDim.x * FS_THREADS) {
        const float v = dy[i] * act_grad(y[i], act);
        dx[i] = acc ? dx[i] + v : v;
    }
}

// out[i] = sum over the CTAs' partials, in CTA order (lanes take every 32nd CTA, then a shuffle tree): the second half
// of every small filter gradient.  One warp per output value, warps numbered over the whole grid.
__device__ void fs_reduce(const fnet_fused_op &op) {
    const int parts = op.d[0], total = op.d[1], split = op.d[2];
    const float *partial = static_cast<const float *>(op.p[0]);
    float *first = static_cast<float *>(op.p[1]), *rest = static_cast<float *>(op.p[2]);
    const int lane = threadIdx.x & 31, warp = blockIdx.x * (FS_THREADS / 32) + (threadIdx.x >> 5);
    for (int i = warp; i < total; i += gridDim.x * (FS_THREADS / 32)) {
        float s = 0.f;
        for (int z = lane; z < parts; z += 32) s += partial[(long long)z * total + i];
        s = warp_sum(s);
        if (lane == 0) {
            if (i < split) first[i] = s;
            else if (rest) rest[i - split] = s;
        }
    }
}

// ------------------------------------------------------------------ small-channel convolutions (C, K <= 16)
// Forward: a thread per output pixel, the filter (+ bias row) in shared memory, K sums in registers.
template <int KMAX>
__device__ void fs_conv_fwd(const fnet_fused_op &op, float *smem) {
    const Geom g = geom_of(op);
    const float *x = static_cast<const float *>(op.p[0]), *w = static_cast<const float *>(op.p[1]);
    const float *bias = static_cast<const float *>(op.p[2]);
    float *y = static_cast<float *>(op.p[3]);
    const int K = g.K, C = g.C, taps_c = g.kh * g.kw * C, act = op.act;
    for (int i = threadIdx.x; i < taps_c * K; i += FS_THREADS) smem[i] = w[i];
    for (int i = threadIdx.x; i < K; i += FS_THREADS) smem[taps_c * K + i] = bias ? bias[i] : 0.f;
    __syncthreads();
    const int M = g.B * g.Ho * g.Wo;
    for (int m = blockIdx.x * FS_THREADS + threadIdx.x; m < M; m += gridDim.x * FS_THREADS) {
        const int wo = m % g.Wo, t = m / g.Wo, ho = t % g.Ho, b = t / g.Ho;
        float acc[KMAX];
#pragma unroll
        for (int k = 0; k < KMAX; ++k) acc[k] = k < K ? smem[taps_c * K + k] : 0.f;
        for (int r = 0; r < g.kh; ++r) {
            const int h = ho * g.sh - g.pt + r * g.dh;
            if ((unsigned)h >= (unsigned)g.H) continue;
            for (int s = 0; s < g.kw; ++s) {
                const int ww = wo * g.sw - g.pl + s * g.dw;
                if ((unsigned)ww >= (unsigned)g.W) continue;
                const float *px = x + ((long long)(b * g.H + h) * g.W + ww) * C;
                const float *wr = smem + (r * g.kw + s) * C * K;
                for (int c = 0; c < C; ++c) {
                    const float xv = px[c];
#pragma unroll
                    for (int k = 0; k < KMAX; ++k)
                        if (k < K) acc[k] = fmaf(xv, wr[c * K + k], acc[k]);
                }
            }
        }
        float *py = y + (long long)m * K;
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
            if (k < K) py[k] = act_apply(acc[k], act);
    }
}

// Data gradient: a thread per INPUT pixel gathers the output pixels whose windows cover it; C sums in registers;
// epilogue dx = (accumulate ? dx + v : v) * act'(y of the producing layer).
template <int CMAX>
__device__ void fs_conv_dgrad(const fnet_fused_op &op, float *smem) {
    const Geom g = geom_of(op);
    const float *dz = static_cast<const float *>(op.p[0]), *w = static_cast<const float *>(op.p[1]);
    float *dx = static_cast<float *>(op.p[2]);
    const float *yact = static_cast<const float *>(op.p[3]);
    const int K = g.K, C = g.C, taps = g.kh * g.kw, act = op.act, accumulate = op.flags & 1;
    for (int i = threadIdx.x; i < taps * C * K; i += FS_THREADS) smem[i] = w[i];
    __syncthreads();
    const int M = g.B * g.H * g.W;
    for (int m = blockIdx.x * FS_THREADS + threadIdx.x; m < M; m += gridDim.x * FS_THREADS) {
        const int wi = m % g.W, t = m / g.W, hi = t % g.H, b = t / g.H;
        float acc[CMAX];
#pragma unroll
        for (int c = 0; c < CMAX; ++c) acc[c] = 0.f;
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
                const float *pz = dz + ((long long)(b * g.Ho + ho) * g.Wo + wo) * K;
                const float *wr = smem + (r * g.kw + s) * C * K;
                for (int k = 0; k < K; ++k) {
                    const float gz = pz[k];
#pragma unroll
                    for (int c = 0; c < CMAX; ++c)
                        if (c < C) acc[c] = fmaf(gz, wr[c * K + k], acc[c]);
                }
            }
        }
        const long long at = (long long)m * C;
#pragma unroll
        for (int c = 0; c < CMAX; ++c)
            if (c < C) {
                float v = acc[c];
                if (accumulate) v += dx[at + c];
                if (act) v *= act_grad(yact[at + c], act);
                dx[at + c] = v;
            }
    }
}

// Filter gradient: dw[tap][c][k] = sum over output pixels of x[window][c] * dz[pixel][k], last row = bias gradient.
// A thread walks its pixels with all rows x K sums in registers (rows <= RT, K <= KT, RT * KT <= FS_WGRAD_ACC), the
// CTA folds them (shuffle tree, then the warps' slots in warp order) into partial[blockIdx][rows * K].
template <int KT, int RT>
__device__ void fs_conv_wgrad(const fnet_fused_op &op, float *smem) {
    const Geom g = geom_of(op);
    const float *x = static_cast<const float *>(op.p[0]), *dz = static_cast<const float *>(op.p[1]);
    float *partial = static_cast<float *>(op.p[2]);
    const int K = g.K, C = g.C, taps_c = g.kh * g.kw * C;
    const int rows = taps_c + ((op.flags & 1) ? 1 : 0), total = rows * K;
    float acc[RT][KT];
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int k = 0; k < KT; ++k) acc[i][k] = 0.f;
    const int M = g.B * g.Ho * g.Wo;
    for (int m = blockIdx.x * FS_THREADS + threadIdx.x; m < M; m += gridDim.x * FS_THREADS) {
        const int wo = m % g.Wo, t = m / g.Wo, ho = t % g.Ho, b = t / g.Ho;
        const float *img = x + (long long)b * g.H * g.W * C;
        float gz[KT];
#pragma unroll
        for (int k = 0; k < KT; ++k) gz[k] = k < K ? dz[(long long)m * K + k] : 0.f;
        int r = 0, s = 0, c = 0;                    // rows are walked in (tap, channel) order
#pragma unroll
        for (int row = 0; row < RT; ++row) {
            if (row < rows) {
                float xv = 1.f;                     // the bias row
                if (row < taps_c) {
                    const int h = ho * g.sh - g.pt + r * g.dh, ww = wo * g.sw - g.pl + s * g.dw;
                    xv = ((unsigned)h < (unsigned)g.H && (unsigned)ww < (unsigned)g.W) ? img[((long long)h * g.W + ww) * C + c] : 0.f;
                    if (++c == C) { c = 0; if (++s == g.kw) { s = 0; ++r; } }
                }
#pragma unroll
                for (int k = 0; k < KT; ++k) acc[row][k] = fmaf(xv, gz[k], acc[row][k]);
            }
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *red = smem;                              // [FS_THREADS / 32][RT * KT]
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int k = 0; k < KT; ++k) {
            const float v = warp_sum(acc[i][k]);
            if (lane == 0) red[warp * (RT * KT) + i * KT + k] = v;
        }
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += FS_THREADS) {
        const int row = i / K, k = i - row * K;
        float v = 0.f;
#pragma unroll
        for (int wv = 0; wv < FS_THREADS / 32; ++wv) v += red[wv * (RT * KT) + row * KT + k];
        partial[(long long)blockIdx.x * total + i] = v;
    }
}

__device__ void fs_conv_wgrad_any(const fnet_fused_op &op, float *smem) {
    const int K = op.d[6], rows = op.d[7] * op.d[8] * op.d[3] + ((op.flags & 1) ? 1 : 0);
    if (K <= 4) {
        if (rows <= 4) fs_conv_wgrad<4, 4>(op, smem);
        else if (rows <= 8) fs_conv_wgrad<4, 8>(op, smem);
        else fs_conv_wgrad<4, 16>(op, smem);
    } else {
        if (rows <= 4) fs_conv_wgrad<8, 4>(op, smem);
        else fs_conv_wgrad<8, 8>(op, smem);
    }
}

// ------------------------------------------------------------------ the softmax head (N <= 16 classes)
struct SmemRow {
    const float *q;
    __device__ float operator()(int c) const { return q[c]; }
};

// (value, index) argmax over the first `n` lanes of a warp, first index on ties (what a scan with a strict > gives)
__device__ __forceinline__ int warp_argmax(float v, int lane, int n) {
    int idx = lane < n ? lane : 0x7fffffff;
    if (lane >= n) v = -INFINITY;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    return idx;
}

// Dense(softmax) + categorical cross-entropy: one CTA per sample row.  Threads split the K inputs (four at a time, so
// that forty weight loads are in flight per thread), the N sums are folded (shuffle tree, warp slots in order), and the
// first warp finishes the row with one class per lane: softmax as keras does it (exp(z - max) / sum), then loss / hit /
// scale * dL/dlogits with the formulas of fnet_softmax_cce (loss_row.cuh), sums taken by shuffle trees.
__device__ void fs_head_fwd_loss(const fnet_fused_op &op, float *smem) {
    const int M = op.d[0], K = op.d[1], N = op.d[2];
    const float *x = static_cast<const float *>(op.p[0]), *w = static_cast<const float *>(op.p[1]);
    const float *bias = static_cast<const float *>(op.p[2]);
    float *probs = static_cast<float *>(op.p[3]);
    const float *targets = static_cast<const float *>(op.p[4]);
    float *dz = static_cast<float *>(op.p[5]), *row_loss_out = static_cast<float *>(op.p[6]), *row_hit = static_cast<float *>(op.p[7]);
    const float scale = op.f[0];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *red = smem;                       // [16 warps][16]
    for (int m = blockIdx.x; m < M; m += gridDim.x) {
        float acc[16];
#pragma unroll
        for (int n = 0; n < 16; ++n) acc[n] = 0.f;
        const float *xr = x + (long long)m * K;
        for (int k0 = threadIdx.x; k0 < K; k0 += 4 * FS_THREADS) {
            float xv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) xv[u] = k0 + u * FS_THREADS < K ? xr[k0 + u * FS_THREADS] : 0.f;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = k0 + u * FS_THREADS < K ? k0 + u * FS_THREADS : k0;       // a zero times a valid row
                const float *wr = w + (long long)k * N;
#pragma unroll
                for (int n = 0; n < 16; ++n)
                    if (n < N) acc[n] = fmaf(xv[u], wr[n], acc[n]);
            }
        }
#pragma unroll
        for (int n = 0; n < 16; ++n) {
            const float v = warp_sum(acc[n]);
            if (lane == 0) red[warp * 16 + n] = v;
        }
        __syncthreads();
        if (warp == 0) {
            const bool mine = lane < N;
            float zv = mine && bias ? bias[lane] : 0.f;
            const float tv = mine ? targets[(long long)m * N + lane] : 0.f;
            if (mine) {
#pragma unroll
                for (int wv = 0; wv < FS_THREADS / 32; ++wv) zv += red[wv * 16 + lane];
            }
            const float mx = warp_max(mine ? zv : -INFINITY);
            const float e = mine ? expf(zv - mx) : 0.f;
            const float p = e / warp_sum(e);
            if (mine) probs[(long long)m * N + lane] = p;
            // categorical_crossentropy on the probabilities: renormalise, clip, -sum t log q; gradient through all of it
            const float sum = warp_sum(mine ? p : 0.f);
            const bool hit = warp_argmax(p, lane, N) == warp_argmax(tv, lane, N);
            const float q = p / sum;
            const float cl = fminf(fmaxf(q, kClipEps), 1.f - kClipEps);
            const float loss = warp_sum(mine ? -tv * logf(cl) : 0.f);
            const float g = (mine && q >= kClipEps && q <= 1.f - kClipEps) ? -tv / cl : 0.f;
            const float dot = warp_sum(mine ? g * q : 0.f);
            const float dot2 = warp_sum(mine ? (g - dot) / sum * p : 0.f);
            if (mine && dz) dz[(long long)m * N + lane] = scale * p * ((g - dot) / sum - dot2);
            if (lane == 0) {
                row_loss_out[m] = loss;
                row_hit[m] = hit ? 1.f : 0.f;
            }
        }
        __syncthreads();
    }
}

// dx[m][k] = sum_n dz[m][n] * w[k][n]: one CTA per row, the row of dz in shared memory, a thread per k.
__device__ void fs_head_dgrad(const fnet_fused_op &op, float *smem) {
    const int M = op.d[0], K = op.d[1], N = op.d[2];
    const float *dz = static_cast<const float *>(op.p[0]), *w = static_cast<const float *>(op.p[1]);
    float *dx = static_cast<float *>(op.p[2]);
    const float *yact = static_cast<const float *>(op.p[3]);
    const int act = op.act, accumulate = op.flags & 1;
    for (int m = blockIdx.x; m < M; m += gridDim.x) {
        if (threadIdx.x < N) smem[threadIdx.x] = dz[(long long)m * N + threadIdx.x];
        __syncthreads();
        float gz[16];
#pragma unroll
        for (int n = 0; n < 16; ++n) gz[n] = n < N ? smem[n] : 0.f;
        for (int k0 = threadIdx.x; k0 < K; k0 += 4 * FS_THREADS) {
            float v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = k0 + u * FS_THREADS < K ? k0 + u * FS_THREADS : k0;
                const float *wr = w + (long long)k * N;
                v[u] = 0.f;
#pragma unroll
                for (int n = 0; n < 16; ++n)
                    if (n < N) v[u] = fmaf(gz[n], wr[n], v[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = k0 + u * FS_THREADS;
                if (k < K) {
                    const long long at = (long long)m * K + k;
                    float r = v[u];
                    if (accumulate) r += dx[at];
                    if (act) r *= act_grad(yact[at], act);
                    dx[at] = r;
                }
            }
        }
        __syncthreads();
    }
}

// dw[k][n] = sum_m x[m][k] * dz[m][n], dbias[n] = sum_m dz[m][n].  dz (M * N <= 2048 floats) sits in shared memory;
// a CTA takes 32 consecutive k at a time (one per lane), its 16 warps split the rows and are folded in warp order.
__device__ void fs_head_wgrad(const fnet_fused_op &op, float *smem) {
    const int M = op.d[0], K = op.d[1], N = op.d[2];
    const float *x = static_cast<const float *>(op.p[0]), *dz = static_cast<const float *>(op.p[1]);
    float *dw = static_cast<float *>(op.p[2]), *dbias = static_cast<float *>(op.p[3]);
    float *dzs = smem;                       // [M][N]
    float *red = smem + 2048;                // [16 warps][32 lanes][16]
    for (int i = threadIdx.x; i < M * N; i += FS_THREADS) dzs[i] = dz[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunks = (K + 31) >> 5;
    for (int chunk = blockIdx.x; chunk < chunks; chunk += gridDim.x) {
        const int k = chunk * 32 + lane;
        float acc[16];
#pragma unroll
        for (int n = 0; n < 16; ++n) acc[n] = 0.f;
        if (k < K)
            for (int m = warp; m < M; m += FS_THREADS / 32) {
                const float xv = x[(long long)m * K + k];
                const float *g = dzs + m * N;
#pragma unroll
                for (int n = 0; n < 16; ++n)
                    if (n < N) acc[n] = fmaf(xv, g[n], acc[n]);
            }
#pragma unroll
        for (int n = 0; n < 16; ++n) red[(warp * 32 + lane) * 16 + n] = acc[n];
        __syncthreads();
        {
            const int l = threadIdx.x >> 4, n = threadIdx.x & 15;        // 32 lanes x 16 columns = 512 threads
            const int kk = chunk * 32 + l;
            if (n < N && kk < K) {
                float v = 0.f;
#pragma unroll
                for (int wv = 0; wv < FS_THREADS / 32; ++wv) v += red[(wv * 32 + l) * 16 + n];
                dw[(long long)kk * N + n] = v;
            }
        }
        __syncthreads();
    }
    if (blockIdx.x == gridDim.x - 1 && dbias && threadIdx.x < N) {
        float v = 0.f;
        for (int m = 0; m < M; ++m) v += dzs[m * N + threadIdx.x];
        dbias[threadIdx.x] = v;
    }
}

// ------------------------------------------------------------------ optimizer + end of the step
__device__ void fs_close(const fnet_fused_op &op) {
    // first warp of CTA 0: the batch's loss and hits (lane l takes rows l, l + 32, ...; then a shuffle tree: a fixed
    // order) into the running accumulators, iteration += 1, cursor += advance
    if (blockIdx.x != 0 || threadIdx.x >= 32) return;
    int32_t *state = static_cast<int32_t *>(op.p[4]);
    const int rows = op.d[2], lane = threadIdx.x;
    if (rows > 0) {
        const float *row_loss_in = static_cast<const float *>(op.p[5]), *row_hit = static_cast<const float *>(op.p[6]);
        double loss = 0.0;
        int hits = 0;
        for (int m = lane; m < rows; m += 32) { loss += (double)row_loss_in[m]; hits += row_hit[m] != 0.f ? 1 : 0; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            loss += __shfl_xor_sync(0xffffffffu, loss, o);
            hits += __shfl_xor_sync(0xffffffffu, hits, o);
        }
        if (lane == 0) {
            double *loss_sum = static_cast<double *>(op.p[7]);
            long long *correct = reinterpret_cast<long long *>(loss_sum + 1);
            *loss_sum += loss;
            *correct += hits;
        }
    }
    if (lane == 0 && op.d[1] >= 0) { state[0] += 1; state[1] += op.d[1]; }
}

__device__ void fs_adam(const fnet_fused_op &op) {
    const long long n = op.d[0];
    float *p = static_cast<float *>(op.p[0]), *m = static_cast<float *>(op.p[2]), *v = static_cast<float *>(op.p[3]);
    const float *g = static_cast<const float *>(op.p[1]);
    const int32_t *state = static_cast<const int32_t *>(op.p[4]);
    const float lr_t = __int_as_float(state[3]);
    const float b1 = op.f[0], b2 = op.f[1], eps = op.f[2], c1 = 1.f - b1, c2 = 1.f - b2;
    for (long long i = (long long)blockIdx.x * FS_THREADS + threadIdx.x; i < n; i += (long long)gridDim.x * FS_THREADS) {
        const float gg = g[i];
        const float mm = b1 * m[i] + c1 * gg;
        const float vv = b2 * v[i] + c2 * (gg * gg);
        m[i] = mm;
        v[i] = vv;
        p[i] = p[i] - lr_t * mm / (sqrtf(vv) + eps);
    }
    fs_close(op);
}

__device__ void fs_sgd(const fnet_fused_op &op) {
    const long long n = op.d[0];
    float *p = static_cast<float *>(op.p[0]), *vel = static_cast<float *>(op.p[2]);
    const float *g = static_cast<const float *>(op.p[1]);
    const int32_t *state = static_cast<const int32_t *>(op.p[4]);
    const float lr = __int_as_float(state[2]), momentum = op.f[0];
    const int nesterov = op.d[3];
    for (long long i = (long long)blockIdx.x * FS_THREADS + threadIdx.x; i < n; i += (long long)gridDim.x * FS_THREADS) {
        const float gg = g[i];
        const float vv = momentum * vel[i] - lr * gg;
        vel[i] = vv;
        p[i] = nesterov ? p[i] + momentum * vv - lr * gg : p[i] + vv;
    }
    fs_close(op);
}

// ------------------------------------------------------------------ the kernel
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// stamps (optional, profiling): CTA 0 writes the global timer at kernel entry and after every operation
__global__ void __launch_bounds__(FS_THREADS, 1) fused_step_kernel(const fnet_fused_op *__restrict__ ops, int n_ops, unsigned *sync,
                                                                   unsigned long long *stamps) {
    extern __shared__ __align__(16) float smem[];
    __shared__ fnet_fused_op list[FS_MAX_OPS];
    // the list is constant for the lifetime of the program: one trip to memory for all of it, not one per operation
    for (int i = threadIdx.x; i < n_ops * (int)(sizeof(fnet_fused_op) / 4); i += FS_THREADS)
        reinterpret_cast<uint32_t *>(list)[i] = __ldg(reinterpret_cast<const uint32_t *>(ops) + i);
    __syncthreads();
    if (stamps && blockIdx.x == 0 && threadIdx.x == 0) stamps[0] = global_ns();
    for (int i = 0; i < n_ops; ++i) {
        const fnet_fused_op &op = list[i];
        switch (op.kind) {
            case FNET_FK_BEGIN:
                if (blockIdx.x == 0 && threadIdx.x == 0) lr_t_update(static_cast<int32_t *>(op.p[0]), op.f[0], op.f[1]);
                break;
            case FNET_FK_GATHER: fs_gather(op); break;
            case FNET_FK_FEED_U8: fs_feed_u8(op); break;
            case FNET_FK_CONV_FWD:
                if (op.d[6] <= 8) fs_conv_fwd<8>(op, smem);
                else fs_conv_fwd<16>(op, smem);
                break;
            case FNET_FK_CONV_DGRAD:
                if (op.d[3] <= 8) fs_conv_dgrad<8>(op, smem);
                else fs_conv_dgrad<16>(op, smem);
                break;
            case FNET_FK_CONV_WGRAD: fs_conv_wgrad_any(op, smem); break;
            case FNET_FK_REDUCE: fs_reduce(op); break;
            case FNET_FK_CONCAT_FWD: fs_concat_fwd(op); break;
            case FNET_FK_CONCAT_BWD: fs_concat_bwd(op); break;
            case FNET_FK_HEAD_FWD_LOSS: fs_head_fwd_loss(op, smem); break;
            case FNET_FK_HEAD_DGRAD: fs_head_dgrad(op, smem); break;
            case FNET_FK_HEAD_WGRAD: fs_head_wgrad(op, smem); break;
            case FNET_FK_ACT_BWD: fs_act_bwd(op); break;
            case FNET_FK_ADAM: fs_adam(op); break;
            case FNET_FK_SGD: fs_sgd(op); break;
            default: break;
        }
        if (op.barrier_after) grid_barrier(sync);       // uniform: every CTA reads the same list
        else __syncthreads();                            // shared memory (op, staging) is reused by the next operation
        if (stamps && blockIdx.x == 0 && threadIdx.x == 0) stamps[i + 1] = global_ns();
    }
}

// The head as a launch of its own (the default, unfused path): Dense(softmax) + loss in ONE kernel instead of
// dense_fwd + softmax_cce.  The CTA that finishes last (a ticket) adds the batch's loss and hits, in row order, to the
// running accumulators and re-arms the ticket.
__global__ void __launch_bounds__(FS_THREADS) head_fwd_loss_kernel(fnet_fused_op op, double *loss_sum, long long *correct,
                                                                   unsigned *ticket) {
    pdl_prologue();
    extern __shared__ __align__(16) float smem[];
    __shared__ bool last;
    fs_head_fwd_loss(op, smem);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last || threadIdx.x >= 32) return;
    __threadfence();
    const int rows = op.d[0], lane = threadIdx.x;
    const float *row_loss_in = static_cast<const float *>(op.p[6]), *row_hit = static_cast<const float *>(op.p[7]);
    double loss = 0.0;
    int hits = 0;
    for (int m = lane; m < rows; m += 32) {
        loss += (double)*reinterpret_cast<const volatile float *>(row_loss_in + m);
        hits += *reinterpret_cast<const volatile float *>(row_hit + m) != 0.f ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        loss += __shfl_xor_sync(0xffffffffu, loss, o);
        hits += __shfl_xor_sync(0xffffffffu, hits, o);
    }
    if (lane == 0) {
        *loss_sum += loss;
        *correct += hits;
        *ticket = 0;
    }
}

static bool fused_geom_ok(const fnet_conv2d_desc *d) {
    return d && d->batch > 0 && d->in_c >= 1 && d->in_c <= 16 && d->out_c >= 1 && d->out_c <= 16 && d->kh >= 1 && d->kw >= 1 &&
           (long long)d->batch * d->in_h * d->in_w < 0x7fffffffLL / 16 && (long long)d->batch * d->out_h * d->out_w < 0x7fffffffLL / 16;
}

}  // namespace fnet

using namespace fnet;

extern "C" {

int fnet_fused_step_grid(void) {
    int device = 0, sms = 0;
    if (cudaGetDevice(&device) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return 0;
    return sms;
}

int fnet_fused_conv_supported(const fnet_conv2d_desc *d, int product, int with_bias) {
    if (!fused_geom_ok(d)) return 0;
    const long long taps = (long long)d->kh * d->kw, taps_c = taps * d->in_c;
    if (product == 0) return (taps_c + 1) * d->out_c <= FS_SMEM_FLOATS ? 1 : 0;
    if (product == 1) return taps_c * d->out_c <= FS_SMEM_FLOATS ? 1 : 0;
    const long long rows = taps_c + (with_bias ? 1 : 0);
    if (d->out_c <= 4) return rows <= 16 ? 1 : 0;
    return (d->out_c <= 8 && rows <= 8) ? 1 : 0;
}

int fnet_fused_head_supported(int64_t m, int64_t k, int64_t n) {
    return (m >= 1 && n >= 1 && n <= 16 && m * n <= 2048 && k >= 1 && m * k < 0x7fffffffLL) ? 1 : 0;
}

size_t fnet_head_fwd_loss_workspace_bytes(int64_t m) { return m > 0 ? (size_t)(2 * ((m + 63) / 64) * 64) * sizeof(float) : 0; }

int fnet_head_fwd_loss(int64_t m, int64_t k, int64_t n, const float *x, const float *w, const float *bias, float *probs,
                       const float *targets, float scale, float *dlogits, double *loss_sum, int64_t *correct, uint32_t *ticket,
                       void *ws, size_t ws_bytes, void *stream) {
    FNET_REQUIRE(fnet_fused_head_supported(m, k, n), "fnet_head_fwd_loss: shape m=%lld k=%lld n=%lld not supported (n <= 16, m * n <= 2048)",
                 (long long)m, (long long)k, (long long)n);
    FNET_REQUIRE(x && w && probs && targets && loss_sum && correct && ticket, "fnet_head_fwd_loss: null pointer");
    const size_t need = fnet_head_fwd_loss_workspace_bytes(m);
    if (!ws || ws_bytes < need) { set_error("fnet_head_fwd_loss: workspace of %zu bytes needed, %zu given", need, ws_bytes); return FNET_ERR_WORKSPACE; }
    fnet_fused_op op = {};
    op.kind = FNET_FK_HEAD_FWD_LOSS;
    op.d[0] = (int32_t)m; op.d[1] = (int32_t)k; op.d[2] = (int32_t)n;
    op.f[0] = scale;
    float *rows = reinterpret_cast<float *>(ws);
    op.p[0] = const_cast<float *>(x); op.p[1] = const_cast<float *>(w); op.p[2] = const_cast<float *>(bias); op.p[3] = probs;
    op.p[4] = const_cast<float *>(targets); op.p[5] = dlogits; op.p[6] = rows; op.p[7] = rows + ((m + 63) / 64) * 64;
    launch_pdl(head_fwd_loss_kernel, dim3((unsigned)m), dim3(FS_THREADS), 2048, as_stream(stream), op, loss_sum,
               reinterpret_cast<long long *>(correct), reinterpret_cast<unsigned *>(ticket));
    return check_launch("fnet_head_fwd_loss");
}

int fnet_fused_step(const fnet_fused_op *ops_dev, int32_t n_ops, uint32_t *sync, void *stream) {
    return fnet_fused_step_timed(ops_dev, n_ops, sync, nullptr, stream);
}

int fnet_fused_step_timed(const fnet_fused_op *ops_dev, int32_t n_ops, uint32_t *sync, uint64_t *stamps, void *stream) {
    FNET_REQUIRE(ops_dev != nullptr && sync != nullptr && n_ops > 0, "fnet_fused_step: null operation list or barrier word");
    FNET_REQUIRE(n_ops <= FS_MAX_OPS, "fnet_fused_step: at most %d operations per step", FS_MAX_OPS);
    const int grid = fnet_fused_step_grid();
    FNET_REQUIRE(grid > 0, "fnet_fused_step: no device");
    static std::atomic<uint64_t> checked{0};             // one bit per device: the whole grid must be co-resident
    int device = 0;
    FNET_CUDA(cudaGetDevice(&device));
    const size_t smem = FS_SMEM_FLOATS * sizeof(float);
    if (!(checked.load(std::memory_order_acquire) & (1ull << (device & 63)))) {
        int per_sm = 0;
        FNET_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused_step_kernel, FS_THREADS, smem));
        if (per_sm < 1) { set_error("fnet_fused_step: the kernel does not fit an SM"); return FNET_ERR_UNSUPPORTED; }
        checked.fetch_or(1ull << (device & 63), std::memory_order_release);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(FS_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = as_stream(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;         // all CTAs resident at once, or the launch fails: the barrier cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    unsigned *counter = sync;
    (void)cudaLaunchKernelEx(&cfg, fused_step_kernel, ops_dev, (int)n_ops, counter, reinterpret_cast<unsigned long long *>(stamps));
    return check_launch("fnet_fused_step");
}

}  // extern "C"

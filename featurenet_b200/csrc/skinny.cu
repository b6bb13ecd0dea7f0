// Direct kernels for the two GEMM shapes that every FeatureNet candidate has and that no tile fits:
//
//  * the classifier head: Dense(flatten -> n_classes) with N <= 16 columns and K = 6 144 ... 524 288
//    (model/keras_model.py:121-124).  Forward = one CTA per sample, threads stride over K with N
//    accumulators each; filter gradient = one thread per weight row k (coalesced over k, dz staged in
//    shared memory); data gradient = one thread per dx element.  One launch each, no workspace, a fixed
//    reduction order.
//  * convolutions with few channels on both sides (out <= 16, taps*in <= 256), the only kind the
//    main_1block_nas vocabulary produces (6 features) and the first layers of everything else:
//    one thread per pixel with the whole filter in shared memory -- no tiles, no barriers in the main
//    loop, fully coalesced stores.  The filter gradient of the smallest ones ((taps*in+1)*out <= 64)
//    keeps every partial sum in registers and reduces deterministically through the workspace.
//
// All of them are memory- or latency-bound by construction (<= 32 flops per byte moved).
#include "igemm_problems.cuh"

namespace fnet {

constexpr int kSkinnyN = 16;

// ------------------------------------------------------------------ dense, N <= 16
// one CTA of 1024 threads per sample: 32 warps keep ~4 independent k-steps each in flight, which is what hides
// the L2 latency of the weight rows (the whole reduction is only K/1024 steps per thread)
template <int NMAX>
__global__ void __launch_bounds__(1024) dense_skinny_fwd_kernel(int K, int N, int act, const float *__restrict__ x,
                                                                const float *__restrict__ w, const float *__restrict__ bias,
                                                                float *__restrict__ y) {
    pdl_prologue();
    __shared__ float red[32][NMAX];
    const int m = blockIdx.x;
    const float *xr = x + (long long)m * K;
    float acc[NMAX];
#pragma unroll
    for (int n = 0; n < NMAX; ++n) acc[n] = 0.f;
    if ((N & 1) == 0 && (reinterpret_cast<uintptr_t>(w) & 7) == 0) {
        // even N: a weight row is N/2 aligned float2 loads (10 classes: 5 instead of 10 load instructions)
#pragma unroll 4
        for (int k = threadIdx.x; k < K; k += 1024) {
            float xv = __ldg(xr + k);
            const float2 *wr = reinterpret_cast<const float2 *>(w + (long long)k * N);
#pragma unroll
            for (int n = 0; n < NMAX / 2; ++n)
                if (2 * n < N) {
                    float2 wv = __ldg(wr + n);
                    acc[2 * n] = fmaf(xv, wv.x, acc[2 * n]);
                    acc[2 * n + 1] = fmaf(xv, wv.y, acc[2 * n + 1]);
                }
        }
    } else {
#pragma unroll 4
        for (int k = threadIdx.x; k < K; k += 1024) {
            float xv = __ldg(xr + k);
            const float *wr = w + (long long)k * N;
#pragma unroll
            for (int n = 0; n < NMAX; ++n)
                if (n < N) acc[n] = fmaf(xv, __ldg(wr + n), acc[n]);
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int n = 0; n < NMAX; ++n) {
        float v = warp_sum(acc[n]);
        if (lane == 0) red[warp][n] = v;
    }
    __syncthreads();
    if (threadIdx.x < N) {
        float v = 0.f;
#pragma unroll
        for (int wv = 0; wv < 32; ++wv) v += red[wv][threadIdx.x];
        if (bias) v += __ldg(bias + threadIdx.x);
        y[(long long)m * N + threadIdx.x] = act_apply(v, act);
    }
}

// dW[k][n] = sum_m x[m][k] * dz[m][n];  block 0 also writes dbias[n] = sum_m dz[m][n]
template <int NMAX>
__global__ void __launch_bounds__(256) dense_skinny_wgrad_kernel(int M, int K, int N, const float *__restrict__ x,
                                                                 const float *__restrict__ dz, float *__restrict__ dw,
                                                                 float *__restrict__ dbias) {
    pdl_prologue();
    // a CTA owns 32 consecutive weight rows k; its 8 warps split the samples (warp g takes rows g, g+8, ...), so a
    // warp reads 128 contiguous bytes of x per sample and there are K/32 CTAs instead of K/256
    extern __shared__ float dz_s[];                 // a chunk of dz rows, [rows][N]
    __shared__ float red[8][32][NMAX + 1];
    const int kk = threadIdx.x & 31, mg = threadIdx.x >> 5;
    const int k = blockIdx.x * 32 + kk;
    float acc[NMAX];
#pragma unroll
    for (int n = 0; n < NMAX; ++n) acc[n] = 0.f;
    float bsum = 0.f;                               // thread n < N of block 0 sums column n
    constexpr int CHUNK = 256;
    for (int m0 = 0; m0 < M; m0 += CHUNK) {
        int rows = min(CHUNK, M - m0);
        __syncthreads();
        for (int i = threadIdx.x; i < rows * N; i += 256) dz_s[i] = __ldg(dz + (long long)m0 * N + i);
        __syncthreads();
        if (k < K) {
#pragma unroll 4
            for (int r = mg; r < rows; r += 8) {
                float xv = __ldg(x + (long long)(m0 + r) * K + k);
#pragma unroll
                for (int n = 0; n < NMAX; ++n)
                    if (n < N) acc[n] = fmaf(xv, dz_s[r * N + n], acc[n]);
            }
        }
        if (dbias && blockIdx.x == 0 && threadIdx.x < N)
            for (int r = 0; r < rows; ++r) bsum += dz_s[r * N + threadIdx.x];
    }
#pragma unroll
    for (int n = 0; n < NMAX; ++n) red[mg][kk][n] = acc[n];
    __syncthreads();
    if (mg == 0 && k < K) {
#pragma unroll
        for (int n = 0; n < NMAX; ++n) {
            if (n < N) {
                float v = 0.f;
#pragma unroll
                for (int g2 = 0; g2 < 8; ++g2) v += red[g2][kk][n];      // fixed order: deterministic
                dw[(long long)k * N + n] = v;
            }
        }
    }
    if (dbias && blockIdx.x == 0 && threadIdx.x < N) dbias[threadIdx.x] = bsum;
}

// dx[m][k] (+)= sum_n dz[m][n] * w[k][n]
template <int NMAX>
__global__ void __launch_bounds__(256) dense_skinny_dgrad_kernel(int M, int K, int N, const float *__restrict__ dz,
                                                                 const float *__restrict__ w, float *__restrict__ dx,
                                                                 int accumulate) {
    pdl_prologue();
    const int m = blockIdx.y;
    float g[NMAX];
#pragma unroll
    for (int n = 0; n < NMAX; ++n) g[n] = n < N ? __ldg(dz + (long long)m * N + n) : 0.f;
    for (int k = blockIdx.x * 256 + threadIdx.x; k < K; k += gridDim.x * 256) {
        const float *wr = w + (long long)k * N;
        float v = 0.f;
#pragma unroll
        for (int n = 0; n < NMAX; ++n)
            if (n < N) v = fmaf(g[n], __ldg(wr + n), v);
        long long at = (long long)m * K + k;
        dx[at] = accumulate ? dx[at] + v : v;
    }
}

// ------------------------------------------------------------------ dense, N <= 16, streaming variants
// The "keras" template flattens 32x32x512 activations into a 524 288-wide head: x is 268 MB per batch, and each
// of the three products has to stream it (or dx) through the SMs exactly once while the 21 MB weight matrix is
// read from DRAM once and reused out of shared memory / registers.  Lanes always sit on consecutive k, so every
// x / dx access of a warp is one 128-byte line; dz (a few KB) is staged in shared memory, zero-padded to NMAX
// columns, and read back as broadcast float4s.

constexpr int kStreamMinK = 32768;      // heads at least this long take the streaming kernels
constexpr int kStreamChunk = 512;       // weight rows staged per step of the forward kernel
constexpr int kStreamRows = 64;         // samples per CTA of the forward kernel: 16 warps x 4 samples

// forward: partial[b][m][n] = sum over the k-chunks CTA b walks of x[m][k] * w[k][n]
template <int NMAX>
__global__ void __launch_bounds__(512) dense_stream_fwd_kernel(int M, int K, int N, const float *__restrict__ x,
                                                               const float *__restrict__ w, float *__restrict__ partial) {
    pdl_prologue();
    __shared__ float w_s[kStreamChunk * (NMAX + 1)];          // odd row stride: conflict-free when lanes walk k
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int m0 = blockIdx.y * kStreamRows + warp * 4;
    float acc[4][NMAX];
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int n = 0; n < NMAX; ++n) acc[s][n] = 0.f;
    for (int i = threadIdx.x; i < kStreamChunk * (NMAX + 1); i += 512) w_s[i] = 0.f;   // padding columns stay zero
    const float *xr[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) xr[s] = x + (long long)min(m0 + s, M - 1) * K;         // clamped rows are never stored
    for (int kb = blockIdx.x * kStreamChunk; kb < K; kb += gridDim.x * kStreamChunk) {
        const int kc = min(kStreamChunk, K - kb);
        __syncthreads();
        for (int i = threadIdx.x; i < kc * N; i += 512) {
            int kk = i / N, n = i - kk * N;
            w_s[kk * (NMAX + 1) + n] = __ldg(w + (long long)kb * N + i);
        }
        __syncthreads();
#pragma unroll 4                                    // 16 loads of x in flight per thread: the kernel is bound by bytes in flight
        for (int kk = lane; kk < kc; kk += 32) {
            float xv[4];
#pragma unroll
            for (int s = 0; s < 4; ++s) xv[s] = __ldg(xr[s] + kb + kk);
#pragma unroll
            for (int n = 0; n < NMAX; ++n) {
                float wv = w_s[kk * (NMAX + 1) + n];
#pragma unroll
                for (int s = 0; s < 4; ++s) acc[s][n] = fmaf(xv[s], wv, acc[s][n]);
            }
        }
    }
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int n = 0; n < NMAX; ++n) {
            float v = warp_sum(acc[s][n]);
            if (lane == 0 && m0 + s < M && n < N) partial[((long long)blockIdx.x * M + m0 + s) * N + n] = v;
        }
}

// The same with FOUR consecutive k per lane (K % 4 == 0, x 16-byte aligned): x arrives 16 bytes per lane and sample, all
// sixteen loads of a 512-row chunk (4 samples x 4 trips) in flight at once -- 256 bytes per thread instead of 64; the
// kernel is bound by the bytes it keeps in flight.  The staged weights are transposed ([n][k], rows padded by four) so
// that a lane reads its four k of one column as one conflict-free 16-byte load.
template <int NMAX>
__global__ void __launch_bounds__(512) dense_stream_fwd4_kernel(int M, int K, int N, const float *__restrict__ x,
                                                                const float *__restrict__ w, float *__restrict__ partial) {
    pdl_prologue();
    constexpr int PITCH = kStreamChunk + 4;
    __shared__ __align__(16) float w_t[NMAX * PITCH];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int m0 = blockIdx.y * kStreamRows + warp * 4;
    float acc[4][NMAX];
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int n = 0; n < NMAX; ++n) acc[s][n] = 0.f;
    const float4 *xr[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) xr[s] = reinterpret_cast<const float4 *>(x + (long long)min(m0 + s, M - 1) * K);   // clamped rows are never stored
    for (int kb = blockIdx.x * kStreamChunk; kb < K; kb += gridDim.x * kStreamChunk) {
        const int kc = min(kStreamChunk, K - kb);
        __syncthreads();
        for (int i = threadIdx.x; i < NMAX * PITCH; i += 512) w_t[i] = 0.f;          // padding columns and a short last chunk stay zero
        __syncthreads();
        for (int i = threadIdx.x; i < kc * N; i += 512) {
            int kk = i / N, n = i - kk * N;
            w_t[n * PITCH + kk] = __ldg(w + (long long)kb * N + i);
        }
        __syncthreads();
        float4 xv[4][4];                                      // [trip][sample]
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int kk = 4 * lane + 128 * j;
#pragma unroll
            for (int s = 0; s < 4; ++s) xv[j][s] = kk < kc ? __ldg(xr[s] + ((kb + kk) >> 2)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int kk = 4 * lane + 128 * j;
#pragma unroll
            for (int n = 0; n < NMAX; ++n) {
                const float4 wv = *reinterpret_cast<const float4 *>(&w_t[n * PITCH + kk]);
#pragma unroll
                for (int s = 0; s < 4; ++s) {
                    acc[s][n] = fmaf(xv[j][s].x, wv.x, acc[s][n]);
                    acc[s][n] = fmaf(xv[j][s].y, wv.y, acc[s][n]);
                    acc[s][n] = fmaf(xv[j][s].z, wv.z, acc[s][n]);
                    acc[s][n] = fmaf(xv[j][s].w, wv.w, acc[s][n]);
                }
            }
        }
    }
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int n = 0; n < NMAX; ++n) {
            float v = warp_sum(acc[s][n]);
            if (lane == 0 && m0 + s < M && n < N) partial[((long long)blockIdx.x * M + m0 + s) * N + n] = v;
        }
}

// filter gradient: one thread per weight row k, all samples in sequence -- no cross-thread reduction at all
template <int NMAX>
__global__ void __launch_bounds__(256) dense_stream_wgrad_kernel(int M, int K, int N, const float *__restrict__ x,
                                                                 const float *__restrict__ dz, float *__restrict__ dw,
                                                                 float *__restrict__ dbias) {
    pdl_prologue();
    constexpr int ROWS = 128;
    __shared__ __align__(16) float dz_s[ROWS][NMAX];
    const int k = blockIdx.x * 256 + threadIdx.x;
    const float *xk = x + min(k, K - 1);
    float acc[NMAX];
#pragma unroll
    for (int n = 0; n < NMAX; ++n) acc[n] = 0.f;
    float bsum = 0.f;
    for (int m0 = 0; m0 < M; m0 += ROWS) {
        const int rows = min(ROWS, M - m0);
        __syncthreads();
        for (int i = threadIdx.x; i < ROWS * NMAX; i += 256) {
            int r = i / NMAX, n = i - r * NMAX;
            dz_s[r][n] = (r < rows && n < N) ? __ldg(dz + (long long)(m0 + r) * N + n) : 0.f;
        }
        __syncthreads();
#pragma unroll 16
        for (int r = 0; r < rows; ++r) {
            float xv = __ldg(xk + (long long)(m0 + r) * K);
#pragma unroll
            for (int q = 0; q < NMAX / 4; ++q) {
                float4 d = *reinterpret_cast<const float4 *>(&dz_s[r][4 * q]);
                acc[4 * q] = fmaf(xv, d.x, acc[4 * q]);
                acc[4 * q + 1] = fmaf(xv, d.y, acc[4 * q + 1]);
                acc[4 * q + 2] = fmaf(xv, d.z, acc[4 * q + 2]);
                acc[4 * q + 3] = fmaf(xv, d.w, acc[4 * q + 3]);
            }
        }
        if (dbias && blockIdx.x == 0 && threadIdx.x < N)
            for (int r = 0; r < rows; ++r) bsum += dz_s[r][threadIdx.x];
    }
    if (k < K) {
#pragma unroll
        for (int n = 0; n < NMAX; ++n)
            if (n < N) dw[(long long)k * N + n] = acc[n];
    }
    if (dbias && blockIdx.x == 0 && threadIdx.x < N) dbias[threadIdx.x] = bsum;
}

// data gradient (every K): a thread keeps its weight row in registers and walks the 16 samples of its CTA
template <int NMAX, bool EPI>
__global__ void __launch_bounds__(256) dense_rows_dgrad_kernel(int M, int K, int N, const float *__restrict__ dz,
                                                               const float *__restrict__ w, float *__restrict__ dx,
                                                               int accumulate, int act, const float *__restrict__ yact) {
    pdl_prologue();
    constexpr int ROWS = 16;
    __shared__ __align__(16) float g_s[ROWS][NMAX];
    const int m0 = blockIdx.y * ROWS, rows = min(ROWS, M - m0);
    for (int i = threadIdx.x; i < ROWS * NMAX; i += 256) {
        int r = i / NMAX, n = i - r * NMAX;
        g_s[r][n] = (r < rows && n < N) ? __ldg(dz + (long long)(m0 + r) * N + n) : 0.f;
    }
    __syncthreads();
    for (int k = blockIdx.x * 256 + threadIdx.x; k < K; k += gridDim.x * 256) {
        float wr[NMAX];
#pragma unroll
        for (int n = 0; n < NMAX; ++n) wr[n] = n < N ? __ldg(w + (long long)k * N + n) : 0.f;
        float *out = dx + (long long)m0 * K + k;
        const float *yin = yact ? yact + (long long)m0 * K + k : nullptr;
        // four rows at a time (eight measured slower: 205 -> 254 us), what the epilogue reads (y of the producing layer, the old dx) fetched BEFORE the sums:
        // with the loads inside the row loop every row waited for its own (ncu: 23 stall cycles per issue on the long
        // scoreboard, 289 us for the 524 288-wide head of the `keras` template against 98 us without an epilogue)
        if (!EPI) {                                           // plain store: the row loop as it always was
#pragma unroll 4
            for (int r = 0; r < rows; ++r) {
                float v = 0.f;
#pragma unroll
                for (int q = 0; q < NMAX / 4; ++q) {
                    float4 d = *reinterpret_cast<const float4 *>(&g_s[r][4 * q]);
                    v = fmaf(d.x, wr[4 * q], v);
                    v = fmaf(d.y, wr[4 * q + 1], v);
                    v = fmaf(d.z, wr[4 * q + 2], v);
                    v = fmaf(d.w, wr[4 * q + 3], v);
                }
                out[(long long)r * K] = v;
            }
            continue;
        }
        for (int r0 = 0; r0 < rows; r0 += 4) {
            float yv[4], old[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const bool in = r0 + u < rows;
                yv[u] = (act && in) ? __ldg(yin + (long long)(r0 + u) * K) : 0.f;
                old[u] = (accumulate && in) ? out[(long long)(r0 + u) * K] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (r0 + u < rows) {
                    float v = 0.f;
#pragma unroll
                    for (int q = 0; q < NMAX / 4; ++q) {
                        float4 d = *reinterpret_cast<const float4 *>(&g_s[r0 + u][4 * q]);
                        v = fmaf(d.x, wr[4 * q], v);
                        v = fmaf(d.y, wr[4 * q + 1], v);
                        v = fmaf(d.z, wr[4 * q + 2], v);
                        v = fmaf(d.w, wr[4 * q + 3], v);
                    }
                    v += old[u];
                    if (act) v *= act_grad(yv[u], act);
                    out[(long long)(r0 + u) * K] = v;
                }
            }
        }
    }
}

// The same for long heads (K % 4 == 0, 16-byte aligned tensors): FOUR consecutive inputs k per thread, so every access
// to dx and to the producing layer's output is 16 bytes per lane (a warp instruction moves 512 bytes instead of 128);
// the four weight rows (4 x NMAX registers) are read once for the 16 samples.  dx of a 524 288-wide head is 268 MB.
template <int NMAX, bool EPI>
__global__ void __launch_bounds__(256) dense_rows_dgrad4_kernel(int M, int K, int N, const float *__restrict__ dz,
                                                                const float *__restrict__ w, float *__restrict__ dx,
                                                                int accumulate, int act, const float *__restrict__ yact) {
    pdl_prologue();
    constexpr int ROWS = 16;
    __shared__ __align__(16) float g_s[ROWS][NMAX];
    const int m0 = blockIdx.y * ROWS, rows = min(ROWS, M - m0);
    for (int i = threadIdx.x; i < ROWS * NMAX; i += 256) {
        int r = i / NMAX, n = i - r * NMAX;
        g_s[r][n] = (r < rows && n < N) ? __ldg(dz + (long long)(m0 + r) * N + n) : 0.f;
    }
    __syncthreads();
    const int quads = K >> 2;
    for (int kq = blockIdx.x * 256 + threadIdx.x; kq < quads; kq += gridDim.x * 256) {
        float wr[4][NMAX];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int n = 0; n < NMAX; ++n) wr[j][n] = n < N ? __ldg(w + (long long)(4 * kq + j) * N + n) : 0.f;
        float4 *out = reinterpret_cast<float4 *>(dx + (long long)m0 * K) + kq;
        const float4 *yin = yact ? reinterpret_cast<const float4 *>(yact + (long long)m0 * K) + kq : nullptr;
        for (int r0 = 0; r0 < rows; r0 += 4) {
            float4 yv[4], old[4];
            if (EPI) {                                        // what the epilogue reads, fetched before the sums
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const bool in = r0 + u < rows;
                    yv[u] = (act && in) ? __ldg(yin + (long long)(r0 + u) * quads) : make_float4(0.f, 0.f, 0.f, 0.f);
                    old[u] = (accumulate && in) ? out[(long long)(r0 + u) * quads] : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (r0 + u < rows) {
                    float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int q = 0; q < NMAX / 4; ++q) {
                        const float4 d = *reinterpret_cast<const float4 *>(&g_s[r0 + u][4 * q]);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            v[j] = fmaf(d.x, wr[j][4 * q], v[j]);
                            v[j] = fmaf(d.y, wr[j][4 * q + 1], v[j]);
                            v[j] = fmaf(d.z, wr[j][4 * q + 2], v[j]);
                            v[j] = fmaf(d.w, wr[j][4 * q + 3], v[j]);
                        }
                    }
                    float4 o = make_float4(v[0], v[1], v[2], v[3]);
                    if (EPI) {
                        o.x += old[u].x; o.y += old[u].y; o.z += old[u].z; o.w += old[u].w;
                        if (act) {
                            o.x *= act_grad(yv[u].x, act); o.y *= act_grad(yv[u].y, act);
                            o.z *= act_grad(yv[u].z, act); o.w *= act_grad(yv[u].w, act);
                        }
                    }
                    out[(long long)(r0 + u) * quads] = o;
                }
            }
        }
    }
}

static bool is_dense_geom(const ConvGeom &g) { return g.H == 1 && g.W == 1 && g.kh == 1 && g.kw == 1 && g.Ho == 1 && g.Wo == 1; }

static int stream_fwd_blocks(const ConvGeom &g) {
    long long chunks = ceil_div<long long>(g.C, kStreamChunk);
    long long per_row_group = 2LL * kNumSMs / ceil_div<long long>(g.B, kStreamRows);
    long long b = chunks < per_row_group ? chunks : per_row_group;
    return (int)(b < 1 ? 1 : b);
}

// ------------------------------------------------------------------ small-channel convolutions
// forward: one thread per output pixel, all out channels (<= NMAX) in registers
template <int NMAX>
__global__ void __launch_bounds__(256) conv_small_fwd_kernel(ConvGeom g, int act, const float *__restrict__ x,
                                                             const float *__restrict__ w, const float *__restrict__ bias,
                                                             float *__restrict__ y) {
    pdl_prologue();
    // [taps*C + 1][NMAX]: the filter (HWIO) with its rows padded to NMAX columns by zeros, then the bias row.  A row is
    // read as NMAX / 4 LDS.128 (the unpadded layout cost one shared load per FMA: lenet5's 6 -> 12 5x5 ran at 10 % of
    // the FFMA rate)
    extern __shared__ __align__(16) float w_s[];
    const int KG = g.kh * g.kw * g.C, N = g.K;
    for (int i = threadIdx.x; i < (KG + 1) * NMAX; i += 256) {
        const int row = i / NMAX, n = i - row * NMAX;
        w_s[i] = n < N ? (row < KG ? __ldg(w + row * N + n) : (bias ? __ldg(bias + n) : 0.f)) : 0.f;
    }
    __syncthreads();
    const long long M = (long long)g.B * g.Ho * g.Wo;
    for (long long m = (long long)blockIdx.x * 256 + threadIdx.x; m < M; m += (long long)gridDim.x * 256) {
        int wo, ho;
        long long b;
        decode3(m, g.Ho, g.Wo, b, ho, wo);
        const float *img = x + b * g.H * g.W * g.C;
        float acc[NMAX];
#pragma unroll
        for (int n = 0; n < NMAX; ++n) acc[n] = w_s[KG * NMAX + n];
        for (int r = 0; r < g.kh; ++r) {
            int h = ho * g.sh - g.pt + r * g.dh;
            if ((unsigned)h >= (unsigned)g.H) continue;
            for (int s = 0; s < g.kw; ++s) {
                int ww = wo * g.sw - g.pl + s * g.dw;
                if ((unsigned)ww >= (unsigned)g.W) continue;
                const float *px = img + ((long long)h * g.W + ww) * g.C;
                const float4 *wt = reinterpret_cast<const float4 *>(w_s + (r * g.kw + s) * g.C * NMAX);
                for (int c = 0; c < g.C; ++c) {
                    float xv = __ldg(px + c);
#pragma unroll
                    for (int q = 0; q < NMAX / 4; ++q) {
                        const float4 wv = wt[c * (NMAX / 4) + q];
                        acc[4 * q] = fmaf(xv, wv.x, acc[4 * q]);
                        acc[4 * q + 1] = fmaf(xv, wv.y, acc[4 * q + 1]);
                        acc[4 * q + 2] = fmaf(xv, wv.z, acc[4 * q + 2]);
                        acc[4 * q + 3] = fmaf(xv, wv.w, acc[4 * q + 3]);
                    }
                }
            }
        }
        float *out = y + m * N;
#pragma unroll
        for (int n = 0; n < NMAX; ++n)
            if (n < N) out[n] = act_apply(acc[n], act);
    }
}

// data gradient: one thread per input pixel, all in channels (<= NMAX) in registers
template <int NMAX>
__global__ void __launch_bounds__(256) conv_small_dgrad_kernel(ConvGeom g, const float *__restrict__ dz,
                                                               const float *__restrict__ w, float *__restrict__ dx,
                                                               int accumulate, int act, const float *__restrict__ yact) {
    pdl_prologue();
    // [taps][K][NMAX]: the filter transposed so that the input channels a thread accumulates are contiguous (padded to
    // NMAX by zeros): NMAX / 4 LDS.128 per (tap, output channel) instead of one shared load per FMA
    extern __shared__ __align__(16) float w_s[];
    const int taps = g.kh * g.kw, C = g.C, K = g.K;
    for (int i = threadIdx.x; i < taps * K * NMAX; i += 256) {
        const int c = i % NMAX, t2 = i / NMAX, ko = t2 % K, tap = t2 / K;
        w_s[i] = c < C ? __ldg(w + (tap * C + c) * K + ko) : 0.f;
    }
    __syncthreads();
    const long long M = (long long)g.B * g.H * g.W;
    for (long long m = (long long)blockIdx.x * 256 + threadIdx.x; m < M; m += (long long)gridDim.x * 256) {
        int wi, hi;
        long long b;
        decode3(m, g.H, g.W, b, hi, wi);
        const float *img = dz + b * g.Ho * g.Wo * K;
        float acc[NMAX];
#pragma unroll
        for (int c = 0; c < NMAX; ++c) acc[c] = 0.f;
        for (int r = 0; r < g.kh; ++r) {
            int hh = hi + g.pt - r * g.dh;
            if (hh < 0) continue;
            int ho = hh / g.sh;
            if (ho * g.sh != hh || ho >= g.Ho) continue;
            for (int s = 0; s < g.kw; ++s) {
                int ww = wi + g.pl - s * g.dw;
                if (ww < 0) continue;
                int wo = ww / g.sw;
                if (wo * g.sw != ww || wo >= g.Wo) continue;
                const float *pz = img + ((long long)ho * g.Wo + wo) * K;
                const float4 *wt = reinterpret_cast<const float4 *>(w_s + (r * g.kw + s) * K * NMAX);
                for (int ko = 0; ko < K; ++ko) {
                    float gv = __ldg(pz + ko);
#pragma unroll
                    for (int q = 0; q < NMAX / 4; ++q) {
                        const float4 wv = wt[ko * (NMAX / 4) + q];
                        acc[4 * q] = fmaf(gv, wv.x, acc[4 * q]);
                        acc[4 * q + 1] = fmaf(gv, wv.y, acc[4 * q + 1]);
                        acc[4 * q + 2] = fmaf(gv, wv.z, acc[4 * q + 2]);
                        acc[4 * q + 3] = fmaf(gv, wv.w, acc[4 * q + 3]);
                    }
                }
            }
        }
        float *out = dx + m * C;
#pragma unroll
        for (int c = 0; c < NMAX; ++c)
            if (c < C) store_dx(out, c, acc[c], accumulate, act, yact ? yact + (out - dx) : nullptr);
    }
}

// Data gradient of a 1x1 / Dense layer with very few OUTPUT channels (the "Reg_" aligners back to 3 channels,
// model/operation.py:186): dx[p][c] = sum over k <= 8 of dz[p][k] * w[c][k].  Nothing to tile: dz is tiny, dx is the
// traffic, so a thread produces four consecutive channels of one pixel and a warp writes 512 contiguous bytes.
// (The implicit-GEMM kernel spent 174 us on 262 144 x 192 <- 3, i.e. 1.2 TB/s for a pure store stream.)
constexpr int kPointwiseK = 8;

__global__ void __launch_bounds__(256) pointwise_dgrad_kernel(int M, int C, int K, const float *__restrict__ dz,
                                                              const float *__restrict__ w, float *__restrict__ dx,
                                                              int accumulate, int act, const float *__restrict__ yact,
                                                              int vec_store) {
    pdl_prologue();
    extern __shared__ __align__(16) float wsm[];             // w transposed to [K][C]: a thread's four channels are one LDS.128
    for (int i = threadIdx.x; i < C * K; i += 256) {
        const int c = i / K, k = i - c * K;
        wsm[k * C + c] = __ldg(w + i);
    }
    __syncthreads();
    const int quads = C >> 2, total = M * quads;
    for (int idx = blockIdx.x * 256 + threadIdx.x; idx < total; idx += gridDim.x * 256) {
        const int p = idx / quads, c = (idx - p * quads) << 2;
        float g[kPointwiseK];
#pragma unroll
        for (int k = 0; k < kPointwiseK; ++k) g[k] = k < K ? __ldg(dz + (long long)p * K + k) : 0.f;
        float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < kPointwiseK; ++k)
            if (k < K) {
                const float4 wv = *reinterpret_cast<const float4 *>(wsm + k * C + c);
                v[0] = fmaf(g[k], wv.x, v[0]);
                v[1] = fmaf(g[k], wv.y, v[1]);
                v[2] = fmaf(g[k], wv.z, v[2]);
                v[3] = fmaf(g[k], wv.w, v[3]);
            }
        const long long at = (long long)p * C + c;
        if (vec_store) {
            // the epilogue with 16-byte accesses as well: with the producer's activation derivative folded in (which is
            // how the engine calls the `Reg_` aligners' data gradient) the scalar path ran at 1.5 TB/s, 137 us on 192 <- 3
            if (accumulate) {
                const float4 o = *reinterpret_cast<const float4 *>(dx + at);
                v[0] += o.x; v[1] += o.y; v[2] += o.z; v[3] += o.w;
            }
            if (act) {
                const float4 yy = __ldg(reinterpret_cast<const float4 *>(yact + at));
                v[0] *= act_grad(yy.x, act); v[1] *= act_grad(yy.y, act); v[2] *= act_grad(yy.z, act); v[3] *= act_grad(yy.w, act);
            }
            *reinterpret_cast<float4 *>(dx + at) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) store_dx(dx, at + j, v[j], accumulate, act, yact);
        }
    }
}

// Forward of the same layers: y[p][k] = act(sum over c of x[p][c] * w[c][k] + b[k]), K <= 8.  LP lanes share a pixel,
// each takes every LP-th group of four channels (one LDG.128 of x, K LDS.128 of the weights, which are stored as
// w[c][k] so that four channels are 4K consecutive floats), then a shuffle tree: x is read once, coalesced.
template <int K, int LP>
__global__ void __launch_bounds__(256) pointwise_fwd_kernel(int M, int C, int act, const float *__restrict__ x,
                                                            const float *__restrict__ w, const float *__restrict__ bias,
                                                            float *__restrict__ y) {
    pdl_prologue();
    extern __shared__ __align__(16) float wsm[];             // w [C][K]
    for (int i = threadIdx.x; i < C * K; i += 256) wsm[i] = __ldg(w + i);
    __syncthreads();
    const int quads = C >> 2, sub = threadIdx.x % LP;
    const int per_block = 256 / LP;
    // the trip count is uniform over the block (the shuffles below need whole warps); pixels past the end do no loads
    for (int base = blockIdx.x * per_block; base < M; base += gridDim.x * per_block) {
        const int p = base + threadIdx.x / LP;
        float acc[K];
#pragma unroll
        for (int k = 0; k < K; ++k) acc[k] = 0.f;
        const float4 *xp = reinterpret_cast<const float4 *>(x + (long long)p * C);
        const int mine = p < M ? quads : 0;
        for (int q0 = sub; q0 < mine; q0 += 4 * LP) {
            float4 xv[4];                                     // four loads in flight before the first use
#pragma unroll
            for (int u = 0; u < 4; ++u) xv[u] = q0 + u * LP < mine ? __ldg(xp + q0 + u * LP) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int q = q0 + u * LP < mine ? q0 + u * LP : sub;      // zeros times any valid weights
                float wf[4 * K];
#pragma unroll
                for (int j = 0; j < K; ++j)
                    *reinterpret_cast<float4 *>(wf + 4 * j) = *reinterpret_cast<const float4 *>(wsm + 4 * q * K + 4 * j);
#pragma unroll
                for (int k = 0; k < K; ++k)
                    acc[k] = fmaf(xv[u].x, wf[k], fmaf(xv[u].y, wf[K + k], fmaf(xv[u].z, wf[2 * K + k], fmaf(xv[u].w, wf[3 * K + k], acc[k]))));
            }
        }
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int o = LP / 2; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
        if (sub == 0 && p < M) {
#pragma unroll
            for (int k = 0; k < K; ++k) y[(long long)p * K + k] = act_apply(acc[k] + (bias ? __ldg(bias + k) : 0.f), act);
        }
    }
}

// Filter gradient of the same layers: dw[c][k] = sum over pixels of x[p][c] * dz[p][k] (+ the bias row).  A thread owns
// four channels and keeps their 4K sums in registers while it walks the pixels; consecutive threads take consecutive
// channel groups of one pixel (coalesced), 256 / (C/4) pixels per block step.  partial[block][(C + 1) * K], fixed-order reduce.
template <int K>
__global__ void __launch_bounds__(256) pointwise_wgrad_kernel(int M, int C, const float *__restrict__ x,
                                                              const float *__restrict__ dz, float *__restrict__ partial) {
    pdl_prologue();
    extern __shared__ __align__(16) float red[];             // [pixel group][quad][4K] (+ [pixel group][K] bias sums)
    const int quads = C >> 2, groups = 256 / quads;
    const int pg = threadIdx.x / quads, q = threadIdx.x - pg * quads;
    float acc[4 * K], bsum[K];
#pragma unroll
    for (int i = 0; i < 4 * K; ++i) acc[i] = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) bsum[k] = 0.f;
    if (pg < groups) {
        const int step = gridDim.x * groups;
        for (int p0 = blockIdx.x * groups + pg; p0 < M; p0 += 4 * step) {
            float4 xv[4];                                     // four pixels per trip: all loads issued before the first FMA
            float g[4][K];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int p = p0 + u * step;
                const bool in = p < M;
                xv[u] = in ? __ldg(reinterpret_cast<const float4 *>(x + (long long)p * C) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int k = 0; k < K; ++k) g[u][k] = in ? __ldg(dz + (long long)p * K + k) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    acc[k] = fmaf(xv[u].x, g[u][k], acc[k]);
                    acc[K + k] = fmaf(xv[u].y, g[u][k], acc[K + k]);
                    acc[2 * K + k] = fmaf(xv[u].z, g[u][k], acc[2 * K + k]);
                    acc[3 * K + k] = fmaf(xv[u].w, g[u][k], acc[3 * K + k]);
                    bsum[k] += g[u][k];
                }
        }
        float *mine = red + ((long long)pg * quads + q) * 4 * K;
#pragma unroll
        for (int i = 0; i < 4 * K; ++i) mine[i] = acc[i];
        if (q == 0) {
#pragma unroll
            for (int k = 0; k < K; ++k) red[groups * quads * 4 * K + pg * K + k] = bsum[k];
        }
    }
    __syncthreads();
    // rows c = 4q + j of dw are red[.][q][j * K + k]: flat index (4q + j) * K + k = q * 4K + (j * K + k), the same order
    const int total = C * K;
    float *out = partial + (long long)blockIdx.x * (total + K);
    for (int i = threadIdx.x; i < total + K; i += 256) {
        float v = 0.f;
        if (i < total) {
            for (int z = 0; z < groups; ++z) v += red[(long long)z * total + i];
        } else {
            for (int z = 0; z < groups; ++z) v += red[groups * total + z * K + (i - total)];
        }
        out[i] = v;
    }
}

constexpr int kPointwiseBlocks = 4 * kNumSMs;

static bool pointwise_shape_ok(const ConvGeom &g) {
    const long long M = (long long)g.B * g.H * g.W;
    return g.kh == 1 && g.kw == 1 && g.sh == 1 && g.sw == 1 && g.pt == 0 && g.pl == 0 && g.Ho == g.H && g.Wo == g.W &&
           g.K <= kPointwiseK && !(g.C & 3) && g.C >= 32 && g.C <= 1024 && (long long)g.C * g.K * 4 <= 40 * 1024 && M >= 4096 &&
           M * (g.C >> 2) < 0x7fffffffLL;
}

static bool pointwise_dgrad_ok(const ConvGeom &g) {
    const long long M = (long long)g.B * g.H * g.W;
    return g.kh == 1 && g.kw == 1 && g.sh == 1 && g.sw == 1 && g.pt == 0 && g.pl == 0 && g.Ho == g.H && g.Wo == g.W &&
           g.K <= kPointwiseK && !(g.C & 3) && g.C >= 32 && (long long)g.C * g.K * 4 <= 40 * 1024 && M >= 4096 &&
           M * (g.C >> 2) < 0x7fffffffLL;
}

// filter gradient of the smallest convolutions: every (tap, c, k) partial sum (and the bias sums) lives in
// registers; partial[block][row][k] goes to the workspace and is reduced in a fixed order afterwards
constexpr int kWgradRegs = 64;
constexpr int kWgradBlocks = 4 * kNumSMs;

// KT = number of output channels, RT = bound on the filter rows (both compile time, so that every partial sum has
// a fixed register and no more registers are held than the layer needs: 142 registers for a 4 x 6 gradient meant
// one CTA per SM and 340 GB/s on a kernel whose only job is to stream x and dz once)
template <int KT, int RT>
__global__ void __launch_bounds__(256) conv_small_wgrad_kernel(ConvGeom g, int rows, const float *__restrict__ x,
                                                               const float *__restrict__ dz, float *__restrict__ partial) {
    pdl_prologue();
    constexpr int ROWS_MAX = RT < kWgradRegs / KT ? RT : kWgradRegs / KT;
    __shared__ float red[8][kWgradRegs];
    const int C = g.C, taps_c = g.kh * g.kw * C;
    const int total = rows * KT;                    // rows = taps*C (+1 bias row) <= ROWS_MAX
    float acc[ROWS_MAX][KT];
#pragma unroll
    for (int i = 0; i < ROWS_MAX; ++i)
#pragma unroll
        for (int k = 0; k < KT; ++k) acc[i][k] = 0.f;
    const long long M = (long long)g.B * g.Ho * g.Wo;
    for (long long m = (long long)blockIdx.x * 256 + threadIdx.x; m < M; m += (long long)gridDim.x * 256) {
        int wo, ho;
        long long b;
        decode3(m, g.Ho, g.Wo, b, ho, wo);
        const float *img = x + b * g.H * g.W * C;
        float gz[KT];
#pragma unroll
        for (int k = 0; k < KT; ++k) gz[k] = __ldg(dz + m * KT + k);
        int r = 0, s = 0, c = 0;                    // rows are walked in (tap, c) order
#pragma unroll
        for (int row = 0; row < ROWS_MAX; ++row) {
            if (row < rows) {
                float xv = 1.f;                     // the bias row
                if (row < taps_c) {
                    int h = ho * g.sh - g.pt + r * g.dh, ww = wo * g.sw - g.pl + s * g.dw;
                    xv = ((unsigned)h < (unsigned)g.H && (unsigned)ww < (unsigned)g.W)
                             ? __ldg(img + ((long long)h * g.W + ww) * C + c) : 0.f;
                    if (++c == C) { c = 0; if (++s == g.kw) { s = 0; ++r; } }
                }
#pragma unroll
                for (int k = 0; k < KT; ++k) acc[row][k] = fmaf(xv, gz[k], acc[row][k]);
            }
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < ROWS_MAX; ++i)
#pragma unroll
        for (int k = 0; k < KT; ++k) {
            float v = warp_sum(acc[i][k]);
            if (lane == 0) red[warp][i * KT + k] = v;
        }
    __syncthreads();
    if (threadIdx.x < total) {
        float v = 0.f;
#pragma unroll
        for (int wv = 0; wv < 8; ++wv) v += red[wv][threadIdx.x];
        partial[(long long)blockIdx.x * total + threadIdx.x] = v;
    }
}

__global__ void __launch_bounds__(256) conv_small_wgrad_reduce_kernel(int blocks, int taps_c, int K, int total,
                                                                      const float *__restrict__ partial,
                                                                      float *__restrict__ dw, float *__restrict__ dbias) {
    pdl_prologue();
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = blockIdx.x * 8 + warp; i < total; i += gridDim.x * 8) {
        float s = 0.f;
        for (int z = lane; z < blocks; z += 32) s += partial[(long long)z * total + i];
        s = warp_sum(s);
        if (lane == 0) {
            if (i < taps_c * K) dw[i] = s;
            else if (dbias) dbias[i - taps_c * K] = s;
        }
    }
}

void launch_small_wgrad_reduce(int blocks, int taps_c, int K, int total, const float *partial, float *dw, float *dbias, cudaStream_t s) {
    launch_pdl(conv_small_wgrad_reduce_kernel, dim3((unsigned)ceil_div<int>(total, 8)), dim3(256), 0, s, blocks, taps_c, K, total, partial, dw, dbias);
}

// ------------------------------------------------------------------ dispatch (called from conv_api.cu)
// "small" = few channels on BOTH sides: a thread walks its pixel's channels serially, which only coalesces
// (through L1) when a pixel is a few dozen bytes
static bool small_fwd_ok(const ConvGeom &g) {
    return !is_dense_geom(g) && g.K <= kSkinnyN && g.C <= kSkinnyN && (long long)g.kh * g.kw * g.C <= 256 &&
           ((long long)g.kh * g.kw * g.C + 1) * g.K * 4 <= 40 * 1024;
}
static bool small_dgrad_ok(const ConvGeom &g) {
    return !is_dense_geom(g) && g.C <= kSkinnyN && g.K <= kSkinnyN && (long long)g.kh * g.kw * g.K <= 256 &&
           (long long)g.kh * g.kw * g.C * g.K * 4 <= 40 * 1024;
}
static bool small_wgrad_ok(const ConvGeom &g, bool with_bias) {
    return !is_dense_geom(g) && g.K <= 8 && ((long long)g.kh * g.kw * g.C + (with_bias ? 1 : 0)) * g.K <= kWgradRegs;
}

bool skinny_supported(const ConvGeom &g, int product, bool with_bias) {
    // head-like only: a long reduction axis and a modest number of rows (one CTA per row / one thread per k)
    if (product == 1 && pointwise_dgrad_ok(g)) return true;
    if (product != 1 && pointwise_shape_ok(g)) return true;
    if (is_dense_geom(g)) return g.K <= kSkinnyN && g.C >= 512 && g.B >= 1 && g.B <= 8192;
    if (product == 0) return small_fwd_ok(g);
    if (product == 1) return small_dgrad_ok(g);
    return small_wgrad_ok(g, with_bias);
}

size_t skinny_workspace(const ConvGeom &g) {
    if (is_dense_geom(g))
        return (skinny_supported(g, 0, true) && g.C >= kStreamMinK) ? (size_t)stream_fwd_blocks(g) * g.B * g.K * sizeof(float) : 0;
    if (pointwise_shape_ok(g)) return (size_t)kPointwiseBlocks * ((size_t)g.C + 1) * g.K * sizeof(float);
    if (small_wgrad_ok(g, true)) return (size_t)kWgradBlocks * kWgradRegs * sizeof(float);
    return 0;
}

static int pixel_grid(long long pixels) {
    long long blocks = ceil_div<long long>(pixels, 256);
    long long cap = (long long)kNumSMs * 8;
    return (int)(blocks > cap ? cap : (blocks < 1 ? 1 : blocks));
}

#define FNET_BY_COLUMNS(N_, CALL)                 \
    do {                                          \
        if ((N_) <= 4) { CALL(4); }               \
        else if ((N_) <= 8) { CALL(8); }          \
        else if ((N_) <= 12) { CALL(12); }        \
        else { CALL(16); }                        \
    } while (0)

int skinny_fwd(const ConvGeom &g, int act, const float *x, const float *w, const float *bias, float *y, void *ws,
               size_t ws_bytes, cudaStream_t s) {
    if (act == FNET_ACT_SOFTMAX) act = FNET_ACT_NONE;      // the caller runs the softmax pass
    if (pointwise_shape_ok(g)) {
        const int M = (int)((long long)g.B * g.H * g.W);
        const size_t smem = (size_t)g.C * g.K * sizeof(float);
        const bool narrow = (g.C >> 2) <= 64;                // <= 256 channels: 16 lanes per pixel keep every lane busy
        const int per_block = narrow ? 16 : 8;
        long long blocks = ceil_div<long long>(M, per_block);
        if (blocks > (long long)kNumSMs * 16) blocks = (long long)kNumSMs * 16;
#define FNET_PW_FWD(K_)                                                                                                              \
    case K_:                                                                                                                         \
        if (narrow) launch_pdl(pointwise_fwd_kernel<K_, 16>, dim3((unsigned)blocks), dim3(256), smem, s, M, g.C, act, x, w, bias, y); \
        else launch_pdl(pointwise_fwd_kernel<K_, 32>, dim3((unsigned)blocks), dim3(256), smem, s, M, g.C, act, x, w, bias, y);        \
        break;
        switch (g.K) { FNET_PW_FWD(1) FNET_PW_FWD(2) FNET_PW_FWD(3) FNET_PW_FWD(4) FNET_PW_FWD(5) FNET_PW_FWD(6) FNET_PW_FWD(7) default: FNET_PW_FWD(8) }
#undef FNET_PW_FWD
        return check_launch("conv2d_fwd[pointwise]");
    }
    if (is_dense_geom(g) && g.C >= kStreamMinK) {
        int blocks = stream_fwd_blocks(g);
        size_t need = (size_t)blocks * g.B * g.K * sizeof(float);
        if (!ws || ws_bytes < need) {
            set_error("dense_fwd[stream]: workspace of %zu bytes needed, %zu given", need, ws_bytes);
            return FNET_ERR_WORKSPACE;
        }
        float *partial = reinterpret_cast<float *>(ws);
        dim3 grid(blocks, (unsigned)ceil_div<long long>(g.B, kStreamRows));
        if (!(g.C & 3) && !(reinterpret_cast<uintptr_t>(x) & 15)) {
#define FNET_CALL(NM) launch_pdl(dense_stream_fwd4_kernel<NM>, dim3(grid), dim3(512), 0, s, g.B, g.C, g.K, x, w, partial)
            FNET_BY_COLUMNS(g.K, FNET_CALL);
#undef FNET_CALL
        } else {
#define FNET_CALL(NM) launch_pdl(dense_stream_fwd_kernel<NM>, dim3(grid), dim3(512), 0, s, g.B, g.C, g.K, x, w, partial)
            FNET_BY_COLUMNS(g.K, FNET_CALL);
#undef FNET_CALL
        }
        int rc = check_launch("dense_fwd[stream]");
        return rc ? rc : simt_reduce_fwd_partials(g, act, bias, y, blocks, partial, s);
    }
    if (is_dense_geom(g)) {
        launch_pdl(dense_skinny_fwd_kernel<kSkinnyN>, dim3(g.B), dim3(1024), 0, s, g.C, g.K, act, x, w, bias, y);
        return check_launch("dense_fwd[skinny]");
    }
    const int np = g.K <= 4 ? 4 : (g.K <= 8 ? 8 : (g.K <= 12 ? 12 : 16));
    size_t smem = ((size_t)g.kh * g.kw * g.C + 1) * np * sizeof(float);
    long long M = (long long)g.B * g.Ho * g.Wo;
    if (np == 4) launch_pdl(conv_small_fwd_kernel<4>, dim3(pixel_grid(M)), dim3(256), smem, s, g, act, x, w, bias, y);
    else if (np == 8) launch_pdl(conv_small_fwd_kernel<8>, dim3(pixel_grid(M)), dim3(256), smem, s, g, act, x, w, bias, y);
    else if (np == 12) launch_pdl(conv_small_fwd_kernel<12>, dim3(pixel_grid(M)), dim3(256), smem, s, g, act, x, w, bias, y);
    else launch_pdl(conv_small_fwd_kernel<16>, dim3(pixel_grid(M)), dim3(256), smem, s, g, act, x, w, bias, y);
    return check_launch("conv2d_fwd[small]");
}

int skinny_dgrad(const ConvGeom &g, const float *dz, const float *w, float *dx, DxEpilogue ep, cudaStream_t s) {
    if (pointwise_dgrad_ok(g)) {
        const int vec_store = !((reinterpret_cast<uintptr_t>(dx) | (ep.act ? reinterpret_cast<uintptr_t>(ep.y) : 0)) & 15);
        const long long M = (long long)g.B * g.H * g.W, total = M * (g.C >> 2);
        long long blocks = ceil_div<long long>(total, 256);
        const long long cap = (long long)kNumSMs * 16;
        launch_pdl(pointwise_dgrad_kernel, dim3((unsigned)(blocks > cap ? cap : blocks)), dim3(256), (size_t)g.C * g.K * sizeof(float), s,
                   (int)M, g.C, g.K, dz, w, dx, ep.accumulate, ep.act, ep.y, vec_store);
        return check_launch("conv2d_bwd_data[pointwise]");
    }
    if (is_dense_geom(g) && g.C >= kStreamMinK && !(g.C & 3) &&
        !((reinterpret_cast<uintptr_t>(dx) | (ep.act ? reinterpret_cast<uintptr_t>(ep.y) : 0)) & 15)) {
        long long bx = ceil_div<long long>(g.C / 4, 256);
        if (bx > 4096) bx = 4096;
        dim3 grid((unsigned)bx, (unsigned)ceil_div<long long>(g.B, 16));
#define FNET_CALL(NM)                                                                                                                  \
    do {                                                                                                                                \
        if (ep.accumulate || ep.act) launch_pdl(dense_rows_dgrad4_kernel<NM, true>, dim3(grid), dim3(256), 0, s, g.B, g.C, g.K, dz, w, dx, ep.accumulate, ep.act, ep.y); \
        else launch_pdl(dense_rows_dgrad4_kernel<NM, false>, dim3(grid), dim3(256), 0, s, g.B, g.C, g.K, dz, w, dx, 0, 0, ep.y);           \
    } while (0)
        FNET_BY_COLUMNS(g.K, FNET_CALL);
#undef FNET_CALL
        return check_launch("dense_bwd_data[skinny, 4 k per thread]");
    }
    if (is_dense_geom(g)) {
        long long bx = ceil_div<long long>(g.C, 256);
        if (bx > 4096) bx = 4096;
        dim3 grid((unsigned)bx, (unsigned)ceil_div<long long>(g.B, 16));
#define FNET_CALL(NM)                                                                                                                  \
    do {                                                                                                                                \
        if (ep.accumulate || ep.act) launch_pdl(dense_rows_dgrad_kernel<NM, true>, dim3(grid), dim3(256), 0, s, g.B, g.C, g.K, dz, w, dx, ep.accumulate, ep.act, ep.y); \
        else launch_pdl(dense_rows_dgrad_kernel<NM, false>, dim3(grid), dim3(256), 0, s, g.B, g.C, g.K, dz, w, dx, 0, 0, ep.y);           \
    } while (0)
        FNET_BY_COLUMNS(g.K, FNET_CALL);
#undef FNET_CALL
        return check_launch("dense_bwd_data[skinny]");
    }
    const int cp = g.C <= 4 ? 4 : (g.C <= 8 ? 8 : (g.C <= 12 ? 12 : 16));
    size_t smem = (size_t)g.kh * g.kw * g.K * cp * sizeof(float);
    long long M = (long long)g.B * g.H * g.W;
    if (cp == 4) launch_pdl(conv_small_dgrad_kernel<4>, dim3(pixel_grid(M)), dim3(256), smem, s, g, dz, w, dx, ep.accumulate, ep.act, ep.y);
    else if (cp == 8) launch_pdl(conv_small_dgrad_kernel<8>, dim3(pixel_grid(M)), dim3(256), smem, s, g, dz, w, dx, ep.accumulate, ep.act, ep.y);
    else if (cp == 12) launch_pdl(conv_small_dgrad_kernel<12>, dim3(pixel_grid(M)), dim3(256), smem, s, g, dz, w, dx, ep.accumulate, ep.act, ep.y);
    else launch_pdl(conv_small_dgrad_kernel<16>, dim3(pixel_grid(M)), dim3(256), smem, s, g, dz, w, dx, ep.accumulate, ep.act, ep.y);
    return check_launch("conv2d_bwd_data[small]");
}

int skinny_wgrad(const ConvGeom &g, const float *x, const float *dz, float *dw, float *dbias, void *ws, size_t ws_bytes,
                 cudaStream_t s) {
    if (pointwise_shape_ok(g)) {
        const int M = (int)((long long)g.B * g.H * g.W), quads = g.C >> 2, groups = 256 / quads;
        const int total = (g.C + 1) * g.K;                   // the kernel always produces the bias row
        long long blocks = ceil_div<long long>(M, groups);
        if (blocks > kPointwiseBlocks) blocks = kPointwiseBlocks;
        const size_t need = (size_t)blocks * total * sizeof(float);
        if (!ws || ws_bytes < need) {
            set_error("conv2d_bwd_filter[pointwise]: workspace of %zu bytes needed, %zu given", need, ws_bytes);
            return FNET_ERR_WORKSPACE;
        }
        float *partial = reinterpret_cast<float *>(ws);
        const size_t smem = ((size_t)groups * quads * 4 * g.K + (size_t)groups * g.K) * sizeof(float);
#define FNET_PW_WGRAD(K_) case K_: launch_pdl(pointwise_wgrad_kernel<K_>, dim3((unsigned)blocks), dim3(256), smem, s, M, g.C, x, dz, partial); break;
        switch (g.K) { FNET_PW_WGRAD(1) FNET_PW_WGRAD(2) FNET_PW_WGRAD(3) FNET_PW_WGRAD(4) FNET_PW_WGRAD(5) FNET_PW_WGRAD(6) FNET_PW_WGRAD(7) default: FNET_PW_WGRAD(8) }
#undef FNET_PW_WGRAD
        int rc = check_launch("conv2d_bwd_filter[pointwise]");
        if (rc) return rc;
        launch_pdl(conv_small_wgrad_reduce_kernel, dim3((unsigned)ceil_div<int>(total, 8)), dim3(256), 0, s, (int)blocks, g.C, g.K, total, partial, dw, dbias);
        return check_launch("conv2d_bwd_filter[pointwise]");
    }
    if (is_dense_geom(g) && g.C >= kStreamMinK) {
        unsigned blocks = (unsigned)ceil_div<long long>(g.C, 256);
#define FNET_CALL(NM) launch_pdl(dense_stream_wgrad_kernel<NM>, dim3(blocks), dim3(256), 0, s, g.B, g.C, g.K, x, dz, dw, dbias)
        FNET_BY_COLUMNS(g.K, FNET_CALL);
#undef FNET_CALL
        return check_launch("dense_bwd_filter[stream]");
    }
    if (is_dense_geom(g)) {
        size_t smem = (size_t)256 * g.K * sizeof(float);
        launch_pdl(dense_skinny_wgrad_kernel<kSkinnyN>, dim3((unsigned)ceil_div<long long>(g.C, 32)), dim3(256), smem, s, g.B, g.C, g.K, x, dz,
                                                                                                     dw, dbias);
        return check_launch("dense_bwd_filter[skinny]");
    }
    int taps_c = g.kh * g.kw * g.C, rows = taps_c + (dbias ? 1 : 0), total = rows * g.K;
    long long M = (long long)g.B * g.Ho * g.Wo;
    int blocks = (int)ceil_div<long long>(M, 256);
    if (blocks > kWgradBlocks) blocks = kWgradBlocks;
    if (blocks < 1) blocks = 1;
    size_t need = (size_t)blocks * total * sizeof(float);
    if (!ws || ws_bytes < need) {
        set_error("conv2d_bwd_filter[small]: workspace of %zu bytes needed, %zu given", need, ws_bytes);
        return FNET_ERR_WORKSPACE;
    }
    float *partial = reinterpret_cast<float *>(ws);
#define FNET_WGRAD_ROWS(KT_)                                                                                              \
    do {                                                                                                                  \
        if (rows <= 4) launch_pdl(conv_small_wgrad_kernel<KT_, 4>, dim3(blocks), dim3(256), 0, s, g, rows, x, dz, partial);       \
        else if (rows <= 8) launch_pdl(conv_small_wgrad_kernel<KT_, 8>, dim3(blocks), dim3(256), 0, s, g, rows, x, dz, partial);  \
        else if (rows <= 16) launch_pdl(conv_small_wgrad_kernel<KT_, 16>, dim3(blocks), dim3(256), 0, s, g, rows, x, dz, partial);\
        else launch_pdl(conv_small_wgrad_kernel<KT_, 64>, dim3(blocks), dim3(256), 0, s, g, rows, x, dz, partial);                \
    } while (0)
    switch (g.K) {
        case 1: FNET_WGRAD_ROWS(1); break;
        case 2: FNET_WGRAD_ROWS(2); break;
        case 3: FNET_WGRAD_ROWS(3); break;
        case 4: FNET_WGRAD_ROWS(4); break;
        case 5: FNET_WGRAD_ROWS(5); break;
        case 6: FNET_WGRAD_ROWS(6); break;
        case 7: FNET_WGRAD_ROWS(7); break;
        default: FNET_WGRAD_ROWS(8); break;
    }
#undef FNET_WGRAD_ROWS
    int rc = check_launch("conv2d_bwd_filter[small]");
    if (rc) return rc;
    launch_pdl(conv_small_wgrad_reduce_kernel, dim3(1), dim3(256), 0, s, blocks, taps_c, g.K, total, partial, dw, dbias);
    return check_launch("conv2d_bwd_filter[small]");
}

}  // namespace fnet

// Geometry of the three GEMM-shaped products of a convolution, shared by the FFMA kernels
// (igemm_simt.cu) and the tcgen05 kernels (igemm_tc.cu).  A "problem" maps GEMM coordinates to
// tensor elements:
//   forward         C[m, n]  m = (b, ho, wo)   n = out channel   k = (r, s, c)
//   backward data   C[m, n]  m = (b, h, w)     n = in channel    k = (r, s, out channel)
//   backward filter C[m, n]  m = (r, s, c) (+1 row of ones = bias)  n = out channel  k = (b, ho, wo)
// Each problem also says along which GEMM axis its operands are contiguous in memory
// (kAContigK / kBContigK), so that a loader can put the lanes of a warp on that axis, and offers a
// 16-byte gather (a4) for the cases where four consecutive k are four consecutive channels.
//
// Integer address arithmetic is what these kernels execute most, so the reduction index is never
// divided in the main loop: a column context is decoded once (make_col) and then stepped from one
// k-tile to the next (next_col); the same for the B operand of the data gradient (BCursor).
#pragma once
#include "igemm.cuh"

namespace fnet {

struct PixelCtx {   // one output pixel (forward / filter) or one input pixel (data)
    long long base; // element offset of the image it indexes into
    int h0, w0;     // top-left input coordinate of the window, or padded input coordinate
    int valid;
};

struct TapCtx {     // one (r, s, channel) of the reduction / filter row
    int r, s;       // tap coordinates (undilated)
    int ch;         // channel inside the tap
    int valid;      // 1 = real tap, 2 = the all-ones bias row (filter gradient only), 0 = out of range
};

__device__ __forceinline__ void decode_out_pixel(const ConvGeom &g, int m, int M, PixelCtx &p) {
    p.valid = m < M;
    int mm = p.valid ? m : 0;
    int b = mm / (g.Ho * g.Wo);
    int rem = mm - b * (g.Ho * g.Wo);
    int ho = rem / g.Wo, wo = rem - ho * g.Wo;
    p.base = (long long)b * g.H * g.W * g.C;
    p.h0 = ho * g.sh - g.pt;
    p.w0 = wo * g.sw - g.pl;
}

// the context of output pixel m + step given that of m: carries instead of divisions unless rows are very short
__device__ __forceinline__ void step_out_pixel(const ConvGeom &g, int m_new, int step, int M, PixelCtx &p) {
    if (g.Wo * 4 < step) {
        decode_out_pixel(g, m_new, M, p);
        return;
    }
    p.valid = m_new < M;
    const int row_w = g.Wo * g.sw, img_h = g.Ho * g.sh;
    p.w0 += step * g.sw;
    while (p.w0 >= row_w - g.pl) {
        p.w0 -= row_w;
        p.h0 += g.sh;
        if (p.h0 >= img_h - g.pt) {
            p.h0 -= img_h;
            p.base += (long long)g.H * g.W * g.C;
        }
    }
}

__device__ __forceinline__ void decode_in_pixel(const ConvGeom &g, int m, int M, PixelCtx &p) {
    p.valid = m < M;
    int mm = p.valid ? m : 0;
    int b = mm / (g.H * g.W);
    int rem = mm - b * (g.H * g.W);
    int h = rem / g.W, w = rem - h * g.W;
    p.base = (long long)b * g.Ho * g.Wo * g.K;
    p.h0 = h + g.pt;
    p.w0 = w + g.pl;
}

__device__ __forceinline__ void decode_tap(const ConvGeom &g, int k, int channels, int limit, TapCtx &t) {
    t.valid = k < limit;
    int kk = t.valid ? k : 0;
    int tap = kk / channels;
    t.ch = kk - tap * channels;
    t.r = tap / g.kw;
    t.s = tap - t.r * g.kw;
}

// the context of k + step given the context of k: no division when a tap holds at least a few channels
__device__ __forceinline__ void step_tap(const ConvGeom &g, int k_new, int step, int channels, int limit, TapCtx &t) {
    if (channels * 4 < step) {          // tiny channel counts would loop many times: decode from scratch
        decode_tap(g, k_new, channels, limit, t);
        return;
    }
    t.valid = k_new < limit;
    t.ch += step;
    while (t.ch >= channels) {
        t.ch -= channels;
        if (++t.s == g.kw) { t.s = 0; ++t.r; }
    }
}

// address of x[b, h0 + r*dh, w0 + s*dw, ch], or nullptr in the zero padding
__device__ __forceinline__ const float *input_at(const ConvGeom &g, const float *__restrict__ x, const PixelCtx &p,
                                                 const TapCtx &t) {
    int h = p.h0 + t.r * g.dh, w = p.w0 + t.s * g.dw;
    if ((unsigned)h >= (unsigned)g.H || (unsigned)w >= (unsigned)g.W) return nullptr;
    return x + p.base + ((long long)h * g.W + w) * g.C + t.ch;
}

// o = v / stride when the division is exact and 0 <= o < extent, else -1 (strides are 1 or 2 in practice)
__device__ __forceinline__ int exact_div(int v, int stride, int extent) {
    if (v < 0) return -1;
    int o;
    if (stride == 1) o = v;
    else if (stride == 2) { if (v & 1) return -1; o = v >> 1; }
    else { o = v / stride; if (o * stride != v) return -1; }
    return o < extent ? o : -1;
}

// address of dz[b, (hp - r*dh)/sh, (wp - s*dw)/sw, ch] when the divisions are exact and in range
__device__ __forceinline__ const float *grad_at(const ConvGeom &g, const float *__restrict__ dz, const PixelCtx &p,
                                                const TapCtx &t) {
    int ho = exact_div(p.h0 - t.r * g.dh, g.sh, g.Ho);
    int wo = exact_div(p.w0 - t.s * g.dw, g.sw, g.Wo);
    if ((ho | wo) < 0) return nullptr;
    return dz + p.base + ((long long)ho * g.Wo + wo) * g.K + t.ch;
}

__device__ __forceinline__ float load1(const float *q) { return q ? __ldg(q) : 0.f; }
__device__ __forceinline__ float4 load4(const float *q) {
    return q ? __ldg(reinterpret_cast<const float4 *>(q)) : make_float4(0.f, 0.f, 0.f, 0.f);
}

// ------------------------------------------------------------------ problems
struct FwdProblem {
    ConvGeom g;
    const float *x, *w, *bias;
    float *y;
    int act;
    typedef PixelCtx Row;
    typedef TapCtx Col;
    typedef const float *BCursor;               // B(k, n) = w[k][n]: a pointer that moves down the rows
    static constexpr bool kAContigK = true;     // A(m, k): consecutive k = consecutive input channels
    static constexpr bool kBContigK = false;    // consecutive n
    static constexpr bool kBiasRow = false;
    __host__ __device__ bool has_bias_row() const { return false; }
    __device__ int colsum_row(int) const { return -1; }
    __device__ bool vec4_ok() const { return (g.C & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0; }
    __device__ bool a_rows_vec4_ok() const { return false; }
    __device__ bool b_rows_vec4_ok() const { return (g.K & 3) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0; }
    __device__ bool b_k_vec4_ok() const { return false; }
    __device__ void make_row(int m, int M, Row &r) const { decode_out_pixel(g, m, M, r); }
    __device__ void make_col(int k, int kend, Col &c) const { decode_tap(g, k, g.C, kend, c); }
    __device__ void next_col(int k_new, int step, int kend, Col &c) const { step_tap(g, k_new, step, g.C, kend, c); }
    __device__ const float *a_ptr(const Row &r, const Col &c) const { return input_at(g, x, r, c); }
    __device__ float a(const Row &r, const Col &c) const { return load1(a_ptr(r, c)); }
    __device__ float4 a4(const Row &r, const Col &c) const { return load4(a_ptr(r, c)); }
    __device__ BCursor b_begin(int k, int n) const { return w + (long long)k * g.K + n; }
    __device__ void b_next(BCursor &c, int step) const { c += (long long)step * g.K; }
    __device__ const float *b_ptr(const BCursor &c) const { return c; }
    __device__ long long b_stride_n() const { return 1; }
    __device__ long long b_stride_k() const { return g.K; }
    __device__ const float *any_ptr() const { return x; }
    __device__ void store(int m, int n, float v) const {
        if (bias) v += __ldg(bias + n);
        y[(long long)m * g.K + n] = act_apply(v, act);
    }
    __device__ bool has_pre() const { return false; }
    __device__ float pre(int, int) const { return 1.f; }
    __device__ void store_pre(int m, int n, float v, float) const { store(m, n, v); }
    // A tcgen05.ld.16x256b.x4 fragment: thread t holds, for each of four 8-column blocks i, columns n + 8 i + {0, 1}
    // (n already includes 2 * (t % 4)) of rows m and m + 8 (v[4 i + 0 .. 1] and v[4 i + 2 .. 3]).  Eight-byte stores:
    // the four lanes of a row write one 32-byte sector, a warp instruction eight full sectors -- no trip through
    // shared memory.  cols = columns that exist from n on; K % 2 == 0 and y 8-byte aligned.
    __device__ void store_frag(int m, bool row_a, bool row_b, int n, int cols, const uint32_t (&v)[16]) const {
        float2 bv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            bv[i] = (bias && 8 * i < cols) ? make_float2(__ldg(bias + n + 8 * i), __ldg(bias + n + 8 * i + 1)) : make_float2(0.f, 0.f);
        float *qa = y + (long long)m * g.K + n, *qb = qa + 8LL * g.K;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (8 * i < cols) {
                float2 a = make_float2(__uint_as_float(v[4 * i]) + bv[i].x, __uint_as_float(v[4 * i + 1]) + bv[i].y);
                float2 b = make_float2(__uint_as_float(v[4 * i + 2]) + bv[i].x, __uint_as_float(v[4 * i + 3]) + bv[i].y);
                if (act == FNET_ACT_RELU) {
                    a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); b.x = fmaxf(b.x, 0.f); b.y = fmaxf(b.y, 0.f);
                } else if (act != FNET_ACT_NONE) {
                    a.x = act_apply(a.x, act); a.y = act_apply(a.y, act); b.x = act_apply(b.x, act); b.y = act_apply(b.y, act);
                }
                if (row_a) *reinterpret_cast<float2 *>(qa + 8 * i) = a;
                if (row_b) *reinterpret_cast<float2 *>(qb + 8 * i) = b;
            }
        }
    }
    __host__ __device__ const void *out_ptr() const { return y; }
    __host__ __device__ const void *out_aux_ptr() const { return nullptr; }
    // rows m .. m+count-1 of column n; src[r * src_stride] holds row r.  Bias and the activation switch are
    // taken out of the row loop (the tensor-core epilogue spends most of its instructions here).
    __device__ void store_col(int m, int count, int n, const float *src, int src_stride) const {
        const float bv = bias ? __ldg(bias + n) : 0.f;
        float *q = y + (long long)m * g.K + n;
        const long long ld = g.K;
        if (act == FNET_ACT_NONE) {
#pragma unroll 8
            for (int r = 0; r < count; ++r) q[r * ld] = src[r * src_stride] + bv;
        } else if (act == FNET_ACT_RELU) {
#pragma unroll 8
            for (int r = 0; r < count; ++r) q[r * ld] = fmaxf(src[r * src_stride] + bv, 0.f);
        } else {
#pragma unroll 4
            for (int r = 0; r < count; ++r) q[r * ld] = act_apply(src[r * src_stride] + bv, act);
        }
    }
};

struct DgradCursor {
    int tap, ko, n;
};

struct DgradProblem {
    ConvGeom g;
    const float *dz, *w;
    float *dx;
    int accumulate;
    int act;            // activation of the layer that produced x ...
    const float *yact;  // ... and its output: dx is multiplied by act'(yact) (DxEpilogue)
    typedef PixelCtx Row;
    typedef TapCtx Col;
    typedef DgradCursor BCursor;                // B(k, n) = w[tap][n][ko] with k = tap*K + ko
    static constexpr bool kAContigK = true;     // consecutive k = consecutive output channels of dz
    static constexpr bool kBContigK = true;     // consecutive ko
    static constexpr bool kBiasRow = false;
    __host__ __device__ bool has_bias_row() const { return false; }
    __device__ int colsum_row(int) const { return -1; }
    __device__ bool vec4_ok() const { return (g.K & 3) == 0 && (reinterpret_cast<uintptr_t>(dz) & 15) == 0; }
    __device__ bool a_rows_vec4_ok() const { return false; }
    __device__ bool b_rows_vec4_ok() const { return false; }
    // w[tap][c][ko]: four consecutive ko of one (tap, c) row are 16 contiguous, aligned bytes when K % 4 == 0
    __device__ bool b_k_vec4_ok() const { return (g.K & 3) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0; }
    __device__ void make_row(int m, int M, Row &r) const { decode_in_pixel(g, m, M, r); }
    __device__ void make_col(int k, int kend, Col &c) const { decode_tap(g, k, g.K, kend, c); }
    __device__ void next_col(int k_new, int step, int kend, Col &c) const { step_tap(g, k_new, step, g.K, kend, c); }
    __device__ const float *a_ptr(const Row &r, const Col &c) const { return grad_at(g, dz, r, c); }
    __device__ float a(const Row &r, const Col &c) const { return load1(a_ptr(r, c)); }
    __device__ float4 a4(const Row &r, const Col &c) const { return load4(a_ptr(r, c)); }
    __device__ BCursor b_begin(int k, int n) const {
        BCursor c;
        c.tap = k / g.K;
        c.ko = k - c.tap * g.K;
        c.n = n;
        return c;
    }
    __device__ void b_next(BCursor &c, int step) const {
        c.ko += step;
        if (g.K * 4 < step) { c.tap += c.ko / g.K; c.ko %= g.K; return; }
        while (c.ko >= g.K) { c.ko -= g.K; ++c.tap; }
    }
    __device__ const float *b_ptr(const BCursor &c) const { return w + ((long long)c.tap * g.C + c.n) * g.K + c.ko; }
    __device__ long long b_stride_n() const { return g.K; }
    __device__ long long b_stride_k() const { return 1; }
    __device__ const float *any_ptr() const { return dz; }
    __device__ void store(int m, int n, float v) const {
        store_dx(dx, (long long)m * g.C + n, v, accumulate, act, yact);
    }
    // two-phase epilogue for the FFMA kernel: fetch every multiplier of the thread's tile, then store
    __device__ bool has_pre() const { return act != 0; }
    __device__ float pre(int m, int n) const { return __ldg(yact + (long long)m * g.C + n); }     // the raw y
    __device__ void store_pre(int m, int n, float v, float yv) const {
        long long at = (long long)m * g.C + n;
        if (accumulate) v += dx[at];
        dx[at] = v * act_grad(yv, act);
    }
    // see FwdProblem::store_frag; every load (act'(y) multipliers, old values) before the first store
    __device__ void store_frag(int m, bool row_a, bool row_b, int n, int cols, const uint32_t (&v)[16]) const {
        float *qa = dx + (long long)m * g.C + n, *qb = qa + 8LL * g.C;
        float2 ya[4], yb[4], oa[4], ob[4];
        if (act) {
            const float *pa = yact + (long long)m * g.C + n, *pb = pa + 8LL * g.C;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                ya[i] = (row_a && 8 * i < cols) ? __ldg(reinterpret_cast<const float2 *>(pa + 8 * i)) : make_float2(0.f, 0.f);
                yb[i] = (row_b && 8 * i < cols) ? __ldg(reinterpret_cast<const float2 *>(pb + 8 * i)) : make_float2(0.f, 0.f);
            }
        }
        if (accumulate) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                oa[i] = (row_a && 8 * i < cols) ? *reinterpret_cast<const float2 *>(qa + 8 * i) : make_float2(0.f, 0.f);
                ob[i] = (row_b && 8 * i < cols) ? *reinterpret_cast<const float2 *>(qb + 8 * i) : make_float2(0.f, 0.f);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (8 * i < cols) {
                float2 a = make_float2(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]));
                float2 b = make_float2(__uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
                if (accumulate) { a.x += oa[i].x; a.y += oa[i].y; b.x += ob[i].x; b.y += ob[i].y; }
                if (act) {
                    a.x *= act_grad(ya[i].x, act); a.y *= act_grad(ya[i].y, act);
                    b.x *= act_grad(yb[i].x, act); b.y *= act_grad(yb[i].y, act);
                }
                if (row_a) *reinterpret_cast<float2 *>(qa + 8 * i) = a;
                if (row_b) *reinterpret_cast<float2 *>(qb + 8 * i) = b;
            }
        }
    }
    __host__ __device__ const void *out_ptr() const { return dx; }
    __host__ __device__ const void *out_aux_ptr() const { return act ? yact : nullptr; }
    __device__ void store_col(int m, int count, int n, const float *src, int src_stride) const {
        float *q = dx + (long long)m * g.C + n;
        const long long ld = g.C;
        if (act) {
            // the multipliers act'(y) first, all loads in flight at once: a load inside the store loop would put
            // one memory latency per row on the epilogue warp (1.25 ms instead of 55 us on a 262 144 x 192 tile set)
            const float *yq = yact + (long long)m * g.C + n;
            float yv[32];                                   // raw loads only: no control flow between them
#pragma unroll
            for (int r = 0; r < 32; ++r) yv[r] = r < count ? __ldg(yq + r * ld) : 0.f;
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                if (r < count) {
                    float v = src[r * src_stride];
                    if (accumulate) v += q[r * ld];
                    q[r * ld] = v * act_grad(yv[r], act);
                }
            }
        } else if (accumulate) {
#pragma unroll 8
            for (int r = 0; r < count; ++r) q[r * ld] += src[r * src_stride];
        } else {
#pragma unroll 8
            for (int r = 0; r < count; ++r) q[r * ld] = src[r * src_stride];
        }
    }
};

struct WgradProblem {
    ConvGeom g;
    const float *x, *dz;
    float *dw, *dbias;
    int taps_c;   // kh*kw*C: rows below are filter taps, row == taps_c is the bias row
    typedef TapCtx Row;
    typedef PixelCtx Col;
    typedef const float *BCursor;               // B(k, n) = dz[pixel k][n]
    static constexpr bool kAContigK = false;    // A(m, k) = x[pixel k][tap/channel m]: consecutive m (channels)
    static constexpr bool kBContigK = false;    // consecutive n
    static constexpr bool kBiasRow = true;      // row taps_c is the all-ones row whose products are the bias gradient
    __host__ __device__ bool has_bias_row() const { return dbias != nullptr; }
    // when the tap rows fill bm-row tiles exactly, the bias row is produced from column sums of B instead
    __device__ int colsum_row(int bm) const { return (dbias != nullptr && taps_c % bm == 0) ? taps_c : -1; }
    __device__ bool vec4_ok() const { return false; }
    // four consecutive rows = four consecutive channels of one tap; four consecutive columns of dz
    __device__ bool a_rows_vec4_ok() const { return (g.C & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0; }
    __device__ bool b_rows_vec4_ok() const { return (g.K & 3) == 0 && (reinterpret_cast<uintptr_t>(dz) & 15) == 0; }
    __device__ bool b_k_vec4_ok() const { return false; }
    __device__ void make_row(int k, int M, Row &r) const {
        decode_tap(g, k, g.C, taps_c, r);
        if (k == taps_c && k < M) r.valid = 2;
    }
    __device__ void make_col(int m, int mend, Col &c) const { decode_out_pixel(g, m, mend, c); }
    __device__ void next_col(int m_new, int step, int mend, Col &c) const { step_out_pixel(g, m_new, step, mend, c); }
    __device__ const float *a_ptr(const Row &r, const Col &c) const { return input_at(g, x, c, r); }
    __device__ float a(const Row &r, const Col &c) const { return r.valid == 2 ? 1.f : load1(a_ptr(r, c)); }
    __device__ float4 a4(const Row &, const Col &) const { return make_float4(0.f, 0.f, 0.f, 0.f); }
    __device__ BCursor b_begin(int m, int n) const { return dz + (long long)m * g.K + n; }
    __device__ void b_next(BCursor &c, int step) const { c += (long long)step * g.K; }
    __device__ const float *b_ptr(const BCursor &c) const { return c; }
    __device__ long long b_stride_n() const { return 1; }
    __device__ long long b_stride_k() const { return g.K; }
    __device__ const float *any_ptr() const { return x; }
    __device__ void store(int k, int n, float v) const {
        if (k < taps_c) dw[(long long)k * g.K + n] = v;
        else if (dbias) dbias[n] = v;
    }
    __device__ bool has_pre() const { return false; }
    __device__ float pre(int, int) const { return 1.f; }
    __device__ void store_pre(int m, int n, float v, float) const { store(m, n, v); }
    __device__ void store_col(int k, int count, int n, const float *src, int src_stride) const {
#pragma unroll 4
        for (int r = 0; r < count; ++r) store(k + r, n, src[r * src_stride]);
    }
};

// ------------------------------------------------------------------ split reductions
// partial[z][M][N], z < splits: summed in a fixed order and handed to the problem's epilogue (no atomics, so a
// result never depends on scheduling).  Few outputs and many splits: one warp per output element, lanes over
// the splits and a fixed shuffle tree.  Many outputs: one thread per element, splits in sequence (coalesced).
template <class P>
__global__ void __launch_bounds__(256) reduce_splits_warp_kernel(P p, int M, int N, int splits,
                                                                  const float *__restrict__ partial) {
    pdl_prologue();
    long long total = (long long)M * N;
    int lane = threadIdx.x & 31;
    long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long i = warp; i < total; i += nwarps) {
        float s = 0.f;
        for (int z = lane; z < splits; z += 32) s += partial[(long long)z * total + i];
        s = warp_sum(s);
        if (lane == 0) {
            int m = (int)(i / N), n = (int)(i - (long long)m * N);
            p.store(m, n, s);
        }
    }
}

template <class P>
__global__ void __launch_bounds__(256) reduce_splits_flat_kernel(P p, int M, int N, int splits,
                                                                  const float *__restrict__ partial) {
    pdl_prologue();
    long long total = (long long)M * N;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int z = 0; z < splits; ++z) s += partial[(long long)z * total + i];
        int m = (int)(i / N), n = (int)(i - (long long)m * N);
        p.store(m, n, s);
    }
}

template <class P>
static int launch_reduce_splits(const P &p, int M, int N, int splits, const float *partial, cudaStream_t stream,
                                const char *what) {
    long long total = (long long)M * N;
    if (splits <= 16 || total >= (1LL << 16)) {
        long long blocks = ceil_div<long long>(total, 256), cap = (long long)kNumSMs * 8;
        launch_pdl(reduce_splits_flat_kernel<P>, dim3((unsigned)(blocks > cap ? cap : blocks)), dim3(256), 0, stream, p, M, N, splits, partial);
    } else {
        long long blocks = ceil_div<long long>(total, 8), cap = (long long)kNumSMs * 8;
        launch_pdl(reduce_splits_warp_kernel<P>, dim3((unsigned)(blocks > cap ? cap : blocks)), dim3(256), 0, stream, p, M, N, splits, partial);
    }
    return check_launch(what);
}

}  // namespace fnet

// Memory-bound element-wise and data-movement kernels: activations, softmax rows, dropout
// (Philox, mask recomputed in the backward), merges, row-block copies, zero padding.
// All of them are single-pass, float4-vectorised where alignment allows, grid-stride with the
// grid capped at 8 CTAs per SM; their roofline is HBM (or L2 when the tensors fit in 126 MB).
#include "common.cuh"

namespace fnet {

// ------------------------------------------------------------------ generic vectorised map
// Applies f to n elements; f4 gets (index of first element, float4 inputs...) so that ops with a
// per-element index (dropout) can use it.
// Four independent 16-byte loads per operand are in flight per thread before the first use: with 8 CTAs of 256 threads
// per SM that is ~19 MB outstanding per operand, enough to cover HBM latency at 6.5 TB/s (one load per thread, 4.8 MB,
// was not: relu forward on 128x32x32x64 ran at 50 % of the copy rate).
template <class F>
__device__ __forceinline__ float4 apply4(const F &f, float4 x, int64_t first) {
    return make_float4(f(x.x, first), f(x.y, first + 1), f(x.z, first + 2), f(x.w, first + 3));
}

// quads in flight per thread: 4 for the cheap functors; a functor with a long body (dropout's Philox rounds) keeps 1,
// four copies of it cost 137 registers and left one CTA per SM (ncu: 12 % of the warp slots, 34 us for 67 MB)
template <class F> struct MapUnroll { static constexpr int value = 4; };

template <class F>
__global__ void __launch_bounds__(256) map1_kernel(int64_t n, const float *__restrict__ a, float *__restrict__ y,
                                                   int accumulate, F f) {
    pdl_prologue();
    const int64_t n4 = n >> 2, stride = (int64_t)gridDim.x * blockDim.x, i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const float4 *a4 = reinterpret_cast<const float4 *>(a);
    float4 *y4 = reinterpret_cast<float4 *>(y);
    constexpr int U = MapUnroll<F>::value;
    for (int64_t v = i; v < n4; v += U * stride) {
        float4 x[U], o[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t at = v + u * stride;
            x[u] = at < n4 ? __ldg(a4 + at) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (accumulate && at < n4) o[u] = y4[at];
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t at = v + u * stride;
            if (at < n4) {
                float4 r = apply4(f, x[u], at * 4);
                if (accumulate) { r.x += o[u].x; r.y += o[u].y; r.z += o[u].z; r.w += o[u].w; }
                y4[at] = r;
            }
        }
    }
    for (int64_t e = (n4 << 2) + i; e < n; e += stride) {
        float r = f(__ldg(a + e), e);
        y[e] = accumulate ? y[e] + r : r;
    }
}

template <class F>
__global__ void __launch_bounds__(256) map2_kernel(int64_t n, const float *__restrict__ a, const float *__restrict__ b,
                                                   float *__restrict__ y, int accumulate, F f) {
    pdl_prologue();
    const int64_t n4 = n >> 2, stride = (int64_t)gridDim.x * blockDim.x, i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const float4 *a4 = reinterpret_cast<const float4 *>(a);
    const float4 *b4 = reinterpret_cast<const float4 *>(b);
    float4 *y4 = reinterpret_cast<float4 *>(y);
    for (int64_t v = i; v < n4; v += 4 * stride) {
        float4 p[4], q[4], o[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t at = v + u * stride;
            const bool in = at < n4;
            p[u] = in ? __ldg(a4 + at) : make_float4(0.f, 0.f, 0.f, 0.f);
            q[u] = in ? __ldg(b4 + at) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (accumulate && in) o[u] = y4[at];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t at = v + u * stride;
            if (at < n4) {
                float4 r = make_float4(f(p[u].x, q[u].x), f(p[u].y, q[u].y), f(p[u].z, q[u].z), f(p[u].w, q[u].w));
                if (accumulate) { r.x += o[u].x; r.y += o[u].y; r.z += o[u].z; r.w += o[u].w; }
                y4[at] = r;
            }
        }
    }
    for (int64_t e = (n4 << 2) + i; e < n; e += stride) {
        float r = f(__ldg(a + e), __ldg(b + e));
        y[e] = accumulate ? y[e] + r : r;
    }
}

static inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <class F>
static int launch_map1(int64_t n, const float *a, float *y, int accumulate, F f, void *stream, const char *what) {
    if (n == 0) return FNET_OK;
    if (n < 0 || !a || !y) { set_error("%s: bad arguments", what); return FNET_ERR_ARG; }
    if (!aligned16(a) || !aligned16(y)) { set_error("%s: pointers must be 16-byte aligned", what); return FNET_ERR_ARG; }
    launch_pdl(map1_kernel<F>, dim3(stream_grid(ceil_div<int64_t>(n, 4), 256)), dim3(256), 0, as_stream(stream), n, a, y, accumulate, f);
    return check_launch(what);
}

template <class F>
static int launch_map2(int64_t n, const float *a, const float *b, float *y, int accumulate, F f, void *stream,
                       const char *what) {
    if (n == 0) return FNET_OK;
    if (n < 0 || !a || !b || !y) { set_error("%s: bad arguments", what); return FNET_ERR_ARG; }
    if (!aligned16(a) || !aligned16(b) || !aligned16(y)) {
        set_error("%s: pointers must be 16-byte aligned", what);
        return FNET_ERR_ARG;
    }
    launch_pdl(map2_kernel<F>, dim3(stream_grid(ceil_div<int64_t>(n, 4), 256)), dim3(256), 0, as_stream(stream), n, a, b, y, accumulate, f);
    return check_launch(what);
}

// ------------------------------------------------------------------ functors
struct ActFwd { int act; __device__ float operator()(float x, int64_t) const { return act_apply(x, act); } };
struct ActBwd { int act; __device__ float operator()(float y, float dy) const { return dy * act_grad(y, act); } };
struct AddOp { __device__ float operator()(float a, float b) const { return a + b; } };
struct MulOp { __device__ float operator()(float a, float b) const { return a * b; } };
struct CopyOp { __device__ float operator()(float a, int64_t) const { return a; } };

// Philox-4x32-10 (Salmon et al.), the same stream as oracle.ops.philox4x32_10
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}

struct DropoutOp {
    uint2 key;
    uint32_t stream_id;
    const int32_t *step;
    uint64_t threshold;   // keep <=> word >= threshold
    float keep;           // 1 - rate; kept values are DIVIDED by it, as tf.nn.dropout does
    __device__ float operator()(float x, int64_t i) const {
        uint4 c = make_uint4((uint32_t)(i >> 2), (uint32_t)__ldg(step), stream_id, 0u);
        uint4 r = philox4x32_10(c, key);
        uint32_t lane = (uint32_t)(i & 3);
        uint32_t word = lane == 0 ? r.x : (lane == 1 ? r.y : (lane == 2 ? r.z : r.w));
        return (uint64_t)word >= threshold ? x / keep : 0.f;
    }
};

// the four elements of an aligned quad share one Philox block: compute it once (the element-wise functor above drew
// the same block four times -- dropout ran at 31 % of the copy rate, bound by integer multiplies)
template <> struct MapUnroll<DropoutOp> { static constexpr int value = 1; };

template <>
__device__ __forceinline__ float4 apply4<DropoutOp>(const DropoutOp &f, float4 x, int64_t first) {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)(first >> 2), (uint32_t)__ldg(f.step), f.stream_id, 0u), f.key);
    return make_float4((uint64_t)r.x >= f.threshold ? x.x / f.keep : 0.f, (uint64_t)r.y >= f.threshold ? x.y / f.keep : 0.f,
                       (uint64_t)r.z >= f.threshold ? x.z / f.keep : 0.f, (uint64_t)r.w >= f.threshold ? x.w / f.keep : 0.f);
}

static DropoutOp make_dropout(float rate, uint64_t seed, uint32_t stream_id, const int32_t *step) {
    DropoutOp op;
    op.key = make_uint2((uint32_t)(seed & 0xffffffffu), (uint32_t)(seed >> 32));
    op.stream_id = stream_id;
    op.step = step;
    double t = ceil((double)rate * 4294967296.0);
    op.threshold = t >= 4294967296.0 ? 4294967296ull : (uint64_t)t;
    op.keep = 1.f - rate;
    return op;
}

// ------------------------------------------------------------------ softmax over rows
// one warp per row; cols is the channel / class count (10..2048 here)
__global__ void __launch_bounds__(256) softmax_rows_kernel(int64_t rows, int cols, const float *__restrict__ x,
                                                           float *__restrict__ y) {
    pdl_prologue();
    int lane = threadIdx.x & 31;
    int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int64_t nrow_stride = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (; row < rows; row += nrow_stride) {
        const float *xr = x + row * cols;
        float *yr = y + row * cols;
        float mx = -INFINITY;
        for (int c = lane; c < cols; c += 32) mx = fmaxf(mx, xr[c]);
        mx = warp_max(mx);
        float sum = 0.f;
        for (int c = lane; c < cols; c += 32) sum += expf(xr[c] - mx);
        sum = warp_sum(sum);
        for (int c = lane; c < cols; c += 32) yr[c] = expf(xr[c] - mx) / sum;
    }
}

// dx (+)= y * (dy - sum(dy * y)) per row
__global__ void __launch_bounds__(256) softmax_rows_bwd_kernel(int64_t rows, int cols, const float *__restrict__ y,
                                                               const float *__restrict__ dy, float *__restrict__ dx,
                                                               int accumulate) {
    pdl_prologue();
    int lane = threadIdx.x & 31;
    int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int64_t nrow_stride = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (; row < rows; row += nrow_stride) {
        const float *yr = y + row * cols, *gr = dy + row * cols;
        float *dr = dx + row * cols;
        float dot = 0.f;
        for (int c = lane; c < cols; c += 32) dot += yr[c] * gr[c];
        dot = warp_sum(dot);
        for (int c = lane; c < cols; c += 32) {
            float v = yr[c] * (gr[c] - dot);
            dr[c] = accumulate ? dr[c] + v : v;
        }
    }
}

// ------------------------------------------------------------------ row-block copy (concat axis 1)
__global__ void __launch_bounds__(256) copy_rows_kernel(int batch, int src_rows, int src_row0, int dst_rows,
                                                        int dst_row0, int64_t block_elems, int inner,
                                                        const float *__restrict__ src, float *__restrict__ dst,
                                                        int accumulate) {
    pdl_prologue();
    // per batch item the copied region is one contiguous block of rows*inner floats
    int64_t src_stride = (int64_t)src_rows * inner, dst_stride = (int64_t)dst_rows * inner;
    int64_t src_off = (int64_t)src_row0 * inner, dst_off = (int64_t)dst_row0 * inner;
    int64_t total = (int64_t)batch * block_elems;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t b, e;
        if (i < 0x7fffffffLL && block_elems < 0x7fffffffLL) {
            unsigned q = (unsigned)i / (unsigned)block_elems;
            b = q;
            e = (unsigned)i - q * (unsigned)block_elems;
        } else {
            b = i / block_elems;
            e = i - b * block_elems;
        }
        float v = __ldg(src + b * src_stride + src_off + e);
        float *d = dst + b * dst_stride + dst_off + e;
        *d = accumulate ? *d + v : v;
    }
}

__global__ void __launch_bounds__(256) pad2d_kernel(int batch, int h, int w, int c, int ph, int pw,
                                                    const float *__restrict__ x, float *__restrict__ y) {
    pdl_prologue();
    int H2 = h + 2 * ph, W2 = w + 2 * pw;
    int64_t total = (int64_t)batch * H2 * W2 * c;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int cc, ww, hh;
        int64_t b;
        decode4(i, H2, W2, c, b, hh, ww, cc);
        int sh = hh - ph, sw = ww - pw;
        float v = 0.f;
        if ((unsigned)sh < (unsigned)h && (unsigned)sw < (unsigned)w) v = __ldg(x + ((b * h + sh) * w + sw) * c + cc);
        y[i] = v;
    }
}

__global__ void __launch_bounds__(256) pad2d_bwd_kernel(int batch, int h, int w, int c, int ph, int pw,
                                                        const float *__restrict__ dy, float *__restrict__ dx,
                                                        int accumulate) {
    pdl_prologue();
    int H2 = h + 2 * ph, W2 = w + 2 * pw;
    (void)H2;
    int64_t total = (int64_t)batch * h * w * c;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int cc, ww, hh;
        int64_t b;
        decode4(i, h, w, c, b, hh, ww, cc);
        float v = __ldg(dy + ((b * (h + 2 * ph) + hh + ph) * W2 + ww + pw) * c + cc);
        dx[i] = accumulate ? dx[i] + v : v;
    }
}

// Concatenate(axis=1) of two [outer, rows_*, inner] tensors in one launch, and its backward (either gradient
// may be skipped with a null pointer).  T = float4 when both blocks are multiples of four floats and the
// pointers are 16-byte aligned (the usual case), I = 32-bit indices below 2^31 elements: one cheap division
// and 16 bytes per thread instead of a 64-bit division per float.
__device__ __forceinline__ float4 vadd(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float vadd(float a, float b) { return a + b; }

template <typename T, typename I>
__global__ void __launch_bounds__(256) concat_rows_fwd_kernel(I outer, I blk_a, I blk_b, const T *__restrict__ a,
                                                              const T *__restrict__ b, T *__restrict__ y) {
    pdl_prologue();
    const I blk = blk_a + blk_b, total = outer * blk;
    for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (I)gridDim.x * blockDim.x) {
        I o = i / blk, e = i - o * blk;
        y[i] = e < blk_a ? __ldg(a + o * blk_a + e) : __ldg(b + o * blk_b + (e - blk_a));
    }
}

template <typename T, typename I>
__global__ void __launch_bounds__(256) concat_rows_bwd_kernel(I outer, I blk_a, I blk_b, const T *__restrict__ dy,
                                                              T *__restrict__ da, T *__restrict__ db, int acc_a, int acc_b) {
    pdl_prologue();
    const I blk = blk_a + blk_b, total = outer * blk;
    for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (I)gridDim.x * blockDim.x) {
        I o = i / blk, e = i - o * blk;
        T g = __ldg(dy + i);
        if (e < blk_a) {
            if (da) { T *d = da + o * blk_a + e; *d = acc_a ? vadd(*d, g) : g; }
        } else if (db) {
            T *d = db + o * blk_b + (e - blk_a);
            *d = acc_b ? vadd(*d, g) : g;
        }
    }
}


// concat backward that also applies the activation derivative of the layer(s) that produced the operands:
// da = (acc_a ? da + slice : slice) * act_a'(ya), likewise b.  One pass instead of the copy plus an act_bwd per operand.
__device__ __forceinline__ float4 grad4(float4 g, float4 y, int act) {
    return make_float4(g.x * act_grad(y.x, act), g.y * act_grad(y.y, act), g.z * act_grad(y.z, act), g.w * act_grad(y.w, act));
}
__device__ __forceinline__ float grad4(float g, float y, int act) { return g * act_grad(y, act); }

template <typename T>
__global__ void __launch_bounds__(256) concat_rows_bwd_act_kernel(int outer, int blk_a, int blk_b, const T *__restrict__ dy, T *da, T *db,
                                                                  int acc_a, int acc_b, int act_a, const T *__restrict__ ya, int act_b,
                                                                  const T *__restrict__ yb) {
    pdl_prologue();
    const int blk = blk_a + blk_b, total = outer * blk;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int o = i / blk, e = i - o * blk;
        T g = __ldg(dy + i);
        if (e < blk_a) {
            if (da) {
                const int at = o * blk_a + e;
                if (acc_a) g = vadd(da[at], g);
                da[at] = act_a ? grad4(g, __ldg(ya + at), act_a) : g;
            }
        } else if (db) {
            const int at = o * blk_b + (e - blk_a);
            if (acc_b) g = vadd(db[at], g);
            db[at] = act_b ? grad4(g, __ldg(yb + at), act_b) : g;
        }
    }
}

int run_softmax_rows(int64_t rows, int64_t cols, const float *x, float *y, cudaStream_t stream) {
    int64_t blocks = ceil_div<int64_t>(rows, 8);
    int64_t cap = (int64_t)kNumSMs * 8;
    launch_pdl(softmax_rows_kernel, dim3((int)(blocks > cap ? cap : blocks)), dim3(256), 0, stream, rows, (int)cols, x, y);
    return check_launch("softmax");
}

}  // namespace fnet

using namespace fnet;

extern "C" {

int fnet_act_fwd(int act, int64_t rows, int64_t cols, const float *x, float *y, void *stream) {
    FNET_REQUIRE(rows >= 0 && cols > 0 && cols < (1 << 30), "fnet_act_fwd: bad shape");
    FNET_REQUIRE(act >= FNET_ACT_NONE && act <= FNET_ACT_SOFTMAX, "fnet_act_fwd: unknown activation %d", act);
    if (rows == 0) return FNET_OK;
    if (act == FNET_ACT_SOFTMAX) {
        FNET_REQUIRE(x && y, "fnet_act_fwd: null pointer");
        return run_softmax_rows(rows, cols, x, y, as_stream(stream));
    }
    return launch_map1(rows * cols, x, y, 0, ActFwd{act}, stream, "fnet_act_fwd");
}

int fnet_act_bwd(int act, int64_t rows, int64_t cols, const float *y, const float *dy, float *dx, int accumulate,
                 void *stream) {
    FNET_REQUIRE(rows >= 0 && cols > 0 && cols < (1 << 30), "fnet_act_bwd: bad shape");
    FNET_REQUIRE(act >= FNET_ACT_NONE && act <= FNET_ACT_SOFTMAX, "fnet_act_bwd: unknown activation %d", act);
    if (rows == 0) return FNET_OK;
    if (act == FNET_ACT_SOFTMAX) {
        FNET_REQUIRE(y && dy && dx, "fnet_act_bwd: null pointer");
        int64_t blocks = ceil_div<int64_t>(rows, 8);
        int64_t cap = (int64_t)kNumSMs * 8;
        launch_pdl(softmax_rows_bwd_kernel, dim3((int)(blocks > cap ? cap : blocks)), dim3(256), 0, as_stream(stream), 
            rows, (int)cols, y, dy, dx, accumulate);
        return check_launch("fnet_act_bwd");
    }
    return launch_map2(rows * cols, y, dy, dx, accumulate, ActBwd{act}, stream, "fnet_act_bwd");
}

int fnet_dropout_fwd(int64_t n, float rate, uint64_t seed, uint32_t stream_id, const int32_t *step, const float *x,
                     float *y, void *stream) {
    FNET_REQUIRE(rate >= 0.f && rate < 1.f && step != nullptr, "fnet_dropout_fwd: bad arguments");
    return launch_map1(n, x, y, 0, make_dropout(rate, seed, stream_id, step), stream, "fnet_dropout_fwd");
}

int fnet_dropout_bwd(int64_t n, float rate, uint64_t seed, uint32_t stream_id, const int32_t *step, const float *dy,
                     float *dx, int accumulate, void *stream) {
    FNET_REQUIRE(rate >= 0.f && rate < 1.f && step != nullptr, "fnet_dropout_bwd: bad arguments");
    return launch_map1(n, dy, dx, accumulate, make_dropout(rate, seed, stream_id, step), stream, "fnet_dropout_bwd");
}

int fnet_add_fwd(int64_t n, const float *a, const float *b, float *y, void *stream) {
    return launch_map2(n, a, b, y, 0, AddOp{}, stream, "fnet_add_fwd");
}

int fnet_mul_fwd(int64_t n, const float *a, const float *b, float *y, void *stream) {
    return launch_map2(n, a, b, y, 0, MulOp{}, stream, "fnet_mul_fwd");
}

int fnet_mul_bwd(int64_t n, const float *dy, const float *other, float *dx, int accumulate, void *stream) {
    return launch_map2(n, dy, other, dx, accumulate, MulOp{}, stream, "fnet_mul_bwd");
}

int fnet_accumulate(int64_t n, const float *src, float *dst, int accumulate, void *stream) {
    return launch_map1(n, src, dst, accumulate, CopyOp{}, stream, "fnet_accumulate");
}

int fnet_copy_rows(int32_t batch, int32_t src_rows, int32_t src_row0, int32_t dst_rows, int32_t dst_row0,
                   int32_t rows, int32_t inner, const float *src, float *dst, int accumulate, void *stream) {
    FNET_REQUIRE(batch >= 0 && rows >= 0 && inner > 0 && src_row0 >= 0 && dst_row0 >= 0 &&
                     src_row0 + rows <= src_rows && dst_row0 + rows <= dst_rows,
                 "fnet_copy_rows: row range outside the tensors");
    FNET_REQUIRE(src && dst, "fnet_copy_rows: null pointer");
    int64_t block_elems = (int64_t)rows * inner;
    if (batch == 0 || block_elems == 0) return FNET_OK;
    launch_pdl(copy_rows_kernel, dim3(stream_grid((int64_t)batch * block_elems, 256)), dim3(256), 0, as_stream(stream), 
        batch, src_rows, src_row0, dst_rows, dst_row0, block_elems, inner, src, dst, accumulate);
    return check_launch("fnet_copy_rows");
}

int fnet_concat_rows_fwd(int32_t outer, int32_t rows_a, int32_t rows_b, int32_t inner, const float *a, const float *b,
                         float *y, void *stream) {
    FNET_REQUIRE(outer >= 0 && rows_a >= 0 && rows_b >= 0 && inner > 0 && a && b && y, "fnet_concat_rows_fwd: bad arguments");
    int64_t total = (int64_t)outer * (rows_a + rows_b) * inner;
    if (total == 0) return FNET_OK;
    const int64_t blk_a = (int64_t)rows_a * inner, blk_b = (int64_t)rows_b * inner;
    const bool small = total < 0x7fffffffLL - (1 << 26);
    cudaStream_t s = as_stream(stream);
    if (small && !(blk_a & 3) && !(blk_b & 3) && aligned16(a) && aligned16(b) && aligned16(y))
        launch_pdl(concat_rows_fwd_kernel<float4, int>, dim3(stream_grid(total / 4, 256)), dim3(256), 0, s, 
            outer, (int)(blk_a / 4), (int)(blk_b / 4), reinterpret_cast<const float4 *>(a), reinterpret_cast<const float4 *>(b),
            reinterpret_cast<float4 *>(y));
    else if (small)
        launch_pdl(concat_rows_fwd_kernel<float, int>, dim3(stream_grid(total, 256)), dim3(256), 0, s, outer, (int)blk_a, (int)blk_b, a, b, y);
    else
        launch_pdl(concat_rows_fwd_kernel<float, int64_t>, dim3(stream_grid(total, 256)), dim3(256), 0, s, outer, blk_a, blk_b, a, b, y);
    return check_launch("fnet_concat_rows_fwd");
}

int fnet_concat_rows_bwd(int32_t outer, int32_t rows_a, int32_t rows_b, int32_t inner, const float *dy, float *da,
                         float *db, int acc_a, int acc_b, void *stream) {
    FNET_REQUIRE(outer >= 0 && rows_a >= 0 && rows_b >= 0 && inner > 0 && dy, "fnet_concat_rows_bwd: bad arguments");
    int64_t total = (int64_t)outer * (rows_a + rows_b) * inner;
    if (total == 0 || (!da && !db)) return FNET_OK;
    const int64_t blk_a = (int64_t)rows_a * inner, blk_b = (int64_t)rows_b * inner;
    const bool small = total < 0x7fffffffLL - (1 << 26);
    cudaStream_t s = as_stream(stream);
    if (small && !(blk_a & 3) && !(blk_b & 3) && aligned16(dy) && aligned16(da) && aligned16(db))
        launch_pdl(concat_rows_bwd_kernel<float4, int>, dim3(stream_grid(total / 4, 256)), dim3(256), 0, s, 
            outer, (int)(blk_a / 4), (int)(blk_b / 4), reinterpret_cast<const float4 *>(dy), reinterpret_cast<float4 *>(da),
            reinterpret_cast<float4 *>(db), acc_a, acc_b);
    else if (small)
        launch_pdl(concat_rows_bwd_kernel<float, int>, dim3(stream_grid(total, 256)), dim3(256), 0, s, outer, (int)blk_a, (int)blk_b, dy, da, db, acc_a, acc_b);
    else
        launch_pdl(concat_rows_bwd_kernel<float, int64_t>, dim3(stream_grid(total, 256)), dim3(256), 0, s, outer, blk_a, blk_b, dy, da, db, acc_a, acc_b);
    return check_launch("fnet_concat_rows_bwd");
}

int fnet_concat_rows_bwd_act(int32_t outer, int32_t rows_a, int32_t rows_b, int32_t inner, const float *dy, float *da, float *db,
                             int acc_a, int acc_b, int act_a, const float *ya, int act_b, const float *yb, void *stream) {
    FNET_REQUIRE(outer >= 0 && rows_a >= 0 && rows_b >= 0 && inner > 0 && dy, "fnet_concat_rows_bwd_act: bad arguments");
    FNET_REQUIRE(act_a >= FNET_ACT_NONE && act_a <= FNET_ACT_SIGMOID && act_b >= FNET_ACT_NONE && act_b <= FNET_ACT_SIGMOID &&
                 (!act_a || ya) && (!act_b || yb), "fnet_concat_rows_bwd_act: relu / tanh / sigmoid with the operand's forward output, or none");
    int64_t total = (int64_t)outer * (rows_a + rows_b) * inner;
    if (total == 0 || (!da && !db)) return FNET_OK;
    FNET_REQUIRE(total < 0x7fffffffLL - (1 << 26), "fnet_concat_rows_bwd_act: tensor too large");
    const int64_t blk_a = (int64_t)rows_a * inner, blk_b = (int64_t)rows_b * inner;
    cudaStream_t s = as_stream(stream);
    if (!(blk_a & 3) && !(blk_b & 3) && aligned16(dy) && aligned16(da) && aligned16(db) && aligned16(ya) && aligned16(yb))
        launch_pdl(concat_rows_bwd_act_kernel<float4>, dim3(stream_grid(total / 4, 256)), dim3(256), 0, s, outer, (int)(blk_a / 4), (int)(blk_b / 4),
                   reinterpret_cast<const float4 *>(dy), reinterpret_cast<float4 *>(da), reinterpret_cast<float4 *>(db), acc_a, acc_b,
                   act_a, reinterpret_cast<const float4 *>(ya), act_b, reinterpret_cast<const float4 *>(yb));
    else
        launch_pdl(concat_rows_bwd_act_kernel<float>, dim3(stream_grid(total, 256)), dim3(256), 0, s, outer, (int)blk_a, (int)blk_b, dy, da, db,
                   acc_a, acc_b, act_a, ya, act_b, yb);
    return check_launch("fnet_concat_rows_bwd_act");
}

int fnet_pad2d_fwd(int32_t batch, int32_t h, int32_t w, int32_t c, int32_t pad_h, int32_t pad_w, const float *x,
                   float *y, void *stream) {
    FNET_REQUIRE(batch >= 0 && h > 0 && w > 0 && c > 0 && pad_h >= 0 && pad_w >= 0 && x && y, "fnet_pad2d_fwd: bad arguments");
    if (batch == 0) return FNET_OK;
    int64_t total = (int64_t)batch * (h + 2 * pad_h) * (w + 2 * pad_w) * c;
    launch_pdl(pad2d_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), batch, h, w, c, pad_h, pad_w, x, y);
    return check_launch("fnet_pad2d_fwd");
}

int fnet_pad2d_bwd(int32_t batch, int32_t h, int32_t w, int32_t c, int32_t pad_h, int32_t pad_w, const float *dy,
                   float *dx, int accumulate, void *stream) {
    FNET_REQUIRE(batch >= 0 && h > 0 && w > 0 && c > 0 && pad_h >= 0 && pad_w >= 0 && dy && dx, "fnet_pad2d_bwd: bad arguments");
    if (batch == 0) return FNET_OK;
    launch_pdl(pad2d_bwd_kernel, dim3(stream_grid((int64_t)batch * h * w * c, 256)), dim3(256), 0, as_stream(stream), 
        batch, h, w, c, pad_h, pad_w, dy, dx, accumulate);
    return check_launch("fnet_pad2d_bwd");
}

}  // extern "C"

// Batch normalisation over axis 1 of a tensor viewed as [outer, A, inner]  (Keras axis=1 on
// NHWC: one statistic per image ROW h, over batch x width x channels).
//
// Two launches per direction:
//   1. partial sums per (a, split) -> workspace       (reads x [and dy] once)
//   2. every CTA of row a folds the <= 64 partials in a fixed order, then normalises its chunk
//      (reads x [and dy] again, writes y / dx); the first CTA of the row also writes the saved
//      statistics, the moving averages or dgamma/dbeta.
// Sums are taken on (x - x[first element of the row]) so that one pass in fp32 does not lose
// the variance to cancellation.  Algorithmic traffic: forward 2 reads + 1 write of x,
// backward 2 reads of x, 2 of dy, 1 write; HBM-bound.
#include "common.cuh"

namespace fnet {

constexpr int kBnMaxSplits = 64;

struct BnShape {
    int outer, a, inner, splits;
    int outer_per_split;
};

static BnShape bn_shape(int outer, int a, int inner) {
    BnShape s;
    s.outer = outer; s.a = a; s.inner = inner;
    // eight 256-thread CTAs per SM: every thread has one 16-byte load in flight at a time, so the bytes in flight -- and the
    // bandwidth -- are set by how many threads are resident (two CTAs per SM moved 1.3 TB/s on 25 MB)
    int want = ceil_div(8 * kNumSMs, a);
    if (want > kBnMaxSplits) want = kBnMaxSplits;
    if (want > outer) want = outer;
    if (want < 1) want = 1;
    s.outer_per_split = ceil_div(outer, want);
    s.splits = ceil_div(outer, s.outer_per_split);
    return s;
}

// element e of the chunk [o0, o1) x inner of row a.  32-bit on purpose: a 64-bit division per element made
// these kernels instruction-bound (outer * inner < 2^31 is checked on the host).
__device__ __forceinline__ int64_t bn_index(const BnShape &s, int a, int o0, unsigned e) {
    unsigned q = e / (unsigned)s.inner;
    unsigned j = e - q * (unsigned)s.inner;
    return ((int64_t)(o0 + q) * s.a + a) * s.inner + j;
}
// the same in units of V floats (V = 4: inner % 4 == 0 and 16-byte aligned tensors; one index decode per float4 --
// the scalar kernels are issue-bound at 1 TB/s)
template <int V>
__device__ __forceinline__ int64_t bn_index_v(const BnShape &s, int a, int o0, unsigned e) {
    const unsigned inner_v = (unsigned)s.inner / V;
    unsigned q = e / inner_v;
    unsigned j = e - q * inner_v;
    return ((int64_t)(o0 + q) * s.a + a) * inner_v + j;
}
template <int V> struct BnVec;
template <> struct BnVec<1> {
    typedef float T;
    static __device__ __forceinline__ void get(const float &v, float (&o)[1]) { o[0] = v; }
    static __device__ __forceinline__ float make(const float (&o)[1]) { return o[0]; }
};
template <> struct BnVec<4> {
    typedef float4 T;
    static __device__ __forceinline__ void get(const float4 &v, float (&o)[4]) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
    static __device__ __forceinline__ float4 make(const float (&o)[4]) { return make_float4(o[0], o[1], o[2], o[3]); }
};

template <int V>
__global__ void __launch_bounds__(256) bn_stats_partial_kernel(BnShape s, const float *__restrict__ x,
                                                               float *__restrict__ partial) {
    pdl_prologue();
    typedef typename BnVec<V>::T T;
    __shared__ float red[32];
    int a = blockIdx.x, part = blockIdx.y;
    int o0 = part * s.outer_per_split, o1 = min(s.outer, o0 + s.outer_per_split);
    float shift = __ldg(x + (int64_t)a * s.inner);
    int64_t count = (int64_t)(o1 - o0) * (s.inner / V);
    float s1 = 0.f, s2 = 0.f;
    const T *xv = reinterpret_cast<const T *>(x);
    for (unsigned e = threadIdx.x; e < (unsigned)count; e += blockDim.x) {
        float f[V];
        BnVec<V>::get(__ldg(xv + bn_index_v<V>(s, a, o0, e)), f);
#pragma unroll
        for (int u = 0; u < V; ++u) {
            float v = f[u] - shift;
            s1 += v;
            s2 = fmaf(v, v, s2);
        }
    }
    s1 = block_sum(s1, red);
    s2 = block_sum(s2, red);
    if (threadIdx.x == 0) {
        partial[((int64_t)a * kBnMaxSplits + part) * 2 + 0] = s1;
        partial[((int64_t)a * kBnMaxSplits + part) * 2 + 1] = s2;
    }
}

template <int V>
__global__ void __launch_bounds__(256) bn_normalize_train_kernel(BnShape s, float eps, float momentum,
                                                                 const float *__restrict__ x,
                                                                 const float *__restrict__ gamma,
                                                                 const float *__restrict__ beta,
                                                                 const float *__restrict__ partial, float *__restrict__ y,
                                                                 float *__restrict__ save_mean, float *__restrict__ save_rstd,
                                                                 float *__restrict__ moving_mean,
                                                                 float *__restrict__ moving_var) {
    pdl_prologue();
    __shared__ float stat[2];
    int a = blockIdx.x, part = blockIdx.y;
    if (threadIdx.x == 0) {
        float s1 = 0.f, s2 = 0.f;
        for (int z = 0; z < s.splits; ++z) {
            s1 += partial[((int64_t)a * kBnMaxSplits + z) * 2 + 0];
            s2 += partial[((int64_t)a * kBnMaxSplits + z) * 2 + 1];
        }
        float n = (float)s.outer * (float)s.inner;
        float d = s1 / n;
        float mean = __ldg(x + (int64_t)a * s.inner) + d;
        float var = fmaxf(s2 / n - d * d, 0.f);
        float rstd = rsqrtf(var + eps);
        rstd = rstd * (1.5f - 0.5f * (var + eps) * rstd * rstd);   // one Newton step: full fp32 accuracy
        stat[0] = mean;
        stat[1] = rstd;
        if (part == 0) {
            save_mean[a] = mean;
            save_rstd[a] = rstd;
            // Keras: moving <- moving*momentum + value*(1-momentum); variance corrected by n/(n-(1+eps))
            moving_mean[a] = moving_mean[a] * momentum + mean * (1.f - momentum);
            moving_var[a] = moving_var[a] * momentum + (var * n / (n - (1.f + eps))) * (1.f - momentum);
        }
    }
    __syncthreads();
    // tf.nn.batch_normalization: inv = rsqrt(var+eps)*gamma; y = x*inv + (beta - mean*inv)
    float inv = stat[1] * __ldg(gamma + a);
    float off = __ldg(beta + a) - stat[0] * inv;
    int o0 = part * s.outer_per_split, o1 = min(s.outer, o0 + s.outer_per_split);
    typedef typename BnVec<V>::T T;
    int64_t count = (int64_t)(o1 - o0) * (s.inner / V);
    const T *xv = reinterpret_cast<const T *>(x);
    T *yv = reinterpret_cast<T *>(y);
    for (unsigned e = threadIdx.x; e < (unsigned)count; e += blockDim.x) {
        int64_t at = bn_index_v<V>(s, a, o0, e);
        float f[V];
        BnVec<V>::get(__ldg(xv + at), f);
#pragma unroll
        for (int u = 0; u < V; ++u) f[u] = fmaf(f[u], inv, off);
        yv[at] = BnVec<V>::make(f);
    }
}

__global__ void __launch_bounds__(256) bn_infer_kernel(int64_t total, int a_len, int inner, float eps,
                                                       const float *__restrict__ x, const float *__restrict__ gamma,
                                                       const float *__restrict__ beta, const float *__restrict__ mean,
                                                       const float *__restrict__ var, float *__restrict__ y) {
    pdl_prologue();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int a = (int)(((unsigned)i / (unsigned)inner) % (unsigned)a_len);
        float v = __ldg(var + a) + eps;
        float rstd = rsqrtf(v);
        rstd = rstd * (1.5f - 0.5f * v * rstd * rstd);
        float inv = rstd * __ldg(gamma + a);
        y[i] = fmaf(__ldg(x + i), inv, __ldg(beta + a) - __ldg(mean + a) * inv);
    }
}

// inference-mode input gradient: the statistics are constants, so dx = dy * gamma / sqrt(moving_var + eps)
__global__ void __launch_bounds__(256) bn_infer_bwd_kernel(int64_t total, int a_len, int inner, float eps,
                                                           const float *__restrict__ dy, const float *__restrict__ gamma,
                                                           const float *__restrict__ var, float *__restrict__ dx, int accumulate) {
    pdl_prologue();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int a = (int)(((unsigned)i / (unsigned)inner) % (unsigned)a_len);
        float v = __ldg(var + a) + eps;
        float rstd = rsqrtf(v);
        rstd = rstd * (1.5f - 0.5f * v * rstd * rstd);
        float g = __ldg(dy + i) * rstd * __ldg(gamma + a);
        dx[i] = accumulate ? dx[i] + g : g;
    }
}

template <int V>
__global__ void __launch_bounds__(256) bn_bwd_partial_kernel(BnShape s, const float *__restrict__ x,
                                                             const float *__restrict__ dy,
                                                             const float *__restrict__ save_mean,
                                                             const float *__restrict__ save_rstd,
                                                             float *__restrict__ partial) {
    pdl_prologue();
    __shared__ float red[32];
    int a = blockIdx.x, part = blockIdx.y;
    int o0 = part * s.outer_per_split, o1 = min(s.outer, o0 + s.outer_per_split);
    float mean = __ldg(save_mean + a), rstd = __ldg(save_rstd + a);
    typedef typename BnVec<V>::T T;
    int64_t count = (int64_t)(o1 - o0) * (s.inner / V);
    float sb = 0.f, sg = 0.f;
    const T *xv = reinterpret_cast<const T *>(x), *dyv = reinterpret_cast<const T *>(dy);
    for (unsigned e = threadIdx.x; e < (unsigned)count; e += blockDim.x) {
        int64_t at = bn_index_v<V>(s, a, o0, e);
        float g[V], f[V];
        BnVec<V>::get(__ldg(dyv + at), g);
        BnVec<V>::get(__ldg(xv + at), f);
#pragma unroll
        for (int u = 0; u < V; ++u) {
            sb += g[u];
            sg = fmaf(g[u], (f[u] - mean) * rstd, sg);
        }
    }
    sb = block_sum(sb, red);
    sg = block_sum(sg, red);
    if (threadIdx.x == 0) {
        partial[((int64_t)a * kBnMaxSplits + part) * 2 + 0] = sb;
        partial[((int64_t)a * kBnMaxSplits + part) * 2 + 1] = sg;
    }
}

template <int V>
__global__ void __launch_bounds__(256) bn_bwd_dx_kernel(BnShape s, const float *__restrict__ x,
                                                        const float *__restrict__ dy, const float *__restrict__ gamma,
                                                        const float *__restrict__ save_mean,
                                                        const float *__restrict__ save_rstd,
                                                        const float *__restrict__ partial, float *__restrict__ dx,
                                                        float *__restrict__ dgamma, float *__restrict__ dbeta,
                                                        int accumulate, int act) {
    pdl_prologue();
    __shared__ float sums[2];
    int a = blockIdx.x, part = blockIdx.y;
    if (threadIdx.x == 0) {
        float sb = 0.f, sg = 0.f;
        for (int z = 0; z < s.splits; ++z) {
            sb += partial[((int64_t)a * kBnMaxSplits + z) * 2 + 0];
            sg += partial[((int64_t)a * kBnMaxSplits + z) * 2 + 1];
        }
        sums[0] = sb;
        sums[1] = sg;
        if (part == 0) {
            dbeta[a] = sb;
            dgamma[a] = sg;
        }
    }
    __syncthreads();
    float n = (float)s.outer * (float)s.inner;
    float mean = __ldg(save_mean + a), rstd = __ldg(save_rstd + a);
    float k = __ldg(gamma + a) * rstd / n;
    float sb = sums[0], sg = sums[1];
    int o0 = part * s.outer_per_split, o1 = min(s.outer, o0 + s.outer_per_split);
    typedef typename BnVec<V>::T T;
    int64_t count = (int64_t)(o1 - o0) * (s.inner / V);
    const T *xv = reinterpret_cast<const T *>(x), *dyv = reinterpret_cast<const T *>(dy);
    T *dxv = reinterpret_cast<T *>(dx);
    for (unsigned e = threadIdx.x; e < (unsigned)count; e += blockDim.x) {
        int64_t at = bn_index_v<V>(s, a, o0, e);
        float f[V], g[V], o[V];
        BnVec<V>::get(__ldg(xv + at), f);
        BnVec<V>::get(__ldg(dyv + at), g);
        if (accumulate) BnVec<V>::get(dxv[at], o);
#pragma unroll
        for (int u = 0; u < V; ++u) {
            float xhat = (f[u] - mean) * rstd;
            float v = k * (n * g[u] - sb - xhat * sg);
            v = accumulate ? o[u] + v : v;
            // x is the OUTPUT of the layer below: its activation derivative comes from x itself, no extra read
            o[u] = act ? v * act_grad(f[u], act) : v;
        }
        dxv[at] = BnVec<V>::make(o);
    }
}

}  // namespace fnet

using namespace fnet;

extern "C" {

size_t fnet_bn_workspace_bytes(int32_t a) { return a > 0 ? (size_t)a * kBnMaxSplits * 2 * sizeof(float) : 0; }

int fnet_bn_fwd_train(int32_t outer, int32_t a, int32_t inner, float eps, float momentum, const float *x,
                      const float *gamma, const float *beta, float *y, float *save_mean, float *save_rstd,
                      float *moving_mean, float *moving_var, void *ws, size_t ws_bytes, void *stream) {
    FNET_REQUIRE(outer > 0 && a > 0 && inner > 0 && (int64_t)outer * a * inner < 0x7fffffffLL, "fnet_bn_fwd_train: bad shape");
    FNET_REQUIRE(x && gamma && beta && y && save_mean && save_rstd && moving_mean && moving_var,
                 "fnet_bn_fwd_train: null pointer");
    if (!ws || ws_bytes < fnet_bn_workspace_bytes(a)) {
        set_error("fnet_bn_fwd_train: workspace of %zu bytes needed, %zu given", fnet_bn_workspace_bytes(a), ws_bytes);
        return FNET_ERR_WORKSPACE;
    }
    BnShape s = bn_shape(outer, a, inner);
    dim3 grid(a, s.splits);
    float *partial = reinterpret_cast<float *>(ws);
    const bool vec = !(inner & 3) && !((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15);
    if (vec) launch_pdl(bn_stats_partial_kernel<4>, dim3(grid), dim3(256), 0, as_stream(stream), s, x, partial);
    else launch_pdl(bn_stats_partial_kernel<1>, dim3(grid), dim3(256), 0, as_stream(stream), s, x, partial);
    int rc = check_launch("fnet_bn_fwd_train");
    if (rc) return rc;
    if (vec)
        launch_pdl(bn_normalize_train_kernel<4>, dim3(grid), dim3(256), 0, as_stream(stream), s, eps, momentum, x, gamma, beta, partial, y,
                                                                          save_mean, save_rstd, moving_mean, moving_var);
    else
        launch_pdl(bn_normalize_train_kernel<1>, dim3(grid), dim3(256), 0, as_stream(stream), s, eps, momentum, x, gamma, beta, partial, y,
                                                                          save_mean, save_rstd, moving_mean, moving_var);
    return check_launch("fnet_bn_fwd_train");
}

int fnet_bn_fwd_infer(int32_t outer, int32_t a, int32_t inner, float eps, const float *x, const float *gamma,
                      const float *beta, const float *moving_mean, const float *moving_var, float *y, void *stream) {
    FNET_REQUIRE(outer >= 0 && a > 0 && inner > 0 && (int64_t)outer * a * inner < 0x7fffffffLL, "fnet_bn_fwd_infer: bad shape");
    FNET_REQUIRE(x && gamma && beta && moving_mean && moving_var && y, "fnet_bn_fwd_infer: null pointer");
    int64_t total = (int64_t)outer * a * inner;
    if (total == 0) return FNET_OK;
    launch_pdl(bn_infer_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), total, a, inner, eps, x, gamma, beta,
                                                                            moving_mean, moving_var, y);
    return check_launch("fnet_bn_fwd_infer");
}

int fnet_bn_bwd_infer(int32_t outer, int32_t a, int32_t inner, float eps, const float *dy, const float *gamma,
                      const float *moving_var, float *dx, int accumulate, void *stream) {
    FNET_REQUIRE(outer >= 0 && a > 0 && inner > 0 && (int64_t)outer * a * inner < 0x7fffffffLL, "fnet_bn_bwd_infer: bad shape");
    FNET_REQUIRE(dy && gamma && moving_var && dx, "fnet_bn_bwd_infer: null pointer");
    int64_t total = (int64_t)outer * a * inner;
    if (total == 0) return FNET_OK;
    launch_pdl(bn_infer_bwd_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), total, a, inner, eps, dy, gamma,
               moving_var, dx, accumulate);
    return check_launch("fnet_bn_bwd_infer");
}

int fnet_bn_bwd(int32_t outer, int32_t a, int32_t inner, const float *x, const float *dy, const float *gamma,
                const float *save_mean, const float *save_rstd, float *dx, float *dgamma, float *dbeta,
                int accumulate, void *ws, size_t ws_bytes, void *stream) {
    return fnet_bn_bwd_act(outer, a, inner, x, dy, gamma, save_mean, save_rstd, dx, dgamma, dbeta, accumulate, FNET_ACT_NONE, ws,
                           ws_bytes, stream);
}

int fnet_bn_bwd_act(int32_t outer, int32_t a, int32_t inner, const float *x, const float *dy, const float *gamma,
                    const float *save_mean, const float *save_rstd, float *dx, float *dgamma, float *dbeta,
                    int accumulate, int act, void *ws, size_t ws_bytes, void *stream) {
    FNET_REQUIRE(act >= FNET_ACT_NONE && act <= FNET_ACT_SIGMOID, "fnet_bn_bwd_act: relu / tanh / sigmoid, or none");
    FNET_REQUIRE(outer > 0 && a > 0 && inner > 0 && (int64_t)outer * a * inner < 0x7fffffffLL, "fnet_bn_bwd: bad shape");
    FNET_REQUIRE(x && dy && gamma && save_mean && save_rstd && dx && dgamma && dbeta, "fnet_bn_bwd: null pointer");
    if (!ws || ws_bytes < fnet_bn_workspace_bytes(a)) {
        set_error("fnet_bn_bwd: workspace of %zu bytes needed, %zu given", fnet_bn_workspace_bytes(a), ws_bytes);
        return FNET_ERR_WORKSPACE;
    }
    BnShape s = bn_shape(outer, a, inner);
    dim3 grid(a, s.splits);
    float *partial = reinterpret_cast<float *>(ws);
    const bool vec = !(inner & 3) &&
                     !((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx)) & 15);
    if (vec) launch_pdl(bn_bwd_partial_kernel<4>, dim3(grid), dim3(256), 0, as_stream(stream), s, x, dy, save_mean, save_rstd, partial);
    else launch_pdl(bn_bwd_partial_kernel<1>, dim3(grid), dim3(256), 0, as_stream(stream), s, x, dy, save_mean, save_rstd, partial);
    int rc = check_launch("fnet_bn_bwd");
    if (rc) return rc;
    if (vec)
        launch_pdl(bn_bwd_dx_kernel<4>, dim3(grid), dim3(256), 0, as_stream(stream), s, x, dy, gamma, save_mean, save_rstd, partial, dx,
                                                                 dgamma, dbeta, accumulate, act);
    else
        launch_pdl(bn_bwd_dx_kernel<1>, dim3(grid), dim3(256), 0, as_stream(stream), s, x, dy, gamma, save_mean, save_rstd, partial, dx,
                                                                 dgamma, dbeta, accumulate, act);
    return check_launch("fnet_bn_bwd");
}

}  // extern "C"

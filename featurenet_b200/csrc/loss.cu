// Softmax + categorical cross-entropy + categorical accuracy in one launch.
// Follows Keras' categorical_crossentropy(from_logits=False) on the output of the softmax "out"
// layer: renormalise p, clip to [1e-7, 1-1e-7], -sum t*log p; the optional gradient goes back
// through the clip (zero where clipped), the renormalisation and the softmax, so it is
// dL/d(logits).  With from_logits the kernel computes the softmax itself (writing the
// probabilities, in place if asked), which saves the separate softmax launch of the head.
//
// One thread per sample (rows are 10..100 floats and a batch is <= a few thousand rows), one CTA
// of up to 256 threads: the batch loss is reduced in a fixed order (shuffle tree, then warp
// order) and added to the double accumulator by a single thread -- deterministic.  Rows of up to
// 16 classes (the 10-class heads of MNIST / CIFAR-10) are held in registers: one read, one write.
#include "common.cuh"
#include "loss_row.cuh"

namespace fnet {

template <int CMAX>
struct RegRow {
    float v[CMAX];          // only ever indexed with compile-time constants, so it lives in registers
};

__global__ void __launch_bounds__(256) softmax_cce_kernel(int batch, int classes, const float *input /* may alias probs_out */,
                                                           const float *__restrict__ targets, float scale,
                                                           int from_logits, float *probs_out, float *__restrict__ dlogits,
                                                           double *__restrict__ loss_sum, long long *__restrict__ correct) {
    pdl_prologue();
    __shared__ double warp_loss[32];
    __shared__ int warp_correct[32];
    double my_loss = 0.0;
    int my_correct = 0;
    for (int row = threadIdx.x; row < batch; row += blockDim.x) {
        const float *in = input + (int64_t)row * classes;
        const float *tg = targets + (int64_t)row * classes;
        float *dz = dlogits ? dlogits + (int64_t)row * classes : nullptr;
        bool hit = false;
        float loss;
        if (classes <= 16) {
            RegRow<16> p, t;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                p.v[c] = c < classes ? in[c] : 0.f;
                t.v[c] = c < classes ? __ldg(tg + c) : 0.f;
            }
            if (from_logits) {
                float mx = -INFINITY, se = 0.f;
#pragma unroll
                for (int c = 0; c < 16; ++c) if (c < classes) mx = fmaxf(mx, p.v[c]);
#pragma unroll
                for (int c = 0; c < 16; ++c) if (c < classes) { p.v[c] = expf(p.v[c] - mx); se += p.v[c]; }
                float *q = probs_out + (int64_t)row * classes;
#pragma unroll
                for (int c = 0; c < 16; ++c) if (c < classes) { p.v[c] = p.v[c] / se; q[c] = p.v[c]; }
            }
            float dzr[16];
            // fully unrolled register version of row_loss (indices must be compile-time constants)
            float sum = 0.f, bp = -INFINITY, bt = -INFINITY;
            int ip = 0, it = 0;
#pragma unroll
            for (int c = 0; c < 16; ++c) if (c < classes) {
                sum += p.v[c];
                if (p.v[c] > bp) { bp = p.v[c]; ip = c; }
                if (t.v[c] > bt) { bt = t.v[c]; it = c; }
            }
            hit = ip == it;
            float dot = 0.f, g[16];
            loss = 0.f;
#pragma unroll
            for (int c = 0; c < 16; ++c) if (c < classes) {
                float q = p.v[c] / sum;
                float cl = fminf(fmaxf(q, kClipEps), 1.f - kClipEps);
                loss -= t.v[c] * logf(cl);
                g[c] = (q >= kClipEps && q <= 1.f - kClipEps) ? -t.v[c] / cl : 0.f;
                dot = fmaf(g[c], q, dot);
            }
            if (dz) {
                float dot2 = 0.f;
#pragma unroll
                for (int c = 0; c < 16; ++c) if (c < classes) dot2 = fmaf((g[c] - dot) / sum, p.v[c], dot2);
#pragma unroll
                for (int c = 0; c < 16; ++c) if (c < classes) dzr[c] = scale * p.v[c] * ((g[c] - dot) / sum - dot2);
#pragma unroll
                for (int c = 0; c < 16; ++c) if (c < classes) dz[c] = dzr[c];
            }
        } else {
            const float *pp = in;
            if (from_logits) {
                // keras softmax: exp(z - max) / sum, written where the loss (and the caller) will read it
                float mx = -INFINITY, se = 0.f;
                for (int c = 0; c < classes; ++c) mx = fmaxf(mx, in[c]);
                for (int c = 0; c < classes; ++c) se += expf(in[c] - mx);
                float *q = probs_out + (int64_t)row * classes;
                for (int c = 0; c < classes; ++c) q[c] = expf(in[c] - mx) / se;    // in place is fine: one thread owns the row
                pp = q;
            }
            loss = row_loss(classes, MemRow{pp}, MemRow{tg}, scale, dz, hit);
        }
        my_loss += (double)loss;
        my_correct += hit ? 1 : 0;
    }
    // fixed-order block reduction
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        my_loss += __shfl_xor_sync(0xffffffffu, my_loss, o);
        my_correct += __shfl_xor_sync(0xffffffffu, my_correct, o);
    }
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) { warp_loss[wid] = my_loss; warp_correct[wid] = my_correct; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double l = 0.0;
        int k = 0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { l += warp_loss[i]; k += warp_correct[i]; }
        if (loss_sum) *loss_sum += l;
        if (correct) *correct += (long long)k;
    }
}

// The same computation with G lanes per row (G = 16 or 32 >= classes): every exp / log / divide chain is one
// element long and the row sums are shuffle trees, so the launch costs about one dependent chain instead of
// `classes` of them -- this kernel sits on the critical path of every training step.  One CTA (fixed-order
// loss reduction), rows are processed blockDim/G at a time.
template <int G>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int G>
__device__ __forceinline__ int group_argmax(float v, int idx) {      // first index on ties
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, v, o);
        int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    return idx;
}

// R rows per lane group are processed side by side: their shuffle / exp / log chains are independent, so a batch
// of 128 ten-class rows is ONE dependent chain for the whole (single-CTA) launch instead of two.
template <int G, int R>
__device__ __forceinline__ void group_sum_r(float (&v)[R]) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1)
#pragma unroll
        for (int u = 0; u < R; ++u) v[u] += __shfl_xor_sync(0xffffffffu, v[u], o);
}

template <int G, int R>
__global__ void __launch_bounds__(1024) softmax_cce_group_kernel(int batch, int classes, const float *input /* may alias probs_out */,
                                                                  const float *__restrict__ targets, float scale,
                                                                  int from_logits, float *probs_out, float *__restrict__ dlogits,
                                                                  double *__restrict__ loss_sum, long long *__restrict__ correct) {
    pdl_prologue();
    __shared__ double group_loss[1024 / G];
    __shared__ int group_correct[1024 / G];
    const int c = threadIdx.x % G, group = threadIdx.x / G, groups = blockDim.x / G;
    double my_loss = 0.0;
    int my_correct = 0;
    for (int row0 = 0; row0 < batch; row0 += groups * R) {           // same trip count for every thread (shuffles)
        int row[R];
        bool live[R];
        int64_t at[R];
        float xin[R], tv[R], p[R];
#pragma unroll
        for (int u = 0; u < R; ++u) {
            row[u] = row0 + u * groups + group;
            live[u] = row[u] < batch && c < classes;
            at[u] = (int64_t)row[u] * classes + c;
            xin[u] = live[u] ? input[at[u]] : 0.f;
            tv[u] = live[u] ? __ldg(targets + at[u]) : 0.f;
            p[u] = xin[u];
        }
        if (from_logits) {
            float mx[R], e[R];
#pragma unroll
            for (int u = 0; u < R; ++u) mx[u] = live[u] ? xin[u] : -INFINITY;
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1)
#pragma unroll
                for (int u = 0; u < R; ++u) mx[u] = fmaxf(mx[u], __shfl_xor_sync(0xffffffffu, mx[u], o));
            float se[R];
#pragma unroll
            for (int u = 0; u < R; ++u) { e[u] = live[u] ? expf(xin[u] - mx[u]) : 0.f; se[u] = e[u]; }
            group_sum_r<G, R>(se);
#pragma unroll
            for (int u = 0; u < R; ++u) {
                p[u] = e[u] / se[u];
                if (live[u]) probs_out[at[u]] = p[u];
            }
        }
        float sum[R];
        int ip[R], it[R];
#pragma unroll
        for (int u = 0; u < R; ++u) {
            sum[u] = p[u];
            ip[u] = group_argmax<G>(live[u] ? p[u] : -INFINITY, c);
            it[u] = group_argmax<G>(live[u] ? tv[u] : -INFINITY, c);
        }
        group_sum_r<G, R>(sum);
        float q[R], cl[R], loss[R], g[R], dot[R];
#pragma unroll
        for (int u = 0; u < R; ++u) {
            q[u] = p[u] / sum[u];
            cl[u] = fminf(fmaxf(q[u], kClipEps), 1.f - kClipEps);
            loss[u] = live[u] ? -tv[u] * logf(cl[u]) : 0.f;
            g[u] = (live[u] && q[u] >= kClipEps && q[u] <= 1.f - kClipEps) ? -tv[u] / cl[u] : 0.f;
            dot[u] = g[u] * q[u];
        }
        group_sum_r<G, R>(loss);
        group_sum_r<G, R>(dot);
        if (dlogits) {
            float dp[R], dot2[R];
#pragma unroll
            for (int u = 0; u < R; ++u) {
                dp[u] = (g[u] - dot[u]) / sum[u];                    // dL/dp through the renormalisation
                dot2[u] = live[u] ? dp[u] * p[u] : 0.f;              // softmax backward
            }
            group_sum_r<G, R>(dot2);
#pragma unroll
            for (int u = 0; u < R; ++u)
                if (live[u]) dlogits[at[u]] = scale * p[u] * (dp[u] - dot2[u]);
        }
#pragma unroll
        for (int u = 0; u < R; ++u)
            if (c == 0 && row[u] < batch) { my_loss += (double)loss[u]; my_correct += ip[u] == it[u] ? 1 : 0; }
    }
    if (c == 0) { group_loss[group] = my_loss; group_correct[group] = my_correct; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double l = 0.0;
        int k = 0;
        for (int i = 0; i < groups; ++i) { l += group_loss[i]; k += group_correct[i]; }     // fixed order
        if (loss_sum) *loss_sum += l;
        if (correct) *correct += (long long)k;
    }
}

}  // namespace fnet

using namespace fnet;

extern "C" int fnet_softmax_cce(int32_t batch, int32_t classes, const float *input, const float *targets, float scale,
                                int from_logits, float *probs_out, float *dlogits, double *loss_sum, int64_t *correct,
                                void *stream) {
    FNET_REQUIRE(batch >= 0 && classes > 0 && input && targets, "fnet_softmax_cce: bad arguments");
    FNET_REQUIRE(!from_logits || probs_out != nullptr, "fnet_softmax_cce: from_logits needs a place for the probabilities");
    FNET_REQUIRE(dlogits == nullptr || dlogits != probs_out, "fnet_softmax_cce: dlogits must not alias the probabilities");
    if (batch == 0) return FNET_OK;
    if (classes <= 32) {
        const int lanes = classes <= 16 ? 16 : 32;
        long long want = (long long)batch * lanes;
        int threads = want >= 1024 ? 1024 : (int)((want + 31) / 32) * 32;
        if (lanes == 16)
            launch_pdl(softmax_cce_group_kernel<16, 2>, dim3(1), dim3(threads), 0, as_stream(stream), batch, classes, input, targets, scale, from_logits,
                                                                               probs_out, dlogits, loss_sum,
                                                                               reinterpret_cast<long long *>(correct));
        else
            launch_pdl(softmax_cce_group_kernel<32, 2>, dim3(1), dim3(threads), 0, as_stream(stream), batch, classes, input, targets, scale, from_logits,
                                                                               probs_out, dlogits, loss_sum,
                                                                               reinterpret_cast<long long *>(correct));
        return check_launch("fnet_softmax_cce");
    }
    int threads = batch >= 256 ? 256 : ((batch + 31) / 32) * 32;
    launch_pdl(softmax_cce_kernel, dim3(1), dim3(threads), 0, as_stream(stream), batch, classes, input, targets, scale, from_logits, probs_out,
                                                             dlogits, loss_sum, reinterpret_cast<long long *>(correct));
    return check_launch("fnet_softmax_cce");
}

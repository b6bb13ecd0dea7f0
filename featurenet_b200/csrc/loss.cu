// Softmax cross-entropy on probabilities + categorical accuracy, one warp per sample.
// Follows Keras' categorical_crossentropy(from_logits=False): renormalise, clip to
// [1e-7, 1-1e-7], -sum t*log p; the optional gradient goes back through the clip (zero where
// clipped), the renormalisation and the softmax that produced p, so it is dL/d(logits).
// Loss is accumulated in double and the correct-count as an integer (exact).
#include "common.cuh"

namespace fnet {

__global__ void __launch_bounds__(256) softmax_cce_kernel(int batch, int classes, const float *__restrict__ probs,
                                                          const float *__restrict__ targets, float scale,
                                                          float *__restrict__ dlogits, double *__restrict__ loss_sum,
                                                          unsigned long long *__restrict__ correct) {
    __shared__ double block_loss[8];
    __shared__ int block_correct[8];
    const float kEps = 1e-7f;
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    double my_loss = 0.0;
    int my_correct = 0;
    for (int row = blockIdx.x * 8 + wid; row < batch; row += gridDim.x * 8) {
        const float *p = probs + (int64_t)row * classes;
        const float *t = targets + (int64_t)row * classes;
        // pass 1: sum(p), argmax p, argmax t (first index on ties)
        float sum = 0.f, bp = -INFINITY, bt = -INFINITY;
        int ip = 0x7fffffff, it = 0x7fffffff;
        for (int c = lane; c < classes; c += 32) {
            float pv = p[c], tv = t[c];
            sum += pv;
            if (pv > bp) { bp = pv; ip = c; }
            if (tv > bt) { bt = tv; it = c; }
        }
        sum = warp_sum(sum);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float op = __shfl_xor_sync(0xffffffffu, bp, o);
            int oi = __shfl_xor_sync(0xffffffffu, ip, o);
            if (op > bp || (op == bp && oi < ip)) { bp = op; ip = oi; }
            float ot = __shfl_xor_sync(0xffffffffu, bt, o);
            int oj = __shfl_xor_sync(0xffffffffu, it, o);
            if (ot > bt || (ot == bt && oj < it)) { bt = ot; it = oj; }
        }
        // pass 2: loss, and g = dL/dq masked by the clip; dot = sum g*q
        float loss = 0.f, dot = 0.f;
        for (int c = lane; c < classes; c += 32) {
            float q = p[c] / sum;
            float cl = fminf(fmaxf(q, kEps), 1.f - kEps);
            float tv = t[c];
            loss -= tv * logf(cl);
            float g = (q >= kEps && q <= 1.f - kEps) ? -tv / cl : 0.f;
            dot = fmaf(g, q, dot);
        }
        loss = warp_sum(loss);
        dot = warp_sum(dot);
        if (dlogits) {
            // dL/dp_k = (g_k - dot)/sum ; softmax backward: dz_j = p_j*(dp_j - sum_k dp_k p_k)
            float dot2 = 0.f;
            for (int c = lane; c < classes; c += 32) {
                float q = p[c] / sum;
                float cl = fminf(fmaxf(q, kEps), 1.f - kEps);
                float g = (q >= kEps && q <= 1.f - kEps) ? -t[c] / cl : 0.f;
                dot2 = fmaf((g - dot) / sum, p[c], dot2);
            }
            dot2 = warp_sum(dot2);
            for (int c = lane; c < classes; c += 32) {
                float q = p[c] / sum;
                float cl = fminf(fmaxf(q, kEps), 1.f - kEps);
                float g = (q >= kEps && q <= 1.f - kEps) ? -t[c] / cl : 0.f;
                dlogits[(int64_t)row * classes + c] = scale * p[c] * ((g - dot) / sum - dot2);
            }
        }
        if (lane == 0) {
            my_loss += (double)loss;
            my_correct += (ip == it) ? 1 : 0;
        }
    }
    if (lane == 0) { block_loss[wid] = my_loss; block_correct[wid] = my_correct; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double l = 0.0;
        int k = 0;
        for (int i = 0; i < 8; ++i) { l += block_loss[i]; k += block_correct[i]; }
        if (loss_sum) atomicAdd(loss_sum, l);
        if (correct && k) atomicAdd(correct, (unsigned long long)k);
    }
}

}  // namespace fnet

using namespace fnet;

extern "C" int fnet_softmax_cce(int32_t batch, int32_t classes, const float *probs, const float *targets, float scale,
                                float *dlogits, double *loss_sum, int64_t *correct, void *stream) {
    FNET_REQUIRE(batch >= 0 && classes > 0 && probs && targets, "fnet_softmax_cce: bad arguments");
    if (batch == 0) return FNET_OK;
    // ONE CTA: the double atomicAdd then happens once per launch, so the loss sum does not depend on
    // the order in which CTAs retire (batches are <= a few thousand rows of <= 100 classes)
    softmax_cce_kernel<<<1, 256, 0, as_stream(stream)>>>(batch, classes, probs, targets, scale, dlogits, loss_sum,
                                                         reinterpret_cast<unsigned long long *>(correct));
    return check_launch("fnet_softmax_cce");
}

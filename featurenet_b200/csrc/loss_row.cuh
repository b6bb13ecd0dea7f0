// One row of softmax-cross-entropy (shared by loss.cu and the fused step kernel).
#pragma once
#include "common.cuh"

namespace fnet {

constexpr float kClipEps = 1e-7f;

// Everything for one row whose values are addressable through P (registers or memory):
// returns the loss, sets `hit`, optionally writes scale * dL/dlogits.
template <class P, class T>
__device__ __forceinline__ float row_loss(int classes, const P &p, const T &t, float scale, float *dz, bool &hit) {
    float sum = 0.f, bp = -INFINITY, bt = -INFINITY;
    int ip = 0, it = 0;
    for (int c = 0; c < classes; ++c) {           // argmax: first index on ties (strict > while scanning upwards)
        float pv = p(c), tv = t(c);
        sum += pv;
        if (pv > bp) { bp = pv; ip = c; }
        if (tv > bt) { bt = tv; it = c; }
    }
    hit = ip == it;
    float loss = 0.f, dot = 0.f;
    for (int c = 0; c < classes; ++c) {           // loss, and g = dL/dq masked by the clip; dot = sum g*q
        float q = p(c) / sum;
        float cl = fminf(fmaxf(q, kClipEps), 1.f - kClipEps);
        float tv = t(c);
        loss -= tv * logf(cl);
        float g = (q >= kClipEps && q <= 1.f - kClipEps) ? -tv / cl : 0.f;
        dot = fmaf(g, q, dot);
    }
    if (dz) {
        // dL/dp_k = (g_k - dot)/sum ; softmax backward: dz_j = p_j*(dp_j - sum_k dp_k p_k)
        float dot2 = 0.f;
        for (int c = 0; c < classes; ++c) {
            float pv = p(c), q = pv / sum;
            float cl = fminf(fmaxf(q, kClipEps), 1.f - kClipEps);
            float g = (q >= kClipEps && q <= 1.f - kClipEps) ? -t(c) / cl : 0.f;
            dot2 = fmaf((g - dot) / sum, pv, dot2);
        }
        for (int c = 0; c < classes; ++c) {
            float pv = p(c), q = pv / sum;
            float cl = fminf(fmaxf(q, kClipEps), 1.f - kClipEps);
            float g = (q >= kClipEps && q <= 1.f - kClipEps) ? -t(c) / cl : 0.f;
            dz[c] = scale * pv * ((g - dot) / sum - dot2);
        }
    }
    return loss;
}

struct MemRow {
    const float *q;
    __device__ float operator()(int c) const { return q[c]; }
};

}  // namespace fnet

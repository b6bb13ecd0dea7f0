// One step of an L2 gradient attack (FGSM with a growing step, or one PGD iteration): the device half of the
// robustness fitness (tensorflow_generator.py:151-218).  One CTA per input row: two block reductions (the
// gradient's norm, then the norm of the new perturbation for the projection), both in a fixed order.
#include "common.cuh"

namespace fnet {

__global__ void __launch_bounds__(256) l2_attack_step_kernel(int64_t row, const float *__restrict__ x0,
                                                             const float *__restrict__ grad, const float *adv_in,
                                                             float *adv_out, float step, float radius, float clip_lo,
                                                             float clip_hi, const uint8_t *__restrict__ active) {
    pdl_prologue();
    __shared__ float red[32];
    const int64_t base = (int64_t)blockIdx.x * row;
    if (active && !active[blockIdx.x]) return;      // finished rows keep the adversarial example they already have
    float s = 0.f;
    for (int64_t i = threadIdx.x; i < row; i += blockDim.x) { float g = __ldg(grad + base + i); s = fmaf(g, g, s); }
    const float gscale = step / (sqrtf(block_sum(s, red)) + 1e-7f);
    float d2 = 0.f;
    for (int64_t i = threadIdx.x; i < row; i += blockDim.x) {
        float d = (adv_in[base + i] - __ldg(x0 + base + i)) + gscale * __ldg(grad + base + i);
        d2 = fmaf(d, d, d2);
    }
    const float dnorm = sqrtf(block_sum(d2, red));
    const float shrink = (radius > 0.f && dnorm > radius) ? radius / (dnorm + 1e-7f) : 1.f;
    for (int64_t i = threadIdx.x; i < row; i += blockDim.x) {
        float x = __ldg(x0 + base + i);
        float d = ((adv_in[base + i] - x) + gscale * __ldg(grad + base + i)) * shrink;
        adv_out[base + i] = fminf(fmaxf(x + d, clip_lo), clip_hi);
    }
}

}  // namespace fnet

using namespace fnet;

extern "C" int fnet_l2_attack_step(int32_t batch, int64_t row, const float *x0, const float *grad, const float *adv_in,
                                   float *adv_out, float step, float radius, float clip_lo, float clip_hi,
                                   const uint8_t *active, void *stream) {
    FNET_REQUIRE(batch >= 0 && row > 0 && x0 && grad && adv_in && adv_out, "fnet_l2_attack_step: bad arguments");
    FNET_REQUIRE(clip_lo <= clip_hi, "fnet_l2_attack_step: empty clip range");
    if (batch == 0) return FNET_OK;
    launch_pdl(l2_attack_step_kernel, dim3(batch), dim3(256), 0, as_stream(stream), row, x0, grad, adv_in, adv_out, step, radius,
               clip_lo, clip_hi, active);
    return check_launch("fnet_l2_attack_step");
}

// Optimizer updates as one launch over the flat parameter arena (Keras runs ~10 small TF ops
// per weight tensor).  Adam: 4 reads (p, g, m, v) + 3 writes per parameter = 28 B/param, the
// HBM roofline of this kernel; lr_t comes from the device-side loop state so the launch is
// identical every step and can live in a CUDA graph.  With advance >= 0 the same launch also
// closes the step (iteration += 1, epoch cursor += advance): only lr / lr_t are read by the
// other CTAs, so one thread may update the two counters without a race.
#include "common.cuh"

namespace fnet {

__device__ __forceinline__ void close_step(int32_t *state, int32_t advance) {
    if (advance >= 0 && blockIdx.x == 0 && threadIdx.x == 0) {
        state[0] += 1;
        state[1] += advance;
    }
}

__global__ void __launch_bounds__(256) adam_kernel(int64_t n, float *__restrict__ p, const float *__restrict__ g,
                                                   float *__restrict__ m, float *__restrict__ v, int32_t *__restrict__ state,
                                                   float b1, float b2, float eps, int32_t advance) {
    pdl_prologue();
    const float lr_t = __int_as_float(state[3]);
    const float c1 = 1.f - b1, c2 = 1.f - b2;
    int64_t n4 = n >> 2;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    float4 *p4 = reinterpret_cast<float4 *>(p), *m4 = reinterpret_cast<float4 *>(m), *v4 = reinterpret_cast<float4 *>(v);
    const float4 *g4 = reinterpret_cast<const float4 *>(g);
#define FNET_ADAM_ONE(P, G, M, V)                  \
    {                                              \
        float mm = b1 * (M) + c1 * (G);            \
        float vv = b2 * (V) + c2 * ((G) * (G));    \
        (M) = mm;                                  \
        (V) = vv;                                  \
        (P) = (P) - lr_t * mm / (sqrtf(vv) + eps); \
    }
    for (int64_t k = i; k < n4; k += stride) {
        float4 pp = p4[k], gg = __ldg(g4 + k), mm4 = m4[k], vv4 = v4[k];
        FNET_ADAM_ONE(pp.x, gg.x, mm4.x, vv4.x)
        FNET_ADAM_ONE(pp.y, gg.y, mm4.y, vv4.y)
        FNET_ADAM_ONE(pp.z, gg.z, mm4.z, vv4.z)
        FNET_ADAM_ONE(pp.w, gg.w, mm4.w, vv4.w)
        p4[k] = pp; m4[k] = mm4; v4[k] = vv4;
    }
    for (int64_t e = (n4 << 2) + i; e < n; e += stride) {
        float pp = p[e], gg = g[e], mm1 = m[e], vv1 = v[e];
        FNET_ADAM_ONE(pp, gg, mm1, vv1)
        p[e] = pp; m[e] = mm1; v[e] = vv1;
    }
#undef FNET_ADAM_ONE
    close_step(state, advance);
}

// Keras SGD: vel = momentum*vel - lr*g ; p += vel  (nesterov: p += momentum*vel - lr*g)
__global__ void __launch_bounds__(256) sgd_kernel(int64_t n, float *__restrict__ p, const float *__restrict__ g,
                                                  float *__restrict__ vel, int32_t *__restrict__ state, float momentum,
                                                  int nesterov, int32_t advance) {
    pdl_prologue();
    const float lr = __int_as_float(state[2]);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float gg = __ldg(g + i);
        float vv = momentum * vel[i] - lr * gg;
        vel[i] = vv;
        p[i] = nesterov ? p[i] + momentum * vv - lr * gg : p[i] + vv;
    }
    close_step(state, advance);
}

}  // namespace fnet

using namespace fnet;

extern "C" {

int fnet_adam_step(int64_t n, float *params, const float *grads, float *m, float *v, int32_t *state, float beta1,
                   float beta2, float eps, int32_t advance, void *stream) {
    FNET_REQUIRE(n >= 0 && params && grads && m && v && state, "fnet_adam_step: bad arguments");
    FNET_REQUIRE(((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) |
                   reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0,
                 "fnet_adam_step: arenas must be 16-byte aligned");
    if (n == 0) return advance >= 0 ? fnet_step_end(state, advance, stream) : FNET_OK;
    launch_pdl(adam_kernel, dim3(stream_grid(ceil_div<int64_t>(n, 4), 256)), dim3(256), 0, as_stream(stream), n, params, grads, m, v, state,
                                                                                       beta1, beta2, eps, advance);
    return check_launch("fnet_adam_step");
}

int fnet_sgd_step(int64_t n, float *params, const float *grads, float *velocity, int32_t *state, float momentum,
                  int nesterov, int32_t advance, void *stream) {
    FNET_REQUIRE(n >= 0 && params && grads && velocity && state, "fnet_sgd_step: bad arguments");
    if (n == 0) return advance >= 0 ? fnet_step_end(state, advance, stream) : FNET_OK;
    launch_pdl(sgd_kernel, dim3(stream_grid(n, 256)), dim3(256), 0, as_stream(stream), n, params, grads, velocity, state, momentum, nesterov,
                                                                   advance);
    return check_launch("fnet_sgd_step");
}

}  // extern "C"

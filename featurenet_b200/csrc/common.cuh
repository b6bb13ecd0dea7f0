// Shared helpers for libfnet's kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include "../../include/fnet.h"

namespace fnet {

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs; grids are sized in multiples of this

void set_error(const char *fmt, ...);

inline int check_launch(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return FNET_ERR_CUDA;
    }
    return FNET_OK;
}

#define FNET_REQUIRE(cond, ...)                 \
    do {                                        \
        if (!(cond)) {                          \
            fnet::set_error(__VA_ARGS__);       \
            return FNET_ERR_ARG;                \
        }                                       \
    } while (0)

#define FNET_CUDA(call)                                                        \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) {                                               \
            fnet::set_error("%s: %s", #call, cudaGetErrorString(e_));          \
            return FNET_ERR_CUDA;                                              \
        }                                                                      \
    } while (0)

inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

template <typename T>
__host__ __device__ constexpr T ceil_div(T a, T b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------- programmatic dependent launch
// A FeatureNet step is a chain of small kernels, so the gap between one kernel's last CTA and the next
// kernel's first instruction is a large part of the step.  Every kernel is launched with the programmatic
// stream-serialisation attribute and opens with pdl_prologue(): `launch_dependents` lets the NEXT kernel's CTAs
// be scheduled as soon as all CTAs of this one have started, and `wait` holds this kernel until its predecessor
// has completed and flushed its writes.  The wait is the first thing a kernel does, so nothing is read or
// written early: the ordering semantics are exactly those of a plain stream, only the launch latency overlaps.
// Captured into a CUDA graph the attribute becomes a programmatic edge.  FNET_PDL=0 turns it off.
__device__ __forceinline__ void pdl_prologue() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// An on/off switch read from the environment.  Callers keep the result in a function-local `static const`, whose
// initialisation the language makes thread-safe: libfnet is driven by several host threads at once (one per resident
// candidate, featurenet_b200/population.py), and ctypes releases the GIL around every call.
inline int env_switch(const char *name, int fallback) {
    const char *e = getenv(name);
    if (!e || !e[0]) return fallback;
    return e[0] == '0' ? 0 : 1;
}

inline bool pdl_enabled() {
    static const int mode = env_switch("FNET_PDL", 1);
    return mode == 1;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to the (kernel, device) pair.  `done` is one bit per device
// ordinal, one mask per kernel instantiation; setting the attribute twice from two racing threads is harmless.
template <typename Kernel>
inline cudaError_t allow_dynamic_smem(Kernel kernel, int bytes, std::atomic<uint64_t> &done) {
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return e;
    const uint64_t bit = 1ull << (device & 63);
    if (done.load(std::memory_order_acquire) & bit) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) done.fetch_or(bit, std::memory_order_release);
    return e;
}

template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    (void)cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);     // errors surface in check_launch()
}

// grid for a grid-stride element-wise kernel: enough CTAs to cover n, capped at 8 per SM
inline int stream_grid(int64_t work_items, int threads) {
    int64_t blocks = ceil_div<int64_t>(work_items, threads);
    int64_t cap = (int64_t)kNumSMs * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// i -> (b, h, w, c) of a dense [B][d1][d2][d3] tensor.  32-bit arithmetic whenever the index allows: a 64-bit
// division is a long software routine, and the streaming kernels do three per element to decode their index.
__device__ __forceinline__ void decode4(int64_t i, int d1, int d2, int d3, int64_t &b, int &h, int &w, int &c) {
    if (i < 0x7fffffffLL) {
        unsigned u = (unsigned)i, t = u / (unsigned)d3;
        c = (int)(u - t * (unsigned)d3);
        unsigned t2 = t / (unsigned)d2;
        w = (int)(t - t2 * (unsigned)d2);
        unsigned t3 = t2 / (unsigned)d1;
        h = (int)(t2 - t3 * (unsigned)d1);
        b = t3;
    } else {
        c = (int)(i % d3);
        int64_t t = i / d3;
        w = (int)(t % d2);
        t /= d2;
        h = (int)(t % d1);
        b = t / d1;
    }
}

// m -> (b, h, w) of [B][d1][d2] pixels, 32-bit whenever possible (see decode4)
__device__ __forceinline__ void decode3(long long m, int d1, int d2, long long &b, int &h, int &w) {
    if (m < 0x7fffffffLL) {
        unsigned u = (unsigned)m, t = u / (unsigned)d2;
        w = (int)(u - t * (unsigned)d2);
        unsigned bb = t / (unsigned)d1;
        h = (int)(t - bb * (unsigned)d1);
        b = bb;
    } else {
        w = (int)(m % d2);
        long long t = m / d2;
        h = (int)(t % d1);
        b = t / d1;
    }
}

__device__ __forceinline__ float act_apply(float z, int act) {
    switch (act) {
        case FNET_ACT_RELU: return fmaxf(z, 0.f);
        case FNET_ACT_TANH: return tanhf(z);
        case FNET_ACT_SIGMOID: return 1.f / (1.f + expf(-z));
        default: return z;
    }
}

// derivative factor from the activation OUTPUT y
__device__ __forceinline__ float act_grad(float y, int act) {
    switch (act) {
        case FNET_ACT_RELU: return y > 0.f ? 1.f : 0.f;
        case FNET_ACT_TANH: return 1.f - y * y;
        case FNET_ACT_SIGMOID: return y * (1.f - y);
        default: return 1.f;
    }
}

// epilogue of every data gradient: dx = (accumulate ? dx + v : v) * act'(y)   (see DxEpilogue in igemm.cuh)
__device__ __forceinline__ void store_dx(float *dx, long long at, float v, int accumulate, int act, const float *y) {
    if (accumulate) v += dx[at];
    if (act) v *= act_grad(__ldg(y + at), act);
    dx[at] = v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// block-wide sum for blockDim.x a multiple of 32 (<= 1024); result valid in every thread
__device__ __forceinline__ float block_sum(float v, float *smem32) {
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) smem32[wid] = v;
    __syncthreads();
    float r = (lane < nw) ? smem32[lane] : 0.f;
    r = warp_sum(r);
    return r;
}

}  // namespace fnet

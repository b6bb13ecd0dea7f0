// libfnet: error reporting, CUDA-graph helpers, loop-state kernels and batch gathering.
#include "common.cuh"

namespace fnet {

static thread_local char g_error[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

// ---------------------------------------------------------------- loop state
// state[0] t (completed iterations)  state[1] cursor  state[2] lr  state[3] lr_t
__global__ void step_begin_kernel(int32_t *state, float beta1, float beta2) {
    pdl_prologue();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double t = (double)(state[0] + 1);
        double lr = (double)__int_as_float(state[2]);
        // Keras Adam: lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t); evaluated once per step
        // (in double, rounded to fp32) instead of once per parameter tensor
        double lr_t = lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t));
        state[3] = __float_as_int((float)lr_t);
    }
}

__global__ void step_end_kernel(int32_t *state, int32_t batch) {
    pdl_prologue();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        state[0] += 1;
        state[1] += batch;
    }
}

__global__ void state_set_lr_kernel(int32_t *state, float lr) {
    pdl_prologue();
    if (threadIdx.x == 0 && blockIdx.x == 0) state[2] = __float_as_int(lr);
}

__global__ void fill_kernel(int64_t n, float value, float *__restrict__ dst) {
    pdl_prologue();
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) dst[i] = value;
}

__global__ void u8_to_f32_kernel(int64_t n, const uint8_t *__restrict__ src, float scale, float *__restrict__ dst) {
    pdl_prologue();
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // same rounding as numpy: x.astype(float32) then an fp32 divide by 255 (scale is the divisor)
    for (; i < n; i += stride) dst[i] = (float)src[i] / scale;
}

// One CTA per (row, chunk); rows are contiguous so float4 copies are coalesced.
__global__ void gather_rows_kernel(int32_t batch, int64_t row_elems, const int32_t *__restrict__ perm,
                                   const int32_t *__restrict__ state, const float *__restrict__ src,
                                   float *__restrict__ dst) {
    pdl_prologue();
    int32_t cursor = state[1];
    for (int32_t row = blockIdx.y; row < batch; row += gridDim.y) {
        int64_t from = (int64_t)perm[cursor + row] * row_elems;
        int64_t to = (int64_t)row * row_elems;
        if ((row_elems & 3) == 0) {
            const float4 *s4 = reinterpret_cast<const float4 *>(src + from);
            float4 *d4 = reinterpret_cast<float4 *>(dst + to);
            for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < row_elems / 4;
                 i += (int64_t)gridDim.x * blockDim.x)
                d4[i] = __ldg(s4 + i);
        } else {
            for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < row_elems;
                 i += (int64_t)gridDim.x * blockDim.x)
                dst[to + i] = __ldg(src + from + i);
        }
    }
}

// Inputs and targets of one batch in ONE launch (+ the per-step lr_t of fnet_step_begin when asked):
// grid = (chunks of the x row, batch); the first chunk of every row also copies the target row.
__global__ void gather_batch_kernel(int32_t batch, int64_t x_row, int64_t y_row, const int32_t *__restrict__ perm,
                                    int32_t *__restrict__ state, const float *__restrict__ x_src,
                                    const float *__restrict__ y_src, float *__restrict__ x_dst, float *__restrict__ y_dst,
                                    int begin_step, float beta1, float beta2) {
    pdl_prologue();
    const int32_t cursor = state[1];
    if (begin_step && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
        double t = (double)(state[0] + 1);
        double lr = (double)__int_as_float(state[2]);
        state[3] = __float_as_int((float)(lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t))));
    }
    for (int32_t row = blockIdx.y; row < batch; row += gridDim.y) {
        const int64_t from = (int64_t)perm[cursor + row];
        const float *xs = x_src + from * x_row;
        float *xd = x_dst + (int64_t)row * x_row;
        if ((x_row & 3) == 0) {
            const float4 *s4 = reinterpret_cast<const float4 *>(xs);
            float4 *d4 = reinterpret_cast<float4 *>(xd);
            for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < x_row / 4; i += (int64_t)gridDim.x * blockDim.x)
                d4[i] = __ldg(s4 + i);
        } else {
            for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < x_row; i += (int64_t)gridDim.x * blockDim.x)
                xd[i] = __ldg(xs + i);
        }
        if (blockIdx.x == 0)
            for (int64_t i = threadIdx.x; i < y_row; i += blockDim.x) y_dst[(int64_t)row * y_row + i] = __ldg(y_src + from * y_row + i);
    }
}

// A batch that arrived from the host as uint8 pixels: x_dst = float(x_src) / scale (numpy's rounding, see
// u8_to_f32_kernel), targets copied as they are, lr_t of the step computed by the first thread -- the first launch
// of a streamed train step.  16 pixels per thread per pass (one 16-byte load, four 16-byte stores).
__global__ void feed_batch_u8_kernel(int64_t x_elems, int64_t y_elems, int32_t *__restrict__ state,
                                     const uint8_t *__restrict__ x_src, float scale, const float *__restrict__ y_src,
                                     float *__restrict__ x_dst, float *__restrict__ y_dst, int begin_step, float beta1,
                                     float beta2) {
    pdl_prologue();
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
    if (begin_step && tid == 0) {
        double t = (double)(state[0] + 1);
        double lr = (double)__int_as_float(state[2]);
        state[3] = __float_as_int((float)(lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t))));
    }
    const int64_t vec = x_elems >> 4;
    const uint4 *s16 = reinterpret_cast<const uint4 *>(x_src);
    float4 *d4 = reinterpret_cast<float4 *>(x_dst);
    for (int64_t i = tid; i < vec; i += stride) {
        const uint4 q = __ldg(s16 + i);
        const uint32_t words[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t wd = words[k];
            d4[i * 4 + k] = make_float4((float)(wd & 255u) / scale, (float)((wd >> 8) & 255u) / scale,
                                        (float)((wd >> 16) & 255u) / scale, (float)(wd >> 24) / scale);
        }
    }
    for (int64_t i = (vec << 4) + tid; i < x_elems; i += stride) x_dst[i] = (float)x_src[i] / scale;
    for (int64_t i = tid; i < y_elems; i += stride) y_dst[i] = __ldg(y_src + i);
}

}  // namespace fnet

using namespace fnet;

extern "C" {

int fnet_gather_batch(int32_t batch, int64_t x_row, int64_t y_row, const int32_t *perm, int32_t *state,
                      const float *x_src, const float *y_src, float *x_dst, float *y_dst, int begin_step, float beta1,
                      float beta2, void *stream) {
    FNET_REQUIRE(batch >= 0 && x_row > 0 && y_row > 0, "fnet_gather_batch: bad sizes");
    FNET_REQUIRE(perm && state && x_src && y_src && x_dst && y_dst, "fnet_gather_batch: null pointer");
    FNET_REQUIRE((x_row & 3) != 0 || ((reinterpret_cast<uintptr_t>(x_src) | reinterpret_cast<uintptr_t>(x_dst)) & 15) == 0,
                 "fnet_gather_batch: 16-byte aligned buffers needed for rows that are a multiple of 4 floats");
    if (batch == 0) return begin_step ? fnet_step_begin(state, beta1, beta2, stream) : FNET_OK;
    int64_t vec = (x_row & 3) == 0 ? x_row / 4 : x_row;
    int bx = (int)ceil_div<int64_t>(vec, 256);
    if (bx > 8) bx = 8;
    dim3 grid(bx, batch > 4096 ? 4096 : batch);
    launch_pdl(gather_batch_kernel, dim3(grid), dim3(256), 0, as_stream(stream), batch, x_row, y_row, perm, state, x_src, y_src, x_dst, y_dst,
                                                             begin_step, beta1, beta2);
    return check_launch("fnet_gather_batch");
}

int fnet_version(void) { return FNET_VERSION; }

const char *fnet_last_error(void) { return g_error; }

int fnet_step_begin(int32_t *state, float beta1, float beta2, void *stream) {
    FNET_REQUIRE(state != nullptr, "fnet_step_begin: null state");
    launch_pdl(step_begin_kernel, dim3(1), dim3(32), 0, as_stream(stream), state, beta1, beta2);
    return check_launch("fnet_step_begin");
}

int fnet_step_end(int32_t *state, int32_t batch, void *stream) {
    FNET_REQUIRE(state != nullptr, "fnet_step_end: null state");
    launch_pdl(step_end_kernel, dim3(1), dim3(32), 0, as_stream(stream), state, batch);
    return check_launch("fnet_step_end");
}

int fnet_state_set_lr(int32_t *state, float lr, void *stream) {
    FNET_REQUIRE(state != nullptr, "fnet_state_set_lr: null state");
    launch_pdl(state_set_lr_kernel, dim3(1), dim3(32), 0, as_stream(stream), state, lr);
    return check_launch("fnet_state_set_lr");
}

int fnet_feed_batch_u8(int32_t batch, int64_t x_row, int64_t y_row, int32_t *state, const uint8_t *x_src, float scale,
                       const float *y_src, float *x_dst, float *y_dst, int begin_step, float beta1, float beta2,
                       void *stream) {
    FNET_REQUIRE(batch >= 0 && x_row > 0 && y_row > 0 && scale != 0.f, "fnet_feed_batch_u8: bad sizes");
    FNET_REQUIRE(state && x_src && y_src && x_dst && y_dst, "fnet_feed_batch_u8: null pointer");
    FNET_REQUIRE(((reinterpret_cast<uintptr_t>(x_src) | reinterpret_cast<uintptr_t>(x_dst)) & 15) == 0,
                 "fnet_feed_batch_u8: 16-byte aligned buffers needed");
    if (batch == 0) return begin_step ? fnet_step_begin(state, beta1, beta2, stream) : FNET_OK;
    const int64_t x_elems = (int64_t)batch * x_row, y_elems = (int64_t)batch * y_row;
    launch_pdl(feed_batch_u8_kernel, dim3(stream_grid(ceil_div<int64_t>(x_elems, 16), 256)), dim3(256), 0, as_stream(stream),
               x_elems, y_elems, state, x_src, scale, y_src, x_dst, y_dst, begin_step, beta1, beta2);
    return check_launch("fnet_feed_batch_u8");
}

int fnet_fill(int64_t n, float value, float *dst, void *stream) {
    FNET_REQUIRE(n >= 0 && (dst != nullptr || n == 0), "fnet_fill: bad arguments");
    if (n == 0) return FNET_OK;
    launch_pdl(fill_kernel, dim3(stream_grid(n, 256)), dim3(256), 0, as_stream(stream), n, value, dst);
    return check_launch("fnet_fill");
}

int fnet_u8_to_f32(int64_t n, const uint8_t *src, float scale, float *dst, void *stream) {
    FNET_REQUIRE(n >= 0 && scale != 0.f, "fnet_u8_to_f32: bad arguments");
    if (n == 0) return FNET_OK;
    launch_pdl(u8_to_f32_kernel, dim3(stream_grid(n, 256)), dim3(256), 0, as_stream(stream), n, src, scale, dst);
    return check_launch("fnet_u8_to_f32");
}

int fnet_gather_rows(int32_t batch, int64_t row_elems, const int32_t *perm, const int32_t *state,
                     const float *src, float *dst, void *stream) {
    FNET_REQUIRE(batch >= 0 && row_elems > 0, "fnet_gather_rows: bad sizes");
    FNET_REQUIRE(perm && state && src && dst, "fnet_gather_rows: null pointer");
    if (batch == 0) return FNET_OK;
    int64_t vec = (row_elems & 3) == 0 ? row_elems / 4 : row_elems;
    int bx = (int)ceil_div<int64_t>(vec, 256);
    if (bx > 8) bx = 8;
    dim3 grid(bx, batch > 4096 ? 4096 : batch);
    launch_pdl(gather_rows_kernel, dim3(grid), dim3(256), 0, as_stream(stream), batch, row_elems, perm, state, src, dst);
    return check_launch("fnet_gather_rows");
}

// ---------------------------------------------------------------- streams
int fnet_stream_create(void **stream_out) {
    FNET_REQUIRE(stream_out != nullptr, "fnet_stream_create: null output");
    cudaStream_t s = nullptr;
    FNET_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream_out = s;
    return FNET_OK;
}

int fnet_stream_destroy(void *stream) {
    if (stream) FNET_CUDA(cudaStreamDestroy(as_stream(stream)));
    return FNET_OK;
}

// ---------------------------------------------------------------- CUDA graphs
int fnet_graph_begin(void *stream) {
    FNET_CUDA(cudaStreamBeginCapture(as_stream(stream), cudaStreamCaptureModeThreadLocal));
    return FNET_OK;
}

int fnet_graph_end(void *stream, void **graph_exec_out) {
    FNET_REQUIRE(graph_exec_out != nullptr, "fnet_graph_end: null output");
    cudaGraph_t graph = nullptr;
    FNET_CUDA(cudaStreamEndCapture(as_stream(stream), &graph));
    cudaGraphExec_t exec = nullptr;
    cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) {
        set_error("cudaGraphInstantiate: %s", cudaGetErrorString(e));
        return FNET_ERR_CUDA;
    }
    *graph_exec_out = exec;
    return FNET_OK;
}

int fnet_graph_launch(void *graph_exec, void *stream) {
    FNET_REQUIRE(graph_exec != nullptr, "fnet_graph_launch: null graph");
    FNET_CUDA(cudaGraphLaunch(reinterpret_cast<cudaGraphExec_t>(graph_exec), as_stream(stream)));
    return FNET_OK;
}

int fnet_graph_launch_n(void *graph_exec, int32_t n, void *stream) {
    FNET_REQUIRE(graph_exec != nullptr && n >= 0, "fnet_graph_launch_n: null graph or negative count");
    cudaGraphExec_t exec = reinterpret_cast<cudaGraphExec_t>(graph_exec);
    cudaStream_t s = as_stream(stream);
    for (int32_t i = 0; i < n; ++i) FNET_CUDA(cudaGraphLaunch(exec, s));
    return FNET_OK;
}

int fnet_graph_step_fed(void *graph_exec, void *x_dev, const void *x_host, size_t x_bytes, void *y_dev,
                        const void *y_host, size_t y_bytes, const void *result_dev, void *result_host,
                        size_t result_bytes, void *stream) {
    FNET_REQUIRE(graph_exec != nullptr, "fnet_graph_step_fed: null graph");
    FNET_REQUIRE((x_host == nullptr || x_dev != nullptr) && (y_host == nullptr || y_dev != nullptr),
                 "fnet_graph_step_fed: host source without a device destination");
    FNET_REQUIRE(result_bytes == 0 || (result_dev != nullptr && result_host != nullptr), "fnet_graph_step_fed: null result buffer");
    cudaStream_t s = as_stream(stream);
    if (x_host && x_bytes) FNET_CUDA(cudaMemcpyAsync(x_dev, x_host, x_bytes, cudaMemcpyHostToDevice, s));
    if (y_host && y_bytes) FNET_CUDA(cudaMemcpyAsync(y_dev, y_host, y_bytes, cudaMemcpyHostToDevice, s));
    FNET_CUDA(cudaGraphLaunch(reinterpret_cast<cudaGraphExec_t>(graph_exec), s));
    if (result_bytes) FNET_CUDA(cudaMemcpyAsync(result_host, result_dev, result_bytes, cudaMemcpyDeviceToHost, s));
    FNET_CUDA(cudaStreamSynchronize(s));
    return FNET_OK;
}

int fnet_graph_step_streamed(void *graph_exec, void *x_slot, const void *x_host, void *x_stage, size_t x_bytes,
                             void *y_slot, const void *y_host, void *y_stage, size_t y_bytes, const void *result_dev,
                             void *result_host, void *result_prev, size_t result_bytes, void *copy_stream,
                             void *compute_stream, void *slot_event, void *copied_event) {
    FNET_REQUIRE(graph_exec != nullptr, "fnet_graph_step_streamed: null graph");
    FNET_REQUIRE(x_slot && x_host && y_slot && y_host, "fnet_graph_step_streamed: null batch buffer");
    FNET_REQUIRE(slot_event && copied_event, "fnet_graph_step_streamed: null event");
    FNET_REQUIRE(result_bytes == 0 || (result_dev != nullptr && result_host != nullptr),
                 "fnet_graph_step_streamed: null result buffer");
    FNET_REQUIRE(copy_stream != compute_stream, "fnet_graph_step_streamed: the copies need a stream of their own");
    cudaStream_t copy = as_stream(copy_stream), compute = as_stream(compute_stream);
    cudaEvent_t slot = reinterpret_cast<cudaEvent_t>(slot_event), copied = reinterpret_cast<cudaEvent_t>(copied_event);
    FNET_CUDA(cudaEventSynchronize(slot));                    // the slot's previous step is over: everything it used is free
    if (result_prev && result_bytes) memcpy(result_prev, result_host, result_bytes);
    if (x_stage) { memcpy(x_stage, x_host, x_bytes); x_host = x_stage; }
    if (y_stage) { memcpy(y_stage, y_host, y_bytes); y_host = y_stage; }
    FNET_CUDA(cudaMemcpyAsync(x_slot, x_host, x_bytes, cudaMemcpyHostToDevice, copy));
    FNET_CUDA(cudaMemcpyAsync(y_slot, y_host, y_bytes, cudaMemcpyHostToDevice, copy));
    FNET_CUDA(cudaEventRecord(copied, copy));
    FNET_CUDA(cudaStreamWaitEvent(compute, copied, 0));
    FNET_CUDA(cudaGraphLaunch(reinterpret_cast<cudaGraphExec_t>(graph_exec), compute));
    if (result_bytes) FNET_CUDA(cudaMemcpyAsync(result_host, result_dev, result_bytes, cudaMemcpyDeviceToHost, compute));
    FNET_CUDA(cudaEventRecord(slot, compute));
    return FNET_OK;
}

int fnet_graph_destroy(void *graph_exec) {
    if (graph_exec) FNET_CUDA(cudaGraphExecDestroy(reinterpret_cast<cudaGraphExec_t>(graph_exec)));
    return FNET_OK;
}

}  // extern "C"

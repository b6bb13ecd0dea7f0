// How many kernel nodes per second does one B200 retire when S streams each replay a straight-line CUDA graph?
// Answers whether a co-resident population (about a million small launches per pass) is bound by the launch front end
// or by the SMs.  Build and run:  nvcc -arch=sm_100a -O2 -o /tmp/launch_rate scripts/launch_rate.cu && /tmp/launch_rate
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <vector>

__global__ void spin(float *p, int iters) {
    float v = p[threadIdx.x & 31];
    for (int i = 0; i < iters; ++i) v = v * 1.0001f + 0.5f;
    if (v == 123.456f) p[0] = v;
}

struct Fat { float *p; int iters; char pad[480]; };          // ~0.5 KB of kernel parameters, like a descriptor-carrying launch
__global__ void spin_fat(Fat f) {
    float v = f.p[threadIdx.x & 31];
    for (int i = 0; i < f.iters; ++i) v = v * 1.0001f + 0.5f;
    if (v == 123.456f) f.p[0] = v;
}

static int g_pdl = 0, g_fat = 0;

static void launch(cudaStream_t st, float *buf, int blocks, int threads, int iters) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(threads); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = g_pdl ? 1 : 0;
    if (g_fat) { Fat f; f.p = buf; f.iters = iters; cudaLaunchKernelEx(&cfg, spin_fat, f); }
    else cudaLaunchKernelEx(&cfg, spin, buf, iters);
}

static double run(int streams, int nodes, int blocks, int threads, int iters, int replays) {
    float *buf;
    cudaMalloc(&buf, 4096);
    cudaMemset(buf, 0, 4096);
    std::vector<cudaStream_t> st(streams);
    std::vector<cudaGraphExec_t> ex(streams);
    for (int s = 0; s < streams; ++s) {
        cudaStreamCreateWithFlags(&st[s], cudaStreamNonBlocking);
        cudaGraph_t g;
        cudaStreamBeginCapture(st[s], cudaStreamCaptureModeThreadLocal);
        for (int n = 0; n < nodes; ++n) launch(st[s], buf, blocks, threads, iters);
        cudaStreamEndCapture(st[s], &g);
        cudaGraphInstantiate(&ex[s], g, 0);
        cudaGraphDestroy(g);
    }
    for (int s = 0; s < streams; ++s) cudaGraphLaunch(ex[s], st[s]);
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < replays; ++r)
        for (int s = 0; s < streams; ++s) cudaGraphLaunch(ex[s], st[s]);
    cudaDeviceSynchronize();
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (int s = 0; s < streams; ++s) { cudaGraphExecDestroy(ex[s]); cudaStreamDestroy(st[s]); }
    cudaFree(buf);
    return sec / ((double)replays * streams * nodes) * 1e6;      // us per kernel node, whole GPU
}

int main() {
    printf("%8s %6s %4s %4s %7s %7s %7s | us per kernel node over the whole GPU\n", "streams", "nodes", "pdl", "fat", "blocks", "threads", "iters");
    const int shapes[][3] = {{1, 32, 0}, {148, 256, 0}, {148, 256, 2000}, {37, 256, 2000}};
    for (int fat = 0; fat < 2; ++fat)
        for (int pdl = 0; pdl < 2; ++pdl)
            for (int nodes : {80, 400})
                for (auto &sh : shapes)
                    for (int s : {1, 8}) {
                        g_pdl = pdl; g_fat = fat;
                        double us = run(s, nodes, sh[0], sh[1], sh[2], 8000 / nodes);
                        printf("%8d %6d %4d %4d %7d %7d %7d | %.3f\n", s, nodes, pdl, fat, sh[0], sh[1], sh[2], us);
                    }
    return 0;
}

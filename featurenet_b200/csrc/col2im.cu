// Data gradient of a convolution with few input channels, as a dense GEMM followed by col2im.
//
// The implicit GEMM for dx has N = C_in columns.  When C_in is small (FeatureNet's early layers: 1, 3,
// 6, 12 channels) every gathered dz element feeds only C_in multiply-adds, so the kernel spends its
// time computing addresses.  The same sum can be regrouped:
//     T[(b,ho,wo), (r,s,c)] = sum_ko dz[(b,ho,wo), ko] * w[r,s,c,ko]          (1)
//     dx[b,h,w,c]           = sum_{r,s} T[(b,ho(h,r),wo(w,s)), (r,s,c)]       (2)
// (1) is a plain [M_out x K] x [K x taps*C] product with no gather at all -- it is exactly the data
// gradient of a dense layer whose weight matrix is w viewed as [taps*C][K], which is w's memory
// layout -- and (2) reads every element of T exactly once.  T lives in the caller's workspace.
// The summation order is fixed (ko inside a tap, taps in (r, s) order), so results are reproducible.
#include "igemm.cuh"

namespace fnet {

__global__ void __launch_bounds__(256) col2im_kernel(ConvGeom g, const float *__restrict__ t, float *__restrict__ dx,
                                                     int accumulate, int act, const float *__restrict__ yact, long long total) {
    pdl_prologue();
    const int taps_c = g.kh * g.kw * g.C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        int c, w, h;
        int64_t b;
        decode4(i, g.H, g.W, g.C, b, h, w, c);
        const float *tb = t + b * g.Ho * g.Wo * taps_c + c;
        float sum = 0.f;
        for (int r = 0; r < g.kh; ++r) {
            int hv = h + g.pt - r * g.dh;
            if (hv < 0) break;                               // later taps only move further up
            int ho = hv / g.sh;
            if (ho * g.sh != hv || ho >= g.Ho) continue;
            for (int s = 0; s < g.kw; ++s) {
                int wv = w + g.pl - s * g.dw;
                if (wv < 0) break;
                int wo = wv / g.sw;
                if (wo * g.sw != wv || wo >= g.Wo) continue;
                sum += __ldg(tb + ((long long)ho * g.Wo + wo) * taps_c + (r * g.kw + s) * g.C);
            }
        }
        store_dx(dx, i, sum, accumulate, act, yact);
    }
}

// bytes of T, or 0 when this formulation is not the better one
size_t col2im_t_bytes(const ConvGeom &g) {
    static const int mode = env_switch("FNET_COL2IM", 1);    // FNET_COL2IM=0 disables (A/B measurements)
    long long taps = (long long)g.kh * g.kw;
    if (!mode || taps < 4 || g.C > 16 || g.K < 32) return 0;   // few output channels: the implicit GEMM is cheaper (K = 12: 54 vs 107 us)
    long long mo = (long long)g.B * g.Ho * g.Wo;
    long long bytes = mo * taps * g.C * 4;
    // worth it once the gather it replaces is long enough; T must stay a reasonable fraction of HBM
    if (mo * taps * g.K * g.C < (1LL << 26) || bytes > (1LL << 31)) return 0;
    return (size_t)((bytes + 255) & ~255LL);
}

int col2im_scatter(const ConvGeom &g, const float *t, float *dx, DxEpilogue ep, cudaStream_t stream) {
    long long total = (long long)g.B * g.H * g.W * g.C;
    if (total == 0) return FNET_OK;
    long long blocks = ceil_div<long long>(total, 256);
    long long cap = (long long)kNumSMs * 16;
    launch_pdl(col2im_kernel, dim3((unsigned)(blocks > cap ? cap : blocks)), dim3(256), 0, stream, g, t, dx, ep.accumulate, ep.act, ep.y, total);
    return check_launch("conv2d_bwd_data[col2im]");
}

}  // namespace fnet

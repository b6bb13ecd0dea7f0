// TMA-fed, warp-specialised, persistent tcgen05 GEMM for the layers that ARE plain GEMMs: Dense and 1x1
// stride-1 convolutions (most of a FeatureNet product: the "Reg_" channel aligners, the x8 / x0.25 1x1
// expansions, Dense on rank-4 tensors).  No im2col gather is involved, so both operands are 2-D boxes that
// the TMA engine copies straight into the swizzled shared-memory layouts tcgen05 reads:
//
//   warp 0 (one lane)  producer: waits for a free stage, arms its mbarrier with the byte count and issues
//                      cp.async.bulk.tensor.2d for A (128 x 32) and B (BN x 32);
//   warp 1 (one lane)  issues four tcgen05.mma.kind::tf32 per stage into one of two TMEM accumulators and
//                      releases the stage (tcgen05.commit -> mbarrier);
//   warps 2..5         epilogue: tcgen05.ld their TMEM lane quadrant, bias + activation, coalesced stores --
//                      while the other accumulator is already being filled for the CTA's next tile.
//
// One CTA per SM walks the output tiles (persistent), so barrier setup, TMEM allocation and descriptor
// prefetch are paid once per SM instead of once per tile, and the epilogue of tile i overlaps the main loop
// of tile i+1.  Forward: A = x [M][K] (K-major), B = w [K][N] (MN-major, SWIZZLE_128B_ATOM_32B boxes of
// 32 x 32).  Data gradient: A = dz [M][N] (K-major), B = w [C][N] (K-major).  Rows / columns / k beyond the
// tensor are zero-filled by the TMA unit, so ragged edges need no code.  The tensor core truncates fp32 to tf32.
#include <cuda.h>
#include "igemm_problems.cuh"

namespace fnet {

constexpr int TG_BM = 128, TG_BK = 32, TG_STAGES = 4, TG_THREADS = 192;
constexpr int TG_A_BYTES = TG_BM * TG_BK * 4;

__device__ __forceinline__ uint32_t tg_smem(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tg_bar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tg_smem(bar)), "r"(count));
}
__device__ __forceinline__ void tg_bar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t addr = tg_smem(bar), done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void tg_bar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tg_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tg_bar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tg_smem(bar)) : "memory");
}
__device__ __forceinline__ void tg_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tg_smem(bar)) : "memory");
}
__device__ __forceinline__ void tg_tma_2d(uint32_t dst, const CUtensorMap *map, uint64_t *bar, int c_inner, int c_outer) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(tg_smem(bar)), "r"(c_inner), "r"(c_outer)
                 : "memory");
}
__device__ __forceinline__ void tg_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}

// K-major SWIZZLE_128B tile (rows of 128 bytes, 8-row atoms of 1024 bytes): what a {32, rows} TMA box produces
__device__ __forceinline__ uint64_t tg_desc_k(uint32_t saddr) {
    uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// MN-major SWIZZLE_128B_BASE32B tile made of {32 (mn), 32 (k)} TMA boxes of 4096 bytes: inside a box the 4-row
// atoms follow each other (stride byte offset 512), the next 32 columns are the next box (leading byte offset 4096)
__device__ __forceinline__ uint64_t tg_desc_mn(uint32_t saddr) {
    uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(4096 >> 4) << 16;
    d |= (uint64_t)(512 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)1 << 61;
    return d;
}

template <class P>
__global__ void __launch_bounds__(TG_THREADS, 1) tma_gemm_kernel(const __grid_constant__ CUtensorMap map_a,
                                                                 const __grid_constant__ CUtensorMap map_b, P p, int M, int N,
                                                                 int Kg, int BN, int n_tiles, int total_tiles, int tmem_cols) {
    constexpr bool B_MN = !P::kBContigK;
    extern __shared__ uint8_t tg_smem_raw[];
    __shared__ uint64_t full_bar[TG_STAGES], empty_bar[TG_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_holder;
    __shared__ float xpose[4][32 * 33];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(tg_smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = tg_smem(smem);
    const int b_bytes = BN * TG_BK * 4, stage_bytes = TG_A_BYTES + b_bytes;
    const int KT = (Kg + TG_BK - 1) / TG_BK;

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < TG_STAGES; ++s) { tg_bar_init(&full_bar[s], 1); tg_bar_init(&empty_bar[s], 1); }
            for (int a = 0; a < 2; ++a) { tg_bar_init(&acc_full[a], 1); tg_bar_init(&acc_empty[a], 4); }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tg_smem(&tmem_holder)),
                     "r"((uint32_t)tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_holder;
    const uint32_t acc_stride = (uint32_t)tmem_cols / 2;      // column offset of the second accumulator

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int m0 = (tile / n_tiles) * TG_BM, n0 = (tile % n_tiles) * BN;
                for (int kt = 0; kt < KT; ++kt) {
                    tg_bar_wait(&empty_bar[stage], phase ^ 1);
                    const uint32_t sa = smem_base + stage * stage_bytes, sb = sa + TG_A_BYTES;
                    tg_bar_expect_tx(&full_bar[stage], (uint32_t)stage_bytes);
                    tg_tma_2d(sa, &map_a, &full_bar[stage], kt * TG_BK, m0);
                    if (B_MN) {
                        for (int j = 0; j < BN / 32; ++j) tg_tma_2d(sb + j * 4096, &map_b, &full_bar[stage], n0 + 32 * j, kt * TG_BK);
                    } else {
                        tg_tma_2d(sb, &map_b, &full_bar[stage], kt * TG_BK, n0);
                    }
                    if (++stage == TG_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor: D fp32, A / B tf32, A K-major, B as the problem says, N = BN, M = 128
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (B_MN ? (1u << 16) : 0u) | ((uint32_t)(BN >> 3) << 17) |
                                   ((uint32_t)(TG_BM >> 4) << 24);
            int stage = 0, it = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
                const int acc = it & 1;
                tg_bar_wait(&acc_empty[acc], (uint32_t)(((it >> 1) & 1) ^ 1));      // the epilogue has drained this accumulator
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t d = tmem + acc * acc_stride;
                for (int kt = 0; kt < KT; ++kt) {
                    tg_bar_wait(&full_bar[stage], phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = smem_base + stage * stage_bytes, sb = sa + TG_A_BYTES;
                    const uint64_t da = tg_desc_k(sa), db = B_MN ? tg_desc_mn(sb) : tg_desc_k(sb);
                    const int badv = B_MN ? 64 : 2;          // 16-byte units per 8 k: two 4-row atoms / 32 bytes along the row
#pragma unroll
                    for (int j = 0; j < TG_BK / 8; ++j)
                        tg_mma(d, da + (uint64_t)(2 * j), db + (uint64_t)(badv * j), idesc, (kt > 0 || j > 0) ? 1u : 0u);
                    tg_commit(&empty_bar[stage]);
                    if (++stage == TG_STAGES) { stage = 0; phase ^= 1; }
                }
                tg_commit(&acc_full[acc]);
            }
        }
    } else {
        const int quad = warp & 3;                            // the TMEM lane quadrant this warp may read
        float *xp = xpose[warp - 2];
        int it = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
            const int acc = it & 1;
            const int m0 = (tile / n_tiles) * TG_BM, n0 = (tile % n_tiles) * BN;
            tg_bar_wait(&acc_full[acc], (uint32_t)((it >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int m_first = m0 + quad * 32, count = min(32, M - m_first);
            for (int c0 = 0; c0 < BN; c0 += 32) {
                uint32_t v[32];
                uint32_t taddr = tmem + ((uint32_t)(quad * 32) << 16) + acc * acc_stride + (uint32_t)c0;
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; ++j) xp[lane * 33 + j] = __uint_as_float(v[j]);      // thread = row
                __syncwarp();
                const int n = n0 + c0 + lane;                                                 // thread = column
                if (n < N && count > 0) p.store_col(m_first, count, n, xp + lane, 33);
                __syncwarp();
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) tg_bar_arrive(&acc_empty[acc]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)tmem_cols) : "memory");
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult status;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &status) == cudaSuccess &&
            status == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    return fn;
}

// a row-major fp32 matrix [rows][cols] seen through boxes of box_rows x box_cols
static bool make_map(CUtensorMap *map, const float *base, long long rows, long long cols, int box_rows, int box_cols,
                     CUtensorMapSwizzle swizzle) {
    EncodeTiledFn fn = encode_tiled();
    if (!fn) return false;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * sizeof(float)};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t elem[2] = {1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, elem,
              CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static bool is_plain_gemm(const ConvGeom &g) {
    return g.kh == 1 && g.kw == 1 && g.sh == 1 && g.sw == 1 && g.pt == 0 && g.pl == 0 && g.Ho == g.H && g.Wo == g.W;
}

static int tg_bn(long long n) {
    long long bn = ((n + 31) / 32) * 32;
    return (int)(bn > 256 ? 256 : bn);
}

// product: 0 forward (N = g.K, reduction g.C), 1 data gradient (N = g.C, reduction g.K)
bool tma_gemm_supported(const ConvGeom &g, int math, int product) {
    static int mode = -1;                                    // FNET_TMA=0 disables (A/B measurements)
    if (mode < 0) {
        const char *e = getenv("FNET_TMA");
        mode = (e && e[0] == '0') ? 0 : 1;
    }
    if (!mode || math != FNET_MATH_TF32 || product > 1 || !is_plain_gemm(g) || !encode_tiled()) return false;
    long long M = (long long)g.B * g.H * g.W;
    long long N = product == 0 ? g.K : g.C, Kg = product == 0 ? g.C : g.K;
    if ((g.C & 3) || (g.K & 3) || N < 16 || Kg < 8 || M > 0x7fffffffLL) return false;      // 16-byte global strides
    long long tiles = ceil_div<long long>(M, TG_BM) * ceil_div<long long>(N, tg_bn(N));
    return tiles >= kNumSMs;                                 // no split reduction here: the cp.async kernels do that
}

template <class P>
static int tma_gemm_run(const P &p, const float *a, long long M, long long Kg, const float *b, long long N, bool b_mn,
                        cudaStream_t stream, const char *what) {
    if ((reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15)) {
        set_error("%s: operands must be 16-byte aligned for the TMA path", what);
        return FNET_ERR_ARG;
    }
    const int BN = tg_bn(N);
    CUtensorMap map_a, map_b;
    bool ok = make_map(&map_a, a, M, Kg, TG_BM, TG_BK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (b_mn) ok = ok && make_map(&map_b, b, Kg, N, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);    // w [K][N]
    else ok = ok && make_map(&map_b, b, N, Kg, BN, TG_BK, CU_TENSOR_MAP_SWIZZLE_128B);               // w [C][Kg]
    if (!ok) { set_error("%s: cuTensorMapEncodeTiled failed", what); return FNET_ERR_CUDA; }
    const int n_tiles = (int)ceil_div<long long>(N, BN);
    const long long total = ceil_div<long long>(M, TG_BM) * n_tiles;
    int tmem_cols = 64;
    while (tmem_cols < 2 * BN) tmem_cols *= 2;
    const size_t smem = (size_t)TG_STAGES * (TG_A_BYTES + BN * TG_BK * 4) + 1024;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(tma_gemm_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("%s: cudaFuncSetAttribute: %s", what, cudaGetErrorString(e)); return FNET_ERR_CUDA; }
        configured = true;
    }
    const unsigned grid = (unsigned)(total < kNumSMs ? total : kNumSMs);
    tma_gemm_kernel<P><<<grid, TG_THREADS, smem, stream>>>(map_a, map_b, p, (int)M, (int)N, (int)Kg, BN, n_tiles, (int)total,
                                                          tmem_cols);
    return check_launch(what);
}

int tma_gemm_fwd(const ConvGeom &g, int act, const float *x, const float *w, const float *bias, float *y, cudaStream_t stream) {
    FwdProblem p{g, x, w, bias, y, act == FNET_ACT_SOFTMAX ? FNET_ACT_NONE : act};
    return tma_gemm_run(p, x, (long long)g.B * g.H * g.W, g.C, w, g.K, true, stream, "conv2d_fwd[tf32,tma]");
}

int tma_gemm_dgrad(const ConvGeom &g, const float *dz, const float *w, float *dx, int accumulate, cudaStream_t stream) {
    DgradProblem p{g, dz, w, dx, accumulate};
    return tma_gemm_run(p, dz, (long long)g.B * g.H * g.W, g.K, w, g.C, false, stream, "conv2d_bwd_data[tf32,tma]");
}

}  // namespace fnet
